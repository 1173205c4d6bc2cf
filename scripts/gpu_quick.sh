#!/bin/bash
# Quick GPU pass: parity tests, then (only if they pass) the headline bench without the side measurements.
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests -q -m gpu -x --timeout=120 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
rc=$?
echo "pytest exit $rc"
tail -15 gpurun_out/pytest_gpu.log
[ $rc -ne 0 ] && exit 1
timeout -s KILL 240 python bench.py --no-aux --no-cpu-baseline "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
echo "bench exit $?"
cat gpurun_out/bench_quick.json; tail -5 gpurun_out/bench_quick.err
