#!/bin/bash
# Round-end style pass on one GPU: build check, smoke, GPU tests, default bench, reference arm, side timings.
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout -s KILL 900 python -m pytest tests -q -m gpu -x --timeout=600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -2 gpurun_out/pytest_gpu.log
( time timeout -s KILL 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err ) 2> gpurun_out/bench_time.log
echo "bench exit $?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench_time.log
timeout -s KILL 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "bench ref exit $?"; cat gpurun_out/bench_ref.json
timeout 300 python scripts/time_eval.py > gpurun_out/time_eval.log 2>&1; cat gpurun_out/time_eval.log
timeout 300 python scripts/time_fullsort.py > gpurun_out/time_fullsort.log 2>&1; cat gpurun_out/time_fullsort.log
