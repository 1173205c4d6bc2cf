#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests -q -m gpu -x --timeout=600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
for i in 1 2; do timeout 300 python scripts/prof_gemm.py 100000 1250000 4 2>&1 | tail -1; done | tee gpurun_out/prof_n8shape.log
timeout 300 python scripts/prof_gemm.py 100000 2500000 3 2>&1 | tail -1 | tee -a gpurun_out/prof_n8shape.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/bench_noaux.json 2> gpurun_out/bench_noaux.err
echo "bench exit $?"; cat gpurun_out/bench_noaux.json
