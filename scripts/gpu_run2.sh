#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests -q -m gpu -x --timeout=600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python scripts/time_eval.py 2>&1 | tee gpurun_out/time_eval.log
timeout 300 python scripts/time_fullsort.py 2>&1 | tee gpurun_out/time_fullsort.log
PCMD="python scripts/prof_gemm.py 100000 1250000 2"
$PCMD > gpurun_out/plain_prof.log 2>&1 &&
timeout -s KILL 900 ncu --set full --clock-control none --import-source on -k regex:gemm_topk -s 1 -c 1 \
    -f -o gpurun_out/prof_gemm_topk_n8shape $PCMD > gpurun_out/ncu_full.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/plain_prof.log
