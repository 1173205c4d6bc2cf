#!/bin/bash
# ncu evidence round: (1) tests, (2) --set full captures of the secondary kernels, (3) --set full of gemm_topk_pair AT THE
# BENCH LAUNCH, (4) launch list of the bench command.
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests -q -m gpu -x --timeout=600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
PCMD="python scripts/prof_kernels.py"
$PCMD > gpurun_out/plain_kernels.log 2>&1 &&
timeout -s KILL 1200 ncu --set full --clock-control none --import-source on \
    -k regex:"select_kernel|l2sq_all|ap_rows|moment_nms|topk_merge_keys|rowsort" -c 16 \
    -f -o gpurun_out/prof_kernels $PCMD > gpurun_out/ncu_kernels.log 2>&1
echo "ncu kernels exit $?"; tail -3 gpurun_out/plain_kernels.log
BCMD="python bench.py --steps 1 --warmup 3 --no-aux --no-cpu-baseline --no-parity"
$BCMD > gpurun_out/plain_bench.log 2>&1 &&
timeout -s KILL 1200 ncu --set full --clock-control none --import-source on -k regex:gemm_topk -s 3 -c 1 \
    -f -o gpurun_out/prof_gemm_topk_bench $BCMD > gpurun_out/ncu_gemm_bench.log 2>&1
echo "ncu gemm bench exit $?"
BCMD2="python bench.py --steps 2 --warmup 3 --no-aux --no-cpu-baseline --no-parity"
timeout -s KILL 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv $BCMD2 > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
