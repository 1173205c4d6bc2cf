#!/bin/bash
mkdir -p gpurun_out
PCMD="python scripts/prof_kernels.py"
$PCMD > gpurun_out/plain_kernels2.log 2>&1 &&
timeout -s KILL 1500 ncu --set full --clock-control none --import-source on \
    -k regex:"moment_nms|ap_rows_kernel|segmax_|scan_i32|topk_merge_heads|l2sq_all|nms_masks" -c 24 \
    -f -o gpurun_out/prof_kernels2 $PCMD > gpurun_out/ncu_kernels2.log 2>&1
echo "ncu kernels2 exit $?"; tail -4 gpurun_out/plain_kernels2.log
