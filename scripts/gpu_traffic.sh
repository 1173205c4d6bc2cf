#!/bin/bash
# DRAM bytes of ONE gemm_topk_pair launch of the bench workload (single-pass ncu metrics) -> profiles/gemm_topk_traffic.json
mkdir -p gpurun_out
timeout -s KILL 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:gemm_topk_pair -s 3 -c 1 --csv --log-file gpurun_out/traffic.csv \
    python bench.py --steps 1 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/traffic.log 2>&1
echo "ncu traffic exit $?"
grep -E "dram__bytes|gpu__time" gpurun_out/traffic.csv
