#!/bin/bash
# usage (gpurun --gpus N): bash scripts/gpu_ag_probe.sh N
N=${1:-8}
mkdir -p gpurun_out
run() { tag=$1; shift; env PROBE_TAG=$tag "$@" timeout -s KILL 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
      --master-addr 127.0.0.1 --master-port 29517 scripts/ag_probe.py 2>&1 | grep -v "^W\|warn\|Warn" ; }
{
run default
run ring NCCL_ALGO=Ring
run nvls_off NCCL_NVLS_ENABLE=0
run ring_simple NCCL_ALGO=Ring NCCL_PROTO=Simple
} > gpurun_out/ag_probe_n$N.log 2>&1
tail -60 gpurun_out/ag_probe_n$N.log
