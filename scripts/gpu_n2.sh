#!/bin/bash
# 2-GPU validation: sharded parity script + a short bench at N=2 (phases, parity_subsample)
mkdir -p gpurun_out
N=${1:-2}
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench N=$N exit $?"; cat gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
