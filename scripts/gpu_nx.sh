#!/bin/bash
# N-GPU bench (no CPU legs: box time at N GPUs is charged N-fold)
mkdir -p gpurun_out
N=${1:-8}
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench N=$N exit $?"; cat gpurun_out/bench_n$N.json; tail -2 gpurun_out/bench_n$N.err
