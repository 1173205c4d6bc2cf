#!/bin/bash
# Round GPU pass: parity tests, headline bench, ncu launch list of the bench command, ncu --set full of the top kernel.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout -s KILL 1500 python -m pytest tests -q -m gpu -x --timeout=600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee gpurun_out/status.log
tail -3 gpurun_out/pytest_gpu.log
timeout -s KILL 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?" | tee -a gpurun_out/status.log
cat gpurun_out/bench.json
timeout -s KILL 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "bench ref exit $?" | tee -a gpurun_out/status.log
cat gpurun_out/bench_ref.json
BCMD="python bench.py --steps 2 --warmup 3 --no-aux --no-cpu-baseline"
$BCMD > gpurun_out/plain_bench.log 2>&1 &&
timeout -s KILL 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv $BCMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?" | tee -a gpurun_out/status.log
PCMD="python scripts/prof_gemm.py 18944 1048576 2"
$PCMD > gpurun_out/plain_prof.log 2>&1 &&
timeout -s KILL 900 ncu --set full --clock-control none --import-source on -k regex:gemm_topk -s 1 -c 1 \
    -f -o gpurun_out/prof_gemm_topk $PCMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?" | tee -a gpurun_out/status.log
cat gpurun_out/plain_prof.log | tail -2
# kernel (3): batched batch-hard triplet timing + phase clocks + one ncu --set full capture
timeout -s KILL 300 python scripts/time_triplet.py > gpurun_out/time_triplet.log 2>&1
echo "time_triplet exit $?" | tee -a gpurun_out/status.log
cat gpurun_out/time_triplet.log
timeout -s KILL 300 python scripts/clk_triplet.py > gpurun_out/clk_triplet.log 2>&1
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:triplet_tc -s 48 -c 1 \
    -f -o gpurun_out/prof_triplet_tc python scripts/time_triplet.py > gpurun_out/ncu_full_triplet.log 2>&1
echo "ncu triplet exit $?" | tee -a gpurun_out/status.log
