#!/bin/bash
# A/B of debug hooks on ONE box, interleaved: scripts/ab_env.sh <n> <reps> VAR1 VAR2 ...  (each VAR is tried set to 1 against the baseline)
n=$1; reps=$2; shift 2
for r in $(seq $reps); do
  for v in BASE "$@"; do
    if [ $v = BASE ]; then env python bench.py --steps 3 --warmup 3 --no-aux --no-cpu-baseline --n $n > gpurun_out/ab_tmp.json 2>gpurun_out/ab_tmp.err
    else env $v=1 python bench.py --steps 3 --warmup 3 --no-aux --no-cpu-baseline --n $n > gpurun_out/ab_tmp.json 2>gpurun_out/ab_tmp.err; fi
    python -c "
import json,sys;d=json.loads(open('gpurun_out/ab_tmp.json').read().strip().splitlines()[-1]);print('n=%s rep=%s %-28s step %.2f ms gemm %.2f ms select %.2f  clk %s  checksum %s'%(sys.argv[1],sys.argv[2],sys.argv[3],d['ms_per_step'],d['phases_ms']['gemm_topk_ms'],d['phases_ms']['select_ms'],d['clocks']['sm_mhz'],d['result_checksum']))" $n $r $v
  done
done
