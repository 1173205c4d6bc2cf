#!/bin/bash
# A/B of one debug variable's VALUES on ONE box, interleaved: scripts/ab_val.sh <n> <reps> VAR v1 v2 ...  ("-" = unset)
n=$1; reps=$2; var=$3; shift 3
for r in $(seq $reps); do
  for v in "$@"; do
    if [ "$v" = "-" ]; then python bench.py --steps 3 --warmup 3 --no-aux --no-cpu-baseline --n $n > gpurun_out/ab_tmp.json 2>gpurun_out/ab_tmp.err
    else env $var=$v python bench.py --steps 3 --warmup 3 --no-aux --no-cpu-baseline --n $n > gpurun_out/ab_tmp.json 2>gpurun_out/ab_tmp.err; fi
    python -c "
import json,sys;d=json.loads(open('gpurun_out/ab_tmp.json').read().strip().splitlines()[-1]);print('n=%s rep=%s %s=%-6s step %.2f ms gemm %.2f ms select %.2f  clk %s  checksum %s fallback %s'%(sys.argv[1],sys.argv[2],sys.argv[3],sys.argv[4],d['ms_per_step'],d['phases_ms']['gemm_topk_ms'],d['phases_ms']['select_ms'],d['clocks']['sm_mhz'],d['result_checksum'],d['search_stats']['fallback_queries']))" $n $r $var "$v"
  done
done
