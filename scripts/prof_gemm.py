"""One tensor-core search on a mid-size bf16 problem (148 query blocks x ~1M rows) for ncu captures.
Usage: python scripts/prof_gemm.py [nq] [n] [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import vqwild_b200 as vq

nq = int(sys.argv[1]) if len(sys.argv) > 1 else 148 * 128
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
dev = torch.device("cuda", 0)
xq = bench.gen_rows(torch, "q", 0, nq, 512, dev)
g = bench.gen_rows(torch, "g", 0, n, 512, dev)
ix = vq.IndexFlatL2(512)
ix.add(g)
for _ in range(reps):
    D, I = ix.search(xq, int(os.environ.get("PROF_K", "100")))
torch.cuda.synchronize()
t = ix.last_timing()
print("nq=%d n=%d" % (nq, n), ix.last_stats(), t, "TFLOP/s=%.1f" % (2.0 * nq * n * 512 / t["gemm_topk_ms"] / 1e9))
