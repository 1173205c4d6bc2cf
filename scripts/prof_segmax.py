"""configs[2] segment-max search with a bf16 gallery, three calls (for ncu -k regex:gemm_topk_pair -s 2 -c 1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, vqwild_b200 as vq
nq, V, S, d = 4096, 20480, 64, 512
n = V * S
dev = torch.device("cuda", 0)
xq = bench.gen_rows(torch, "q", 0, nq, d, dev)
g = bench.gen_rows(torch, "g", 0, n, d, dev)
ix = vq.IndexFlatL2(d); ix.add(g); ix.set_segments(np.arange(0, n + 1, S, dtype=np.int64))
for _ in range(3):
    Vv, R, D = ix.search_segmax(xq, 100)
torch.cuda.synchronize()
print(ix.last_timing(), ix.last_stats())
