"""What a rank's local search costs when it asks for only the head of its list (k_local = head_len(k, G) instead of k):
the shard shapes of configs[3] at G = 8 / 4 on one GPU."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import vqwild_b200 as vq
from vqwild_b200 import ops

dev = torch.device("cuda", 0)
nq, d = 100000, 512
xq = bench.gen_rows(torch, "q", 0, nq, d, dev)
for G in (8, 4):
    n = 10_000_000 // G
    ix = vq.IndexFlatL2(d)
    ix.add(bench.gen_rows(torch, "g", 0, n, d, dev))
    for k in (100, ops.head_len(100, G)):
        D = torch.empty((nq, k), dtype=torch.float32, device=dev)
        I = torch.empty((nq, k), dtype=torch.int64, device=dev)
        for _ in range(2):
            ix.search(xq, k, out=(D, I))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            ix.search(xq, k, out=(D, I))
        e1.record()
        torch.cuda.synchronize()
        print("G=%d shard rows %d  k_local=%3d: %.2f ms/search  %s  stats %s" % (
            G, n, k, e0.elapsed_time(e1) / 3, {a: round(b, 2) for a, b in ix.last_timing().items()}, ix.last_stats()), flush=True)
    del ix
