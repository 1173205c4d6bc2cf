"""configs[2] shape: plain top-100 search vs per-video segment-max search, bf16, 4096 queries x 20480 videos x 64 moments."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, vqwild_b200 as vq
nq, V, S, d = int(sys.argv[1]) if len(sys.argv) > 1 else 4096, 20480, 64, 512
n = V * S
dev = torch.device("cuda", 0)
xq = bench.gen_rows(torch, "q", 0, nq, d, dev)
g = bench.gen_rows(torch, "g", 0, n, d, dev)
ix = vq.IndexFlatL2(d); ix.add(g); ix.set_segments(np.arange(0, n + 1, S, dtype=np.int64))
for name, fn in (("search top-100", lambda: ix.search(xq, 100)), ("search_segmax top-100", lambda: ix.search_segmax(xq, 100))):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t = ix.last_timing(); st = ix.last_stats()
    print("%-24s nq=%d: %s  TFLOP/s(gemm)=%.0f  %s" % (name, nq, {k: round(v, 2) for k, v in t.items()}, 2.0 * nq * n * d / t["gemm_topk_ms"] / 1e9, st), flush=True)
