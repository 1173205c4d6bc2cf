"""Timing of the k = ntotal path (l2sq_all + stable radix sort) at BASELINE.json configs[0] and at the reference's
real trimmed size, batched and one query per call (the reference's loop)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vqwild_b200 as vq

dev = torch.device("cuda", 0)
for nq, n in [(4096, 20480), (6338, 7670)]:
    g = torch.randn((n, 512), device=dev)
    g /= g.norm(dim=1, keepdim=True)
    q = torch.randn((nq, 512), device=dev)
    q /= q.norm(dim=1, keepdim=True)
    ix = vq.IndexFlatL2(512)
    ix.add(g)
    D = torch.empty((nq, n), dtype=torch.float32, device=dev)
    I = torch.empty((nq, n), dtype=torch.int64, device=dev)
    for _ in range(2):
        ix.search(q, n, out=(D, I))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ix.search(q, n, out=(D, I))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("full sort batched  nq=%d n=%d: %.2f ms  (%.1f G elements/s sorted, launches %d)" %
          (nq, n, ms, nq * n / ms / 1e6, ix.last_stats()["launches"]))
    qh = q[:200].cpu().numpy()
    t0 = time.perf_counter()
    for i in range(200):
        ix.search(qh[i].reshape(1, 512), n)
    dt = (time.perf_counter() - t0) / 200
    print("  one query per call (host in / host out, the reference's loop): %.3f ms per query" % (dt * 1e3))
