"""Moment-retrieval evaluation (a8) at BASELINE.json configs[2] scale: V videos x S sliding moments, 512-d f32.
Usage: python scripts/prof_moment.py [nq] [V] [S]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import vqwild_b200 as vq

nq = int(sys.argv[1]) if len(sys.argv) > 1 else 409
V = int(sys.argv[2]) if len(sys.argv) > 2 else 20480
S = int(sys.argv[3]) if len(sys.argv) > 3 else 64
d, dev = 512, "cuda"
n = V * S
gen = torch.Generator(device=dev).manual_seed(0)
g = torch.randn(n, d, device=dev, generator=gen)
g = g / g.norm(dim=1, keepdim=True) * (0.6 + 0.4 * torch.rand(n, 1, device=dev, generator=gen))
q = torch.randn(nq, d, device=dev, generator=gen)
q = q / q.norm(dim=1, keepdim=True)
rng = np.random.default_rng(0)
# sliding moments: window lengths 1..~6 clips of 5 s at stride 5 s (dataloader_baseline.py:1116-1137)
s_in = np.arange(n) % S
length = 1 + (s_in % 6)
start = (5 * (s_in // 6)).astype(np.float32)
end = start + 5 * length.astype(np.float32)
video = (np.arange(n) // S).astype(np.int32)
hit = rng.integers(-1, 180, n).astype(np.int32)
iou = np.where(hit >= 0, rng.uniform(0, 1, n), 0).astype(np.float32)
ix = vq.IndexFlatL2(d)
ix.add(g)
ix.set_meta(hit, video)
ix.set_segments(np.arange(0, n + 1, S, dtype=np.int64))
ix.set_moments(start, end, iou, hit)
ql = torch.from_numpy(rng.integers(0, 180, nq).astype(np.int32)).to(dev)
ign = torch.from_numpy(rng.integers(0, V, (nq, 5)).astype(np.int32)).to(dev)
for it in range(6):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ap, rec, nk, _ = ix.eval_moment(q, ql, ign)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("eval_moment nq=%d V=%d S=%d: %.1f ms  (%.3f ms/query, %.3g pairs/s)  mAP=%.4f kept/query=%.0f"
          % (nq, V, S, dt * 1e3, dt * 1e3 / nq, nq * n / dt, float(ap.mean()), float(nk.double().mean())), flush=True)
