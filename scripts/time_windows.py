"""Device timing of vq_window_means (f1) at untrimmed-gallery scale: V videos x T frames x 512 channels, the sliding
moment grid of garner_feat.  Usage: python scripts/time_windows.py [V] [T]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from vqwild_b200 import ops
from vqwild_b200.retrieval import moment_grid

V = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
T = int(sys.argv[2]) if len(sys.argv) > 2 else 600
C = 512
s, e, lo, hi = moment_grid(T / 3.0 + 0.4, T)
nw = len(s)
t0 = np.arange(V + 1, dtype=np.int64) * T
ws = np.arange(V + 1, dtype=np.int64) * nw
feat = torch.rand(C, V * T, device="cuda")
lo_d = torch.from_numpy(np.tile(lo, V).astype(np.int32)).cuda()
hi_d = torch.from_numpy(np.tile(hi, V).astype(np.int32)).cuda()
for _ in range(2):
    out = ops.window_means(feat, t0, ws, lo_d, hi_d)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    out = ops.window_means(feat, t0, ws, lo_d, hi_d)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
nbytes = 4.0 * C * V * T + 4.0 * C * V * nw
print("window_means V=%d T=%d windows/video=%d: %.3f ms  %.1f GB/s algorithmic (%.2f GB)" % (V, T, nw, ms, nbytes / ms / 1e6, nbytes / 1e9))
ref = feat[:, 5 * T + int(lo[7]):5 * T + int(hi[7])].double().mean(1).float()
print("check", float((out[5 * nw + 7] - ref).abs().max()))
