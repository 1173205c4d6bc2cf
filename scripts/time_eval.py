"""Evaluator timings: canonical all-pairs scores (l2sq_all, packed FFMA2/FADD2), the configs[0] trimmed evaluation and
a 1024-query slice of the configs[2] moment evaluation, with the fp32-pipe issue bound beside each."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import vqwild_b200 as vq
from vqwild_b200 import ops

dev = torch.device("cuda", 0)


def ev_time(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


nq, n, d = 4096, 20480, 512
g = torch.randn((n, d), device=dev)
g /= g.norm(dim=1, keepdim=True)
q = g[torch.randperm(n, device=dev)[:nq]].contiguous()
ms = ev_time(lambda: ops.l2sq(q, g))
steps = nq * n * d
# bound: one packed step = FADD2 + FFMA2 for two pairs; 148 SMs x 4 sub-partitions x 1 issue / clk
print("l2sq_all %d x %d x %d: %.3f ms = %.2f T canonical steps/s (%.1f TFLOP/s as 2*Q*N*D); per-SM-subpartition issue "
      "slots used by the packed math at 1.9 GHz: %.0f %%" % (nq, n, d, ms, steps / ms / 1e9, 2 * steps / ms / 1e9,
                                                        100.0 * (steps / 32.0) / (ms * 1e-3 * 148 * 4 * 1.9e9)))
label = torch.randint(0, 180, (n,), device=dev, dtype=torch.int32)
video = torch.arange(n, device=dev, dtype=torch.int32)
ix = vq.IndexFlatL2(d)
ix.add(g)
ix.set_meta(label, video)
ql = torch.randint(0, 180, (nq,), device=dev, dtype=torch.int32)
ign = torch.randint(0, n, (nq, 5), device=dev, dtype=torch.int32)
from vqwild_b200 import _lib as L
import bench
for nm, gg, lab in (("random labels, unit rows", g, label),):
    for hook in (0, 1):
        L.debug_set("NO_TC_EVAL", hook)
        ms = ev_time(lambda: ix.eval_trimmed(q, ql, ign))
        print("eval_trimmed configs[0] shape, %s, %s: %.3f ms (handed back: %d)" % (nm, "canonical all-pairs" if hook else "tensor-core form", ms, ix.last_stats()["segments"]))
L.debug_set("NO_TC_EVAL", 0)
g0, label0, video0, qrow0 = bench.cfg0_problem(np)
gd = torch.from_numpy(g0).to(dev)
ix0 = vq.IndexFlatL2(d)
ix0.add(gd)
ix0.set_meta(torch.from_numpy(label0).to(dev), torch.from_numpy(video0).to(dev))
qd = gd[torch.from_numpy(qrow0).to(dev)].contiguous()
ql0 = torch.from_numpy(label0[qrow0]).to(dev)
ig0 = torch.from_numpy(video0[qrow0][:, None].astype(np.int32)).to(dev)
for hook in (0, 1):
    L.debug_set("NO_TC_EVAL", hook)
    ms = ev_time(lambda: ix0.eval_trimmed(qd, ql0, ig0))
    ap = ix0.eval_trimmed(qd, ql0, ig0)[0]
    print("eval_trimmed configs[0] (bench data, class-structured), %s: %.3f ms (handed back: %d) mAP %.6f" % ("canonical all-pairs" if hook else "tensor-core form", ms, ix0.last_stats()["segments"], float(ap.double().mean())))
L.debug_set("NO_TC_EVAL", 0)
del ix0, gd
qb, gb = q.bfloat16(), g.bfloat16()
ms = ev_time(lambda: ops.l2sq(qb.float(), gb.float()))
# moment slice
V, S, nq2 = 20480, 64, 1024
n2 = V * S
g2 = torch.randn((n2, d), device=dev)
g2 /= g2.norm(dim=1, keepdim=True)
q2 = torch.randn((nq2, d), device=dev)
q2 /= q2.norm(dim=1, keepdim=True)
rng = np.random.default_rng(0)
s_in = np.arange(n2) % S
start = (5 * (s_in // 6)).astype(np.float32)
end = start + 5 * (1 + s_in % 6).astype(np.float32)
vid2 = (np.arange(n2) // S).astype(np.int32)
hit = rng.integers(-1, 180, n2).astype(np.int32)
iou = np.where(hit >= 0, rng.uniform(0, 1, n2), 0).astype(np.float32)
im = vq.IndexFlatL2(d)
im.add(g2)
im.set_meta(hit, vid2)
im.set_segments(np.arange(0, n2 + 1, S, dtype=np.int64))
im.set_moments(start, end, iou, hit)
ql2 = torch.from_numpy(rng.integers(0, 180, nq2).astype(np.int32)).to(dev)
ign2 = torch.from_numpy(rng.integers(0, V, (nq2, 5)).astype(np.int32)).to(dev)
out = torch.empty((256, n2), dtype=torch.float32, device=dev)
ms = ev_time(lambda: ops.l2sq(q2[:256].contiguous(), g2), reps=3, warm=1)
print("l2sq_all 256 x %d x %d: %.2f ms (%.2f T steps/s)" % (n2, d, ms, 256.0 * n2 * d / ms / 1e9))
ms = ev_time(lambda: im.eval_moment(q2, ql2, ign2), reps=2, warm=1)
print("eval_moment %d queries x %d moments: %.1f ms = %.4f ms per query" % (nq2, n2, ms, ms / nq2))
