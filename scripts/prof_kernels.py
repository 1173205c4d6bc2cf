"""One launch of every kernel VERDICT r1 asked ncu evidence for, at representative shapes (for `ncu --set full -k regex:...`):
select_kernel (100 000 x 1 M top-100), l2sq_all + ap_rows (configs[0]), l2sq_all + moment_nms + ap_rows<MOMENT>
(256 queries of configs[2]), topk_merge_keys (G = 8, 100 000 x 100), the full-sort kernels (1024 x 20 480)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
import vqwild_b200 as vq
from vqwild_b200 import ops

dev = torch.device("cuda", 0)
rng = np.random.default_rng(0)
# search: select_kernel
nq, n = 100000, 1 << 20
xq = bench.gen_rows(torch, "q", 0, nq, 512, dev)
ix = vq.IndexFlatL2(512)
ix.add(bench.gen_rows(torch, "g", 0, n, 512, dev))
D, I = ix.search(xq, 100)
torch.cuda.synchronize()
print("search", ix.last_timing())
# merge of 8 shard lists
keys = torch.stack([ops.topk_pack(D + 0.001 * g, I, g * n) for g in range(8)])
od, oi = ops.topk_merge_keys(keys)
hd, hi_, inc = ops.topk_merge_heads(keys[:, :, :40].contiguous(), 100)   # round 1 of the exchange at G = 8
torch.cuda.synchronize()
print("heads incomplete:", int(inc.item()))
del ix, keys, D, I
# configs[0] evaluation
g0, label, video, qrow = bench.cfg0_problem(np)
gd = torch.from_numpy(g0).to(dev)
ix = vq.IndexFlatL2(512)
ix.add(gd)
ix.set_meta(torch.from_numpy(label).to(dev), torch.from_numpy(video).to(dev))
qd = gd[torch.from_numpy(qrow).to(dev)].contiguous()
ap, rec, _ = ix.eval_trimmed(qd, torch.from_numpy(label[qrow]).to(dev), torch.from_numpy(video[qrow][:, None].astype(np.int32)).to(dev))
print("cfg0 mAP", float(ap.mean()))
Df, If = ix.search(qd[:1024].contiguous(), 20480)   # full sort
torch.cuda.synchronize()
del ix, gd, Df, If
# configs[2] moment evaluation, 256 queries
V, S, d, nq2 = 20480, 64, 512, 256
n2 = V * S
g2 = torch.randn((n2, d), device=dev)
g2 /= g2.norm(dim=1, keepdim=True)
q2 = torch.randn((nq2, d), device=dev)
q2 /= q2.norm(dim=1, keepdim=True)
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
Vs, Rs, Ds = im.search_segmax(q2, 100)          # SEGMAX GEMM + video-major re-scoring (f32 gallery)
torch.cuda.synchronize()
apm, recm, _, _ = im.eval_moment(q2, torch.from_numpy(rng.integers(0, 180, nq2).astype(np.int32)).to(dev),
                                 torch.from_numpy(rng.integers(0, V, (nq2, 5)).astype(np.int32)).to(dev))
torch.cuda.synchronize()
print("moment mAP", float(apm.mean()))
