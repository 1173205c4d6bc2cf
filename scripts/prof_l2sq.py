"""l2sq_all at the configs[0] shape, three calls (for ncu -k regex:l2sq_all)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vqwild_b200 import ops
nq, n, d = 4096, 20480, 512
g = torch.randn((n, d), device="cuda"); g /= g.norm(dim=1, keepdim=True)
q = g[torch.randperm(n, device="cuda")[:nq]].contiguous()
for _ in range(3):
    out = ops.l2sq(q, g)
torch.cuda.synchronize()
print(float(out[0, :4].sum()))
