import sys; sys.path.insert(0, "/root/repo")
import torch, vqwild_b200 as vq
from vqwild_b200 import _lib as L
for d in (512, 1024):
    g = torch.randn(20480, d, device="cuda"); g /= g.norm(dim=1, keepdim=True)
    q = g[:4096].contiguous()
    ix = vq.IndexFlatL2(d); ix.add(g)
    for w in (0, 2, 3, 4, 6, 9):
        L.debug_set("WALKERS", w)
        for _ in range(3): ix.search(q, 100)
        torch.cuda.synchronize()
        ts = []
        for _ in range(5):
            ix.search(q, 100); torch.cuda.synchronize(); ts.append(ix.last_timing()["gemm_topk_ms"])
        print(d, "walkers", w, "gemm %.3f ms" % min(ts), ix.last_stats()["segments"], flush=True)
    L.debug_set("WALKERS", 0)
