"""configs[2] segment-max search, bf16 and f32 gallery: phases with the video-major re-scoring and (hook) the per-query one."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, vqwild_b200 as vq
from vqwild_b200 import _lib as L
nq, V, S, d = 4096, 20480, 64, 512
n = V * S
dev = torch.device("cuda", 0)
xq = bench.gen_rows(torch, "q", 0, nq, d, dev)
g = bench.gen_rows(torch, "g", 0, n, d, dev)
for name, gg, qq in (("bf16", g, xq), ("f32", g.float(), xq.float())):
    ix = vq.IndexFlatL2(d); ix.add(gg); ix.set_segments(np.arange(0, n + 1, S, dtype=np.int64))
    ref = None
    for label, hooks in (("unique arg-max: one score", {}), ("video-major re-scoring", {"NO_SEG_UNIQUE": 1}),
                         ("per-query re-scoring", {"NO_VMAJOR": 1})):
        for h in ("NO_SEG_UNIQUE", "NO_VMAJOR"):
            L.debug_set(h, hooks.get(h, 0))
        for _ in range(3):
            Vv, R, D = ix.search_segmax(qq, 100)
        torch.cuda.synchronize()
        t = ix.last_timing()
        cs = (int(R.sum().item()), int(Vv.sum().item()), int(D.view(torch.int32).long().sum().item()))
        ref = cs if ref is None else ref
        print("%-5s %-26s %s checksums %s fallback %d" % (name, label, {k: round(v, 2) for k, v in t.items()},
                                                        "same" if cs == ref else "DIFFERENT", ix.last_stats()["fallback_queries"]), flush=True)
    for h in ("NO_SEG_UNIQUE", "NO_VMAJOR"):
        L.debug_set(h, 0)
    del ix
