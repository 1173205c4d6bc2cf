"""configs[2] segment-max search: parts per query group ("walkers") of the SEGMAX plan, A/B through the WALKERS hook."""
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
    for w in (0, 4, 5, 9, 14, 18):
        L.debug_set("WALKERS", w)
        for _ in range(3):
            Vv, R, D = ix.search_segmax(qq, 100)
        torch.cuda.synchronize()
        t = ix.last_timing()
        cs = (int(R.sum().item()), int(D.view(torch.int32).long().sum().item()))
        ref = cs if ref is None else ref
        print("%-5s walkers %2d: %s %s %s" % (name, w, {k: round(v, 2) for k, v in t.items()}, "same" if cs == ref else "DIFFERENT",
                                             {k: ix.last_stats()[k] for k in ("work_units", "segments", "fallback_queries")}), flush=True)
    L.debug_set("WALKERS", 0)
    del ix
