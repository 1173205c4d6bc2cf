"""A/B of epilogue variants on one box, interleaved: gemm_topk time of the N=8-shaped launch (100k x 1.25M) under the
debug hooks.  NO_INSERT is a timing probe (no candidate handling at all = the floor of the kernel at this shape)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import vqwild_b200 as vq
from vqwild_b200 import _lib as L

nq = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1250000
rounds = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda", 0)
xq = bench.gen_rows(torch, "q", 0, nq, 512, dev)
ix = vq.IndexFlatL2(512)
ix.add(bench.gen_rows(torch, "g", 0, n, 512, dev))
D = torch.empty((nq, 100), dtype=torch.float32, device=dev)
I = torch.empty((nq, 100), dtype=torch.int64, device=dev)
dflt = {"NO_CHAIN": 0, "DENSE_ROWS": -1, "NO_INSERT": 0, "KPRIME": 0, "TRIGGER": 0}
variants = {
    "default": {},
    "no_insert(floor)": {"NO_INSERT": 1},
    "trigger384": {"TRIGGER": 384},
    "trigger512": {"TRIGGER": 512},
    "trigger768": {"TRIGGER": 768},
    "k112": {"KPRIME": 112},
    "k112+trig384": {"KPRIME": 112, "TRIGGER": 384},
    "kprime104": {"KPRIME": 104},
}
for _ in range(3):
    ix.search(xq, 100, out=(D, I))
res = {k: [] for k in variants}
fb = {}
sel = {k: [] for k in variants}
for r in range(rounds):
    for name, hooks in variants.items():
        for k, v in dflt.items():
            L.debug_set(k, hooks.get(k, v))
        ix.search(xq, 100, out=(D, I))
        t = ix.last_timing()
        res[name].append(t["gemm_topk_ms"])
        fb.setdefault(name, []).append(ix.last_stats()["fallback_queries"])
        sel[name].append(t["select_ms"])
for k, v in dflt.items():
    L.debug_set(k, v)
for name, v in res.items():
    m = sum(v) / len(v)
    print("%-18s gemm_topk_ms %s mean %.1f  TFLOP/s %.1f  select %.2f" % (name, ["%.1f" % x for x in v], m, 2.0 * nq * n * 512 / m / 1e9,
                                                                        sum(sel[name]) / len(sel[name])), "fallback", fb[name])
print("stats", ix.last_stats())
