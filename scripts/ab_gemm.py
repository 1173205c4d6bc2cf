"""A/B of gemm_topk variants on ONE box (box-to-box clock differences under the power cap are ~5-8%):
builds the configs[3] index once and alternates debug switches between searches.
Usage: python scripts/ab_gemm.py [n] [nq] [rounds]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
import vqwild_b200 as vq

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
rounds = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda", 0)
xq = bench.gen_rows(torch, "q", 0, nq, 512, dev)
ix = vq.IndexFlatL2(512)
ix.add(bench.gen_rows(torch, "g", 0, n, 512, dev))
D = torch.empty((nq, 100), dtype=torch.float32, device=dev)
I = torch.empty((nq, 100), dtype=torch.int64, device=dev)
variants = {
    "dense8(default)": {},
    "nodense": {"DENSE_ROWS": 33},
    "dense2": {"DENSE_ROWS": 2},
    "dense3": {"DENSE_ROWS": 3},
    "dense4": {"DENSE_ROWS": 4},
    "dense16": {"DENSE_ROWS": 16},
    "nochain+nodense": {"NO_CHAIN": 1, "DENSE_ROWS": 33},
}
keys = {"NO_CHAIN": 0, "DENSE_ROWS": -1}
for _ in range(3):
    ix.search(xq, 100, out=(D, I))
res = {k: [] for k in variants}
ref = None
for r in range(rounds):
    for name, env in variants.items():
        for k, dflt in keys.items():
            vq._lib.debug_set(k, env.get(k, dflt))
        ix.search(xq, 100, out=(D, I))
        t = ix.last_timing()
        res[name].append(t["gemm_topk_ms"])
        cs = int((I.sum() % (1 << 61)).item())
        ref = cs if ref is None else ref
        assert cs == ref, (name, cs, ref)
for name, v in res.items():
    m = sum(v) / len(v)
    print("%-22s gemm_topk_ms %s mean %.1f  TFLOP/s %.1f" % (name, ["%.1f" % x for x in v], m, 2.0 * nq * n * 512 / m / 1e9))
print("stats", ix.last_stats())
