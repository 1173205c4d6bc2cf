"""Exact mode (VQ_PREC_EXACT: canonical CUDA-core streaming top-k, also the fallback of unproven queries) vs the default
tensor-core pre-filter path on the same index.  Usage: python scripts/time_exact.py [nq] [n]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, vqwild_b200 as vq
nq = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1310720
d, dev = 512, torch.device("cuda", 0)
xq = bench.gen_rows(torch, "q", 0, nq, d, dev)
g = bench.gen_rows(torch, "g", 0, n, d, dev)
ix = vq.IndexFlatL2(d); ix.add(g)
res = {}
for name, prec in (("default", None), ("exact", 3)):
    if prec is not None:
        ix.set_precision(prec)
    for _ in range(2):
        D, I = ix.search(xq, 100)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    D, I = ix.search(xq, 100)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    res[name] = (D.clone(), I.clone())
    print("%-8s nq=%d n=%d: %.2f ms  %.3g pairs/s  %s" % (name, nq, n, dt * 1e3, nq * n / dt, ix.last_timing()), flush=True)
print("identical:", bool((res["default"][1] == res["exact"][1]).all()) and bool((res["default"][0] == res["exact"][0]).all()))
