"""Where does the e2e arm's extra time go?  Times the pieces of one step with host clocks (n = 1M rows)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import vqwild_b200 as vq
dev = torch.device("cuda", 0)
nq, n, k = 100_000, (int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20), 100
xq = bench.gen_rows(torch, "q", 0, nq, 512, dev)
g = bench.gen_rows(torch, "g", 0, n, 512, dev)
ix = vq.IndexFlatL2(512); ix.add(g)
D = torch.empty((nq, k), dtype=torch.float32, device=dev); I = torch.empty((nq, k), dtype=torch.int64, device=dev)
xh = torch.empty(xq.shape, dtype=torch.bfloat16, pin_memory=True); xh.copy_(xq)
Dh = torch.empty((nq, k), dtype=torch.float32, pin_memory=True); Ih = torch.empty((nq, k), dtype=torch.int64, pin_memory=True)
def T(name, fn, reps=4):
    for i in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) * 1e3
    print("%-34s %.1f ms" % (name, dt), flush=True)
T("search device in/out", lambda: ix.search(xq, k, out=(D, I)))
T("h2d queries (torch)", lambda: xh.to(dev, non_blocking=True))
T("d2h results (torch)", lambda: (Dh.copy_(D, non_blocking=True), Ih.copy_(I, non_blocking=True)))
T("search host in, device out", lambda: ix.search(xh, k, out=(D, I)))
T("search device in, host out", lambda: ix.search(xq, k, out=(Dh, Ih)))
T("search host in/out (C-ABI e2e)", lambda: ix.search(xh, k, out=(Dh, Ih)))
print(ix.last_timing())
T("search device in/out (again)", lambda: ix.search(xq, k, out=(D, I)))
print(ix.last_timing())
