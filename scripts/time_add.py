"""vq_index_add (a2: prepare_rows_kernel) on device-resident rows: 4 M x 512, bf16 and f32."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, vqwild_b200 as vq
dev = torch.device("cuda", 0)
n, d = 4_000_000, 512
g = bench.gen_rows(torch, "g", 0, n, d, dev)
for name, x in (("bf16", g), ("f32", g.float())):
    for it in range(3):
        ix = vq.IndexFlatL2(d)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        ix.add(x)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        del ix
    planes = 2 if name == "f32" else 1
    nbytes = n * d * x.element_size() + planes * n * (d + 16) * 2 + n * d * x.element_size()   # read + planes + raw copy
    print("index.add %s %d x %d: %.2f ms (%.0f GB/s incl. the raw-row copy)" % (name, n, d, ms, nbytes / ms / 1e6), flush=True)
