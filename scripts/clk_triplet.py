"""Phase clocks of the fused triplet kernel (VQ_DEBUG_TRIPLET_CLOCKS); `nograd` as argv[1] skips the gradient."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import vqwild_b200 as vq
from vqwild_b200 import _lib as L

os.environ["VQ_DEBUG_TRIPLET_CLOCKS"] = "1"
nograd = len(sys.argv) > 1 and sys.argv[1] == "nograd"
lib = vq.lib()
for G in (1, 148, 1480):
    emb = torch.randn(G, 256, 512, device="cuda")
    emb = emb / emb.norm(dim=2, keepdim=True)
    lab = torch.arange(256, device="cuda").div(4, rounding_mode="floor").to(torch.int32).repeat(G, 1).contiguous()
    loss = torch.empty(G, device="cuda")
    grad = torch.empty_like(emb)
    trip = torch.empty((G, 256, 3), dtype=torch.int32, device="cuda")
    for _ in range(3):
        print("G", G, "nograd" if nograd else "", file=sys.stderr)
        L.check(lib.vq_triplet_batchhard(L.ptr(emb), L.ptr(lab), C.c_int64(G), C.c_int64(256), C.c_int(512), C.c_float(1.0),
                                         L.ptr(loss), C.c_void_p(0) if nograd else L.ptr(grad), L.ptr(trip), C.c_void_p(0)))
        torch.cuda.synchronize()
