import os, sys, ctypes as C
sys.path.insert(0, "/root/repo")
import torch, vqwild_b200 as vq
from vqwild_b200 import _lib as L
os.environ["VQ_DEBUG_TRIPLET_CLOCKS"] = "1"
lib = vq.lib()
for G in (1, 148, 1480):
    emb = torch.randn(G, 256, 512, device="cuda"); emb = emb / emb.norm(dim=2, keepdim=True)
    lab = torch.arange(256, device="cuda").div(4, rounding_mode="floor").to(torch.int32).repeat(G, 1).contiguous()
    loss = torch.empty(G, device="cuda"); grad = torch.empty_like(emb); trip = torch.empty((G, 256, 3), dtype=torch.int32, device="cuda")
    for _ in range(3):
        print("G", G, file=sys.stderr)
        L.check(lib.vq_triplet_batchhard(L.ptr(emb), L.ptr(lab), C.c_int64(G), C.c_int64(256), C.c_int(512), C.c_float(1.0), L.ptr(loss), L.ptr(grad), L.ptr(trip), C.c_void_p(0)))
        torch.cuda.synchronize()
