"""One launch shape of the fused triplet kernel for `ncu --set full --import-source on -k regex:triplet_tc` (G = 1480,
B = 256, D = 512: ten groups per SM)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import vqwild_b200 as vq
from vqwild_b200 import _lib as L

G = int(sys.argv[1]) if len(sys.argv) > 1 else 1480
lib = vq.lib()
emb = torch.randn(G, 256, 512, device="cuda")
emb = emb / emb.norm(dim=2, keepdim=True)
lab = torch.arange(256, device="cuda").div(4, rounding_mode="floor").to(torch.int32).repeat(G, 1).contiguous()
loss = torch.empty(G, device="cuda")
grad = torch.empty_like(emb)
trip = torch.empty((G, 256, 3), dtype=torch.int32, device="cuda")
for _ in range(3):
    L.check(lib.vq_triplet_batchhard(L.ptr(emb), L.ptr(lab), C.c_int64(G), C.c_int64(256), C.c_int(512), C.c_float(1.0),
                                     L.ptr(loss), L.ptr(grad), L.ptr(trip), C.c_void_p(0)))
torch.cuda.synchronize()
print("loss mean", float(loss.mean()), "handed back", lib.vq_triplet_last_handed_back())
