"""Device timing of the batch-hard triplet loss (fwd+bwd fused in one C-ABI call): configs[1] B=256 D=512,
G=1 and G=4096, fused tensor-core kernel vs the CUDA-core path (VQ_DEBUG_NO_TC_TRIPLET)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import vqwild_b200 as vq
from vqwild_b200 import _lib as L

dev = "cuda"
lib = vq.lib()
for G in [int(x) for x in os.environ.get("TT_G", "1,148,4096").split(",")]:  # TT_G=4096: that case only
    emb = torch.randn(G, 256, 512, device=dev)
    emb = emb / emb.norm(dim=2, keepdim=True)
    lab = torch.arange(256, device=dev).div(4, rounding_mode="floor").to(torch.int32).repeat(G, 1).contiguous()
    loss = torch.empty(G, device=dev)
    grad = torch.empty_like(emb)
    trip = torch.empty((G, 256, 3), dtype=torch.int32, device=dev)
    outs = {}
    for mode in ("tc", "simt"):
        if mode == "simt":
            if G > 148:
                continue
            L.debug_set("NO_TC_TRIPLET", 1)
        else:
            L.debug_set("NO_TC_TRIPLET", 0)

        def run():
            L.check(lib.vq_triplet_batchhard(L.ptr(emb), L.ptr(lab), C.c_int64(G), C.c_int64(256), C.c_int(512),
                                             C.c_float(1.0), L.ptr(loss), L.ptr(grad), L.ptr(trip), C.c_void_p(0)))
        for _ in range(3):
            run()
        hb = lib.vq_triplet_last_handed_back()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20 if G < 4096 else 10
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            run()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        nbytes = G * (2 * 256 * 512 * 4 + 256 * 8 + 4)
        outs[mode] = (loss.clone(), grad.clone(), trip.clone())
        print("G=%d %s: %.4f ms  %.1f GB/s algorithmic  handed_back=%d" % (G, mode, ms, nbytes / ms / 1e6, hb), flush=True)
    if len(outs) == 2:
        same = all(torch.equal(a, b) for a, b in zip(outs["tc"], outs["simt"]))
        print("   tc == simt bit for bit:", same)
