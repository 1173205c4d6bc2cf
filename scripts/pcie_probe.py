"""Host<->device copy bandwidth from pinned memory (what bounds bench.py's e2e arm)."""
import torch
n = 128 << 20
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for name, fn in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        fn()
    e1.record(); torch.cuda.synchronize()
    print("%s pinned: %.2f GB/s" % (name, 5 * n / e0.elapsed_time(e1) / 1e6))
h64 = torch.empty(120_000_000 // 8, dtype=torch.int64, pin_memory=True)
d64 = torch.empty_like(h64, device="cuda")
import time
torch.cuda.synchronize(); t0 = time.perf_counter(); h64.copy_(d64); torch.cuda.synchronize(); print("d2h 120MB wall %.1f ms" % ((time.perf_counter() - t0) * 1e3))
