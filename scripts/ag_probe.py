"""Exchange probe (torchrun, N ranks): what does ONE all-gather of the packed head keys cost at this world size, and what
would reading the peers' buffers over NVLink peer memory cost instead?  Sizes = MB per rank (32 = 100 000 queries x 40
keys x 8 B, the head exchange at 8 ranks; 80 = full lists).  CUDA events on the launching stream, ranks aligned by a
barrier first, median of 10, max over ranks."""
import os
import subprocess
import sys

import torch
import torch.distributed as dist

rank, ws, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
tag = os.environ.get("PROBE_TAG", "default")
if rank == 0:
    print("== %s  world %d  nccl %s  env %s" % (tag, ws, torch.cuda.nccl.version(),
                                                {k: v for k, v in os.environ.items() if k.startswith("NCCL_")}), flush=True)
    if tag == "default":
        print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout[:3000], flush=True)


def timed(fn, reps=10):
    ts = []
    for _ in range(reps):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = torch.tensor([sorted(ts)[len(ts) // 2], min(ts)], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


for mb in (1, 32, 80):
    n = mb * 2 ** 20 // 8
    x = torch.full((n,), rank, dtype=torch.int64, device="cuda")
    out = torch.empty((ws, n), dtype=torch.int64, device="cuda")
    for _ in range(3):
        dist.all_gather_into_tensor(out, x)
    med, mn = timed(lambda: dist.all_gather_into_tensor(out, x))
    if rank == 0:
        print("nccl all_gather_into_tensor %3d MB/rank: median %.3f ms  min %.3f ms  (%.0f GB/s received per rank)"
              % (mb, med, mn, (ws - 1) * mb * 2 ** 20 / med / 1e6), flush=True)
    assert bool((out[ws - 1] == ws - 1).all())

if tag == "default":
    try:
        import torch.distributed._symmetric_memory as symm
        for mb in (1, 32, 80):
            n = mb * 2 ** 20 // 8
            buf = symm.empty(n, dtype=torch.int64, device="cuda")
            hdl = symm.rendezvous(buf, dist.group.WORLD)
            buf.fill_(rank)
            out = torch.empty((ws, n), dtype=torch.int64, device="cuda")
            peers = [hdl.get_buffer(p, (n,), torch.int64) for p in range(ws)]

            def pull():
                hdl.barrier()
                for p in range(ws):
                    out[(rank + p) % ws].copy_(peers[(rank + p) % ws])
                hdl.barrier()
            for _ in range(3):
                pull()
            med, mn = timed(pull)
            if rank == 0:
                print("peer-memory pull (symm_mem, %d copies + 2 barriers) %3d MB/rank: median %.3f ms  min %.3f ms  multicast %s"
                      % (ws, mb, med, mn, getattr(hdl, "multicast_ptr", None) not in (None, 0)), flush=True)
            assert bool((out[ws - 1] == ws - 1).all())
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print("symmetric memory unavailable:", repr(e)[:400], flush=True)
dist.destroy_process_group()
