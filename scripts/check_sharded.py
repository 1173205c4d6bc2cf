"""torchrun --nproc-per-node G scripts/check_sharded.py : row-sharded search (one NCCL all-gather +
merge) must equal the single-index search bit for bit, duplicates across shards included."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import bench
import vqwild_b200 as vq

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n, nq, d, k = 400_003, 3000, 512, 100
g = bench.gen_rows(torch, "g", 0, n, d, dev)
g[n - 5] = g[17]                      # duplicate rows on different shards: tie -> lowest global row
xq = bench.gen_rows(torch, "q", 0, nq, d, dev)
xq[0] = g[17]
r0, r1 = rank * n // world, (rank + 1) * n // world
shard = vq.IndexFlatL2(d)
shard.add(g[r0:r1].contiguous())
D, I = vq.sharded_search(shard, xq, k, r0)
full = vq.IndexFlatL2(d)
full.add(g)
Df, If = full.search(xq, k)
ok = torch.equal(I, If) and torch.equal(D.view(torch.int32), Df.view(torch.int32))
# the two-round (head) exchange, forced at this world size: final when complete, full round otherwise -- same answer
from vqwild_b200 import ops
Dl, Il = shard.search(xq, k)
for head in (80, 50):   # 50 = the mean share of a shard at 2 ranks: many heads too short -> the full round runs
    ops._head_failed.clear()
    Dh, Ih = ops.exchange_merge(Dl, Il, r0, head=head)
    ok = ok and torch.equal(Ih, If) and torch.equal(Dh.view(torch.int32), Df.view(torch.int32))
    if rank == 0:
        print("head=%d: fell back to the full round: %s" % (head, bool(ops._head_failed)))
# sharded_search with a head-only LOCAL search (forced head length at this world size): final at 80, at 50 the head
# round is refused and the search is repeated for k entries -- same answer, and the refusal is remembered
real_head_len = ops.head_len
for head, want_fail in ((80, False), (50, True)):
    ops._head_failed.clear()
    ops.head_len = lambda k_, ws_: head
    assert ops.local_k(k, world) == head
    Ds, Is = vq.sharded_search(shard, xq, k, r0, n_global=n)
    ok = ok and torch.equal(Is, If) and torch.equal(Ds.view(torch.int32), Df.view(torch.int32))
    ok = ok and bool(ops._head_failed) == want_fail and ops.local_k(k, world) == (k if want_fail else head)
    if rank == 0:
        print("head-only local search, m=%d: repeated for k entries: %s" % (head, bool(ops._head_failed)))
ops.head_len = real_head_len
ops._head_failed.clear()
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("sharded == single-index on %d ranks: %s (I[0,:3]=%s)" % (world, bool(flag.item()), I[0, :3].tolist()))
dist.destroy_process_group()
sys.exit(0 if flag.item() else 1)
