"""eval_trimmed (configs[0] gallery) at the per-rank query counts of the sharded evaluation, before and after a large
search ran in the same process (the N = 8 bench line showed 6.4 ms for 512 queries per rank)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
import vqwild_b200 as vq

dev = torch.device("cuda", 0)
g, label, video, qrow = bench.cfg0_problem(np)
ix = vq.IndexFlatL2(g.shape[1])
gd = torch.from_numpy(g).to(dev)
ix.add(gd)
ix.set_meta(torch.from_numpy(label).to(dev), torch.from_numpy(video).to(dev))
qd = gd[torch.from_numpy(qrow).to(dev)].contiguous()
ql = torch.from_numpy(label[qrow]).to(dev)
igd = torch.from_numpy(video[qrow][:, None].astype(np.int32)).to(dev)


def run(tag):
    for nq in (4096, 512, 1024, 2048):
        ts = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            ap, rec, _ = ix.eval_trimmed(qd[:nq].contiguous(), ql[:nq].contiguous(), igd[:nq].contiguous())
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        print("%s nq=%4d: %s ms  handed back %s  mAP %.6f" % (tag, nq, ["%.2f" % t for t in ts], ix.last_stats()["segments"],
                                                            float(ap.double().mean())), flush=True)


run("fresh   ")
big = vq.IndexFlatL2(512)
big.add(bench.gen_rows(torch, "g", 0, 1_250_000, 512, dev))
xq = bench.gen_rows(torch, "q", 0, 100000, 512, dev)
for _ in range(2):
    big.search(xq, 40)
torch.cuda.synchronize()
run("after   ")
del big
torch.cuda.synchronize()
run("big gone")
