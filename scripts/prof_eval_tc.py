"""configs[0] trimmed evaluation, tensor-core form, a few calls (for an ncu launch list)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, vqwild_b200 as vq
dev = torch.device("cuda", 0)
g0, label0, video0, qrow0 = bench.cfg0_problem(np)
gd = torch.from_numpy(g0).to(dev)
ix0 = vq.IndexFlatL2(512)
ix0.add(gd)
ix0.set_meta(torch.from_numpy(label0).to(dev), torch.from_numpy(video0).to(dev))
qd = gd[torch.from_numpy(qrow0).to(dev)].contiguous()
ql0 = torch.from_numpy(label0[qrow0]).to(dev)
ig0 = torch.from_numpy(video0[qrow0][:, None].astype(np.int32)).to(dev)
for _ in range(3):
    ap = ix0.eval_trimmed(qd, ql0, ig0)[0]
torch.cuda.synchronize()
print("mAP", float(ap.double().mean()), ix0.last_stats())
