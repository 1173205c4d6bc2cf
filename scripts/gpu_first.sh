#!/bin/bash
# First-contact GPU script: SIMT paths, then the tcgen05 path under a hard timeout, then pytest.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== exact path ==" > gpurun_out/first.log
timeout -s KILL 300 python - >> gpurun_out/first.log 2>&1 <<'PY'
import numpy as np, vqwild_b200 as vq
from oracle import oracle as orc
rng = np.random.default_rng(0)
g = rng.standard_normal((3000, 128)).astype(np.float32); q = rng.standard_normal((70, 128)).astype(np.float32)
ix = vq.IndexFlatL2(128); ix.set_precision(vq.PREC_EXACT); ix.add(g)
D, I = ix.search(q, 10); Do, Io = orc.search(q, g, 10)
print("exact idx ok", np.array_equal(I, Io), "bits ok", np.array_equal(D.view(np.uint32), Do.view(np.uint32)))
PY
echo "== tcgen05 path ==" >> gpurun_out/first.log
for prec in BF16 SPLIT3; do
timeout -s KILL 120 python - $prec >> gpurun_out/first.log 2>&1 <<'PY'
import sys, numpy as np, vqwild_b200 as vq
from oracle import oracle as orc
prec = sys.argv[1]
rng = np.random.default_rng(0)
for (nq, n, d, k) in [(70, 3000, 128, 10), (300, 9000, 512, 100)]:
    g = rng.standard_normal((n, d)).astype(np.float32); g /= np.linalg.norm(g, axis=1, keepdims=True)
    q = rng.standard_normal((nq, d)).astype(np.float32); q /= np.linalg.norm(q, axis=1, keepdims=True)
    ix = vq.IndexFlatL2(d); ix.set_precision(getattr(vq, "PREC_" + prec)); ix.add(g)
    D, I = ix.search(q, k); Do, Io = orc.search(q, g, k)
    st = ix.last_stats()
    print(prec, (nq, n, d, k), "idx ok", np.array_equal(I, Io), "bits ok", np.array_equal(D.view(np.uint32), Do.view(np.uint32)), st, flush=True)
    if not np.array_equal(I, Io):
        bad = np.where((I != Io).any(1))[0]
        print("  bad queries", len(bad), bad[:10], "first row got", I[bad[0], :8], "want", Io[bad[0], :8], D[bad[0], :4], Do[bad[0], :4])
PY
echo "exit $?" >> gpurun_out/first.log
done
echo "== pytest ==" >> gpurun_out/first.log
timeout -s KILL 1500 python -m pytest tests -q -m gpu -x --timeout=600 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/first.log
tail -5 gpurun_out/pytest_gpu.log >> gpurun_out/first.log
cat gpurun_out/first.log
