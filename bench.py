#!/usr/bin/env python
"""bench.py -- the hot path's headline measurement (BASELINE.json `metric`): query-gallery
pairs/s on synthetic 512-d embeddings, one process per GPU.

Workload (config.workload): BASELINE.json configs[3] "large-gallery scaling": 100 000 queries x
10 000 000 clips x 512-d bf16, top-100, gallery row-sharded over the ranks (STRONG scaling: the
10 M gallery is fixed, each of N ranks holds 10M/N rows), per-rank top-k merged after ONE NCCL
all-gather.  It fits one B200 (20.5 GB of gallery planes), so it is also the N=1 workload.

A step = one pass of the matching path over all queries: vq_index_search on the local shard
(query operand preparation -> tcgen05 GEMM + streaming top-k -> canonical re-score / verify ->
exact fallback for unproven queries) and, for N>1, all-gather + merge.  `value` is timed with the
queries resident in HBM; `e2e` is the same step through the C-ABI with HOST (pinned) query and
result buffers, copies inside the timed region.  The gallery is index state (the reference's
`index.add(xb)`, dataloader_baseline.py:1507, runs once per evaluation); its one-off cost from
host memory is reported separately as `index_add`.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    torchrun --nproc-per-node N bench.py --gpus N ...

--impl reference: the reference's CPU path for the same metric.  The reference is pure Python
(nothing compiles into oracle/_ref) and is absent on the GPU box, so this arm times the C port
of its matching stage (oracle/vq_oracle.c: canonical squared-L2 + full sort per query, OpenMP
over all host threads) on a bounded sample of the workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SEED = 0
BLOCK_ROWS = 1 << 20  # synthetic data is generated in blocks seeded by (SEED, block id)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    # the named workload; overriding any of these makes config.workload say "custom"
    ap.add_argument("--nq", type=int, default=100_000)
    ap.add_argument("--n", type=int, default=10_000_000)
    ap.add_argument("--d", type=int, default=512)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the 64-query oracle check of the full-size result")
    ap.add_argument("--no-aux", action="store_true", help="skip the cfg-1 mAP-eval and cfg-2 loss side measurements")
    return ap.parse_args()


def workload_name(a):
    if (a.nq, a.n, a.d, a.k) == (100_000, 10_000_000, 512, 100):
        return "configs[3] large-gallery scaling: 100k queries x 10M clips x 512-d bf16, top-100"
    return "custom: %d queries x %d clips x %d-d bf16, top-%d" % (a.nq, a.n, a.d, a.k)


# ---------------------------------------------------------------------------------------------
# synthetic embeddings: x = normalize(randn) * u, u ~ U(0.6, 1): pooled features
# have norm <= 1, not 1, so squared-L2 and cosine rankings differ.
# ---------------------------------------------------------------------------------------------
def gen_block(torch, kind, block, rows, d, device):
    g = torch.Generator(device=device)
    g.manual_seed(SEED * 1_000_003 + (0 if kind == "g" else 500_000) + block)
    x = torch.randn(rows, d, generator=g, device=device, dtype=torch.float32)
    x = x / x.norm(dim=1, keepdim=True)
    u = torch.rand(rows, 1, generator=g, device=device, dtype=torch.float32) * 0.4 + 0.6
    return (x * u).to(torch.bfloat16)


def gen_rows(torch, kind, r0, r1, d, device):
    """Rows [r0, r1) of the synthetic matrix `kind`; identical whatever the sharding."""
    out = torch.empty((r1 - r0, d), dtype=torch.bfloat16, device=device)
    b = r0 // BLOCK_ROWS
    while b * BLOCK_ROWS < r1:
        lo, hi = b * BLOCK_ROWS, (b + 1) * BLOCK_ROWS
        blk = gen_block(torch, kind, b, BLOCK_ROWS, d, device)
        s, e = max(lo, r0), min(hi, r1)
        out[s - r0:e - r0] = blk[s - lo:e - lo]
        del blk
        b += 1
    return out


# ---------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, dev_index):
        self.rows, self.proc, self.th = [], None, None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(dev_index), "--query-gpu=" + self.FIELDS,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, pw, reasons = [], None, [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ts, line in self.rows:
            if ts < t0 or ts > t1 + 0.1:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax = float(f[1])
                pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v == "Active":
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": smax,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
# the CPU arm: C port of the reference's matching stage on a bounded sample
# ---------------------------------------------------------------------------------------------
def cpu_sample(a, steps, warmup):
    """Returns (pairs/s, cores, sample description, ms/step)."""
    import numpy as np
    import torch
    from oracle import oracle as orc
    # all the host threads this process may use: torchrun exports OMP_NUM_THREADS=1 to its ranks, which would time
    # the CPU arm on one core at N>1 (VQ_CPU_THREADS overrides)
    want = int(os.environ.get("VQ_CPU_THREADS", "0")) or len(os.sched_getaffinity(0))
    orc.set_num_threads(want)
    cores = orc.num_threads()
    n_s = min(a.n, 1 << 18)
    nq_s = min(a.nq, 32 * cores)  # ~7 s of work per step on 16 threads
    g = gen_block(torch, "g", 0, n_s, a.d, "cpu").float().numpy()
    q = gen_block(torch, "q", 0, nq_s, a.d, "cpu").float().numpy()
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        D, I = orc.search(q, g, a.k)  # canonical d2 to every row + full sort, as faiss k=N (:1592-1597)
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    assert (np.diff(D, axis=1) >= 0).all()
    dt = sum(times) / len(times)
    sample = ("%d queries x first %d gallery rows x %d-d (bf16 values widened to f32), %d steps; what the reference's faiss call "
              "does -- squared L2 to EVERY row + a FULL sort per query (k = ntotal, :1594), then the first %d are read -- with the "
              "oracle's sequential-fmaf distances (not BLAS): a baseline for context, not a like-for-like top-%d CPU kernel") % (
        nq_s, n_s, a.d, len(times), a.k, a.k)
    return nq_s * n_s / dt, cores, sample, dt * 1e3


def parity_subsample(torch, a, Dm, Im, nsub=64):
    """Outside the timed region: a subsample of the queries of the FULL-size result (after the merge when N > 1)
    against the CPU oracle.  The gallery is regenerated block by block (1 Mi rows, the same seeded generator every
    sharding uses), each block searched by the oracle (canonical d2 to every row of the block, exact top-k), the
    per-block lists merged by the oracle's merge.  Indices exact, distances bit-equal."""
    import numpy as np
    from oracle import oracle as orc
    t0 = time.perf_counter()
    orc.set_num_threads(int(os.environ.get("VQ_CPU_THREADS", "0")) or len(os.sched_getaffinity(0)))
    rs = np.random.RandomState(12345)
    sub = np.sort(rs.choice(a.nq, min(nsub, a.nq), replace=False))
    dev = Dm.device
    xq = gen_rows(torch, "q", 0, a.nq, a.d, dev)[torch.from_numpy(sub).to(dev)].float().cpu().numpy()
    k = a.k
    parts_d, parts_i = [], []
    for b0 in range(0, a.n, BLOCK_ROWS):
        b1 = min(a.n, b0 + BLOCK_ROWS)
        blk = gen_rows(torch, "g", b0, b1, a.d, dev).float().cpu().numpy()
        kk = min(k, b1 - b0)
        d_, i_ = orc.search(xq, blk, kk)
        d_ = np.pad(d_, ((0, 0), (0, k - kk)), constant_values=np.inf)
        i_ = np.pad(i_, ((0, 0), (0, k - kk)), constant_values=-1)
        parts_d.append(d_)
        parts_i.append(np.where(i_ >= 0, i_ + b0, i_))
        del blk
    Do, Io = orc.topk_merge(np.stack(parts_d), np.stack(parts_i))
    sel = torch.from_numpy(sub).to(dev)
    Dg, Ig = Dm[sel].cpu().numpy(), Im[sel].cpu().numpy()
    ok = bool(np.array_equal(Ig, Io) and np.array_equal(Dg.view(np.uint32), np.asarray(Do, np.float32).view(np.uint32)))
    return {"queries": int(len(sub)), "ok": ok, "checked": "top-%d rows exact + d2 bit-equal vs the CPU oracle over all %d gallery rows"
            % (k, a.n), "mismatching_queries": int((Ig != Io).any(axis=1).sum()), "seconds": time.perf_counter() - t0}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    v, cores, sample, ms = cpu_sample(a, max(1, a.steps), max(0, min(a.warmup, 1)))
    emit({
        "impl": "reference", "metric": "query-gallery pairs/sec (512-d)", "value": v, "unit": "pairs/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(a), "nq": a.nq, "n": a.n, "d": a.d, "k": a.k},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference is pure Python and absent on the GPU box; this is the OpenMP C port of its matching stage "
                "(oracle/vq_oracle.c), a faster baseline than the reference's own per-query Python loop",
    })


# ---------------------------------------------------------------------------------------------
# side measurements on rank 0 at N=1 (the second half of BASELINE.json's metric, and configs[1])
# ---------------------------------------------------------------------------------------------
def aux_measurements(torch, vq):
    import numpy as np
    dev = "cuda"
    out = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            pk = json.load(f)
        peaks_hbm, peaks_tc = pk.get("hbm_gbs"), pk.get("bf16_tflops")
    except (OSError, ValueError):
        peaks_hbm, peaks_tc = None, None
    peaks_tc = peaks_tc or 1590.0

    def tc_roof(flops, ms, note):
        ach = flops / (ms * 1e-3) / 1e12
        return {"bound": "tensor", "achieved": ach, "peak": peaks_tc, "unit": "TFLOP/s", "frac": ach / peaks_tc,
                "algorithmic_flops": flops, "note": note}
    # configs[0]: trimmed retrieval eval, 4096 x 20480 x 512 f32, full-list AP + R@N
    rng = np.random.default_rng(0)
    n, nq, d = 20480, 4096, 512
    label = rng.integers(0, 180, n).astype(np.int32)
    centre = rng.standard_normal((180, d)).astype(np.float32)
    g = rng.standard_normal((n, d)).astype(np.float32) + centre[label]
    g /= np.linalg.norm(g, axis=1, keepdims=True)
    g *= rng.uniform(0.6, 1.0, (n, 1)).astype(np.float32)
    video = np.arange(n, dtype=np.int32)
    qrow = rng.choice(n, nq, replace=False)
    ign = video[qrow][:, None].astype(np.int32)
    ix = vq.IndexFlatL2(d)
    gd = torch.from_numpy(g).to(dev)
    ix.add(gd)
    ix.set_meta(torch.from_numpy(label).to(dev), torch.from_numpy(video).to(dev))
    qd = gd[torch.from_numpy(qrow).to(dev)].contiguous()
    ql = torch.from_numpy(label[qrow]).to(dev)
    igd = torch.from_numpy(ign).to(dev)
    ts = []
    for it in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ap, rec, _ = ix.eval_trimmed(qd, ql, igd)
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    out["map_eval_cfg0"] = {"workload": "configs[0]: 4096 queries x 20480 gallery x 512-d f32, full-list AP + R@30/50/100",
                            "wall_ms": min(ts[1:]) * 1e3, "mAP": float(ap.double().mean().item()),
                            "pairs_per_s": nq * n / min(ts[1:]),
                            "handed_back_to_canonical": ix.last_stats()["segments"],
                            "roofline": tc_roof(2.0 * nq * n * d, min(ts[1:]) * 1e3,
                                                "2*Q*N*D against the burst bf16 peak as SURVEY.md 8d scores it, WHOLE evaluation "
                                                "(3-pass split GEMM in STORE mode 0.31 ms = 3x these flops issued, then "
                                                "ap_rows_tc: gathers + rank counting, not tensor work; DESIGN.md 4.4)")}
    # the same evaluation by the CPU port (oracle, all host threads) on a bounded sample of the queries
    try:
        from oracle import oracle as orc
        ns = 4 * orc.num_threads()
        t0 = time.perf_counter()
        apc, _, _, _ = orc.eval_trimmed(g[qrow[:ns]], g, label, video, label[qrow[:ns]], ign[:ns])
        dt = time.perf_counter() - t0
        out["map_eval_cfg0"]["cpu_port"] = {"sample": "%d of the %d queries x full gallery, %d threads" % (ns, nq, orc.num_threads()),
                                            "wall_ms_sample": dt * 1e3, "wall_ms_extrapolated": dt * 1e3 * nq / ns,
                                            "max_abs_ap_diff_vs_gpu": float(np.abs(apc - ap[:ns].double().cpu().numpy()).max())}
    except Exception as e:  # noqa: BLE001
        out["map_eval_cfg0"]["cpu_port"] = {"error": repr(e)}
    del ix, gd
    # configs[1]: batch-hard triplet fwd+bwd, B=256 D=512 (latency bound) and 4096 such groups (HBM bound)
    from vqwild_b200.loss import TripletLoss
    crit = TripletLoss(margin=1.0, sampling_method="batchhard")
    for G in (1, 4096):
        emb = torch.randn(G, 256, 512, device=dev)
        emb = (emb / emb.norm(dim=2, keepdim=True)).requires_grad_(True)
        lab = torch.arange(256, device=dev).div(4, rounding_mode="floor").to(torch.int32).repeat(G, 1)
        x = emb if G > 1 else emb[0].detach().requires_grad_(True)
        lb = lab if G > 1 else lab[0]
        for _ in range(3):
            crit(x, lb)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20 if G == 1 else 5
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            crit(x, lb)   # forward computes loss AND the gradient (fused); backward only scales it
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        nbytes = G * (2 * 256 * 512 * 4 + 256 * 8 + 4)
        out["triplet_cfg1_G%d" % G] = {"workload": "configs[1]: batch-hard triplet fwd+bwd, %d x (256 x 512-d f32)" % G,
                                       "ms": ms, "algorithmic_bytes": nbytes, "GB_per_s": nbytes / ms / 1e6}
        if G == 1:
            try:  # the CPU port of the same group (oracle: canonical pdist + mining + loss + gradient gather)
                from oracle import oracle as orc
                e_np, l_np = x.detach().cpu().numpy(), lb.cpu().numpy()
                orc.batchhard_triplet(e_np, l_np, 1.0)
                t0 = time.perf_counter()
                orc.batchhard_triplet(e_np, l_np, 1.0)
                out["triplet_cfg1_G1"]["cpu_port_ms"] = (time.perf_counter() - t0) * 1e3
            except Exception as e:  # noqa: BLE001
                out["triplet_cfg1_G1"]["cpu_port_ms"] = repr(e)
        elif peaks_hbm:
            out["triplet_cfg1_G%d" % G]["roofline"] = {"bound": "hbm", "achieved": nbytes / ms / 1e6, "peak": peaks_hbm,
                                                      "unit": "GB/s", "frac": nbytes / ms / 1e6 / peaks_hbm}
    # configs[2]: untrimmed moment retrieval, 4096 queries x 20480 videos x 64 moments, 512-d f32: segment
    # max-pool + top-100 videos (tensor cores) and the full moment evaluation (rank all, per-video NMS, AP, R@N)
    try:
        out.update(aux_moment(torch, vq, tc_roof=tc_roof))
    except Exception as e:  # noqa: BLE001 -- a side measurement must not lose the headline line
        out["moment_cfg2_error"] = repr(e)
    # f1: sliding-moment pooling (garner_feat) at untrimmed-gallery scale, HBM-bound
    try:
        out["window_means_f1"] = aux_windows(torch, peaks_hbm)
    except Exception as e:  # noqa: BLE001
        out["window_means_f1"] = {"error": repr(e)}
    # configs[4]: 1024-d (ELMo word space) zero-shot split, 4096 x 20480 f32: exact top-100 + full-list evaluation
    try:
        out["cfg4_1024d"] = aux_cfg4(torch, vq, tc_roof)
    except Exception as e:  # noqa: BLE001
        out["cfg4_1024d"] = {"error": repr(e)}
    return out


def aux_cfg4(torch, vq, tc_roof, nq=4096, n=20480, d=1024, ncls=140):
    """BASELINE.json configs[4]: unit-norm 1024-d embeddings, 140 novel classes; top-100 search (3-pass split-bf16 GEMM,
    single-CTA kernel: the padded dimension exceeds the CTA-pair kernel's 512) and the full-list AP / R@N evaluation."""
    dev = "cuda"
    gen = torch.Generator(device=dev).manual_seed(4)
    label = torch.randint(0, ncls, (n,), device=dev, generator=gen, dtype=torch.int32)
    centre = torch.randn(ncls, d, device=dev, generator=gen)
    g = torch.randn(n, d, device=dev, generator=gen) + centre[label.long()]
    g = g / g.norm(dim=1, keepdim=True)
    qrow = torch.randperm(n, device=dev, generator=gen)[:nq]
    q = g[qrow].contiguous()
    video = torch.arange(n, device=dev, dtype=torch.int32)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    ix.set_meta(label, video)
    ql, ign = label[qrow].contiguous(), video[qrow][:, None].contiguous()

    def timed(fn, reps=5):
        for _ in range(2):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            r = fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps, r
    ms_s, (D, I) = timed(lambda: ix.search(q, 100))
    t = ix.last_timing()
    ms_e, (ap, rec, _) = timed(lambda: ix.eval_trimmed(q, ql, ign))
    flops = 2.0 * nq * n * d
    return {"workload": "configs[4]: %d queries x %d gallery x %d-d f32, top-100 search and full-list AP + R@30/50/100" % (nq, n, d),
            "search_ms": ms_s, "gemm_topk_ms": t["gemm_topk_ms"], "select_ms": t["select_ms"],
            "self_match_first": bool((I[:, 0] == qrow).all().item()), "eval_ms": ms_e, "mAP": float(ap.double().mean().item()),
            "handed_back_to_canonical": ix.last_stats()["segments"],
            "roofline": tc_roof(flops, t["gemm_topk_ms"], "top-100 GEMM launch alone, 2*Q*N*D against the burst bf16 peak; the "
                                                         "3-pass split issues 3x these flops")}


def aux_windows(torch, peaks_hbm, V=2048, T=600, C=512):
    import numpy as np
    from vqwild_b200 import ops
    from vqwild_b200.retrieval import moment_grid
    s, e, lo, hi = moment_grid(T / 3.0 + 0.4, T)
    nw = len(s)
    t0, ws = np.arange(V + 1, dtype=np.int64) * T, np.arange(V + 1, dtype=np.int64) * nw
    feat = torch.rand(C, V * T, device="cuda")
    lo_d = torch.from_numpy(np.tile(lo, V).astype(np.int32)).cuda()
    hi_d = torch.from_numpy(np.tile(hi, V).astype(np.int32)).cuda()
    o = torch.empty((V * nw, C), dtype=torch.float32, device="cuda")   # allocated once: a 2.9 GB cudaMalloc inside the
    for _ in range(2):                                                  # timed loop is what the number would show
        ops.window_means(feat, t0, ws, lo_d, hi_d, out=o)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        ops.window_means(feat, t0, ws, lo_d, hi_d, out=o)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    nbytes = 4.0 * C * V * T + 4.0 * C * V * nw
    # the reference's host loop (np.mean per window, dataloader_baseline.py:1140-1143) on a few videos
    fh = feat[:, :4 * T].cpu().numpy()
    tc = time.perf_counter()
    for v in range(4):
        for a, z in zip(lo, hi):
            np.mean(fh[:, v * T + a:v * T + z], axis=1)
    cpu_ms = (time.perf_counter() - tc) * 1e3 / 4 * V
    r = {"workload": "f1: %d videos x %d frames x %d-d, %d sliding moments per video (garner_feat grid)" % (V, T, C, nw),
         "ms": ms, "algorithmic_bytes": nbytes, "GB_per_s": nbytes / ms / 1e6,
         "numpy_loop_ms_extrapolated": cpu_ms, "numpy_sample": "4 of the %d videos, 1 thread" % V}
    if peaks_hbm:
        r["roofline"] = {"bound": "hbm", "achieved": nbytes / ms / 1e6, "peak": peaks_hbm, "unit": "GB/s",
                         "frac": nbytes / ms / 1e6 / peaks_hbm}
    del feat, o
    return r


def cfg0_problem(np, n=20480, nq=4096, d=512):
    rng = np.random.default_rng(0)
    label = rng.integers(0, 180, n).astype(np.int32)
    centre = rng.standard_normal((180, d)).astype(np.float32)
    g = rng.standard_normal((n, d)).astype(np.float32) + centre[label]
    g /= np.linalg.norm(g, axis=1, keepdims=True)
    g *= rng.uniform(0.6, 1.0, (n, 1)).astype(np.float32)
    video = np.arange(n, dtype=np.int32)
    qrow = rng.choice(n, nq, replace=False)
    return g, label, video, qrow


def sharded_map_eval(torch, dist, vq, ops, dev, G):
    """configs[0] full-list AP + R@N with the queries sharded over the ranks (ops.sharded_eval): every rank holds the
    whole 42 MB gallery, evaluates Q/G queries and one all-gather returns the per-query vectors everywhere."""
    import numpy as np
    g, label, video, qrow = cfg0_problem(np)
    nq = len(qrow)
    ix = vq.IndexFlatL2(g.shape[1])
    gd = torch.from_numpy(g).to(dev)
    ix.add(gd)
    ix.set_meta(torch.from_numpy(label).to(dev), torch.from_numpy(video).to(dev))
    qd = gd[torch.from_numpy(qrow).to(dev)].contiguous()
    ql = torch.from_numpy(label[qrow]).to(dev)
    igd = torch.from_numpy(video[qrow][:, None].astype(np.int32)).to(dev)

    mid = []

    def eval_fn(q0, q1):
        ap, rec, _ = ix.eval_trimmed(qd[q0:q1].contiguous(), ql[q0:q1].contiguous(), igd[q0:q1].contiguous())
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        mid.append(e)
        return ap, rec

    best = None
    for it in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        ap, rec = ops.sharded_eval(eval_fn, nq)
        e1.record()
        torch.cuda.synchronize()
        # whole call and its local half (this rank's slice, before the all-gather), each as the max over ranks
        ms = torch.tensor([e0.elapsed_time(e1), e0.elapsed_time(mid[-1])], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if it > 1 and (best is None or float(ms[0].item()) < best[0]):
            best = ms.tolist()
    return {"workload": "configs[0]: 4096 queries x 20480 gallery x 512-d f32, full-list AP + R@30/50/100, queries sharded %d-way"
            % G, "wall_ms": best[0], "local_eval_ms": best[1], "mAP": float(ap.mean().item()),
            "pairs_per_s": nq * g.shape[0] / (best[0] * 1e-3),
            "timing": "CUDA events, max over ranks, best of 4 after 2 warm-ups; local_eval_ms = this rank's slice before the all-gather"}


def aux_moment(torch, vq, nq=4096, V=20480, S=64, d=512, tc_roof=None):
    import numpy as np
    dev, n, out = "cuda", V * S, {}
    gen = torch.Generator(device=dev).manual_seed(0)
    g = torch.randn(n, d, device=dev, generator=gen)
    g = g / g.norm(dim=1, keepdim=True) * (0.6 + 0.4 * torch.rand(n, 1, device=dev, generator=gen))
    q = torch.randn(nq, d, device=dev, generator=gen)
    q = q / q.norm(dim=1, keepdim=True)
    rng = np.random.default_rng(0)
    s_in = np.arange(n) % S  # sliding moments: 1..6 clips of 5 s at stride 5 s (dataloader_baseline.py:1116-1137)
    start = (5 * (s_in // 6)).astype(np.float32)
    end = start + 5 * (1 + s_in % 6).astype(np.float32)
    video = (np.arange(n) // S).astype(np.int32)
    hit = rng.integers(-1, 180, n).astype(np.int32)
    iou = np.where(hit >= 0, rng.uniform(0, 1, n), 0).astype(np.float32)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    gsrc = [g]
    gsrc_host = g.cpu().numpy()
    del g
    ix.set_meta(hit, video)
    ix.set_segments(np.arange(0, n + 1, S, dtype=np.int64))
    ix.set_moments(start, end, iou, hit)
    ql = torch.from_numpy(rng.integers(0, 180, nq).astype(np.int32)).to(dev)
    ign = torch.from_numpy(rng.integers(0, V, (nq, 5)).astype(np.int32)).to(dev)

    def wall(fn, reps):
        ts = []
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = fn()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        return min(ts[1:]) if reps > 1 else ts[0], r

    t, _ = wall(lambda: ix.search_segmax(q, 100), 3)
    tm = ix.last_timing()
    out["segmax_cfg2"] = {"workload": "configs[2]: %d queries x %d videos x %d moments x %d-d f32, per-video max + top-100 videos"
                          % (nq, V, S, d), "wall_ms": t * 1e3, "pairs_per_s": nq * n / t,
                          "gemm_topk_ms": tm["gemm_topk_ms"], "select_ms": tm["select_ms"],
                          "tflops_algorithmic": 2.0 * nq * n * d / tm["gemm_topk_ms"] / 1e9}
    if tc_roof:
        out["segmax_cfg2"]["roofline"] = tc_roof(2.0 * nq * n * d, tm["gemm_topk_ms"], "SEGMAX GEMM launch alone; 3-pass split-bf16 "
                                                 "issues 3x these flops")
    # the same gallery held in bf16 (north_star's bf16-input mode): CTA-pair kernel, one MMA pass
    ixb = vq.IndexFlatL2(d)
    ixb.add(torch.cat([ix_rows.to(torch.bfloat16) for ix_rows in gsrc]))
    ixb.set_segments(np.arange(0, n + 1, S, dtype=np.int64))
    qb = q.to(torch.bfloat16)
    t, _ = wall(lambda: ixb.search_segmax(qb, 100), 3)
    tm = ixb.last_timing()
    out["segmax_cfg2_bf16"] = {"workload": "configs[2] with bf16 gallery and queries", "wall_ms": t * 1e3, "pairs_per_s": nq * n / t,
                               "gemm_topk_ms": tm["gemm_topk_ms"], "select_ms": tm["select_ms"],
                               "tflops_algorithmic": 2.0 * nq * n * d / tm["gemm_topk_ms"] / 1e9}
    if tc_roof:
        out["segmax_cfg2_bf16"]["roofline"] = tc_roof(2.0 * nq * n * d, tm["gemm_topk_ms"], "SEGMAX GEMM launch alone (CTA-pair kernel)")
    del ixb, gsrc
    t, r = wall(lambda: ix.eval_moment(q, ql, ign), 2)
    cpu = None
    try:  # CPU port on a bounded sample: one query per host thread against the full moment gallery
        from oracle import oracle as orc
        ns = orc.num_threads()
        gh = gsrc_host
        t0 = time.perf_counter()
        apc, _ = orc.eval_moment(q[:ns].cpu().numpy(), gh, video, start, end, iou, hit, ql[:ns].cpu().numpy(), ign[:ns].cpu().numpy())
        dt = time.perf_counter() - t0
        cpu = {"sample": "%d of the %d queries x all %d moments, %d threads" % (ns, nq, n, ns), "wall_ms_sample": dt * 1e3,
               "wall_ms_extrapolated": dt * 1e3 * nq / ns,
               "max_abs_ap_diff_vs_gpu": float(np.abs(apc - r[0][:ns].cpu().numpy()).max())}
    except Exception as e:  # noqa: BLE001
        cpu = {"error": repr(e)}
    out["moment_eval_cfg2"] = {"cpu_port": cpu,"workload": "configs[2]: full moment evaluation (rank all %d moments per query, per-video NMS, "
                               "AP@tIoU0.5, grouped R@30/50/100), %d queries" % (n, nq), "wall_ms": t * 1e3,
                               "ms_per_query": t * 1e3 / nq, "pairs_per_s": nq * n / t, "mAP": float(r[0].mean().item())}
    if tc_roof:
        out["moment_eval_cfg2"]["roofline"] = tc_roof(2.0 * nq * n * d, t * 1e3, "whole evaluation (canonical scores on FFMA2/FADD2 CUDA-core pipes + "
                                                      "per-video NMS + AP) against the tensor peak, as SURVEY.md 8d scores it; "
                                                      "DESIGN.md 4.4 says why this one stays canonical")
    return out


# ---------------------------------------------------------------------------------------------
def run_b200(a):
    import torch
    import torch.distributed as dist
    import vqwild_b200 as vq
    from vqwild_b200 import ops

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    G = world
    if a.gpus != world and rank == 0:
        print("note: --gpus %d but WORLD_SIZE=%d; using WORLD_SIZE" % (a.gpus, world), file=sys.stderr)

    # ---- data: gallery shard [r0, r1), all queries ----
    r0, r1 = rank * a.n // G, (rank + 1) * a.n // G
    xq = gen_rows(torch, "q", 0, a.nq, a.d, dev)
    shard = gen_rows(torch, "g", r0, r1, a.d, dev)
    index = vq.IndexFlatL2(a.d)
    # one-off: index.add from pinned HOST memory (the reference's :1507) -- reported, not part of a step
    add_info = None
    if G == 1 and shard.numel() * 2 <= 24 << 30:
        host = torch.empty(shard.shape, dtype=torch.bfloat16, pin_memory=True)
        host.copy_(shard)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        index.add(host)
        torch.cuda.synchronize()
        add_info = {"ms": (time.perf_counter() - t0) * 1e3, "h2d_bytes": shard.numel() * 2}
        del host
    else:
        index.add(shard)
    del shard
    torch.cuda.synchronize()

    k = a.k
    D = torch.empty((a.nq, k), dtype=torch.float32, device=dev)
    I = torch.empty((a.nq, k), dtype=torch.int64, device=dev)

    xevents = []

    def step_device():
        if G == 1:
            index.search(xq, k, out=(D, I))
            return D, I
        # local search for the head of the shard's list (ops.local_k: 40 of 100 at 8 ranks), one all-gather, merge
        return ops.sharded_search(index, xq, k, r0, n_global=a.n, events=xevents)

    xq_host = torch.empty(xq.shape, dtype=torch.bfloat16, pin_memory=True)
    xq_host.copy_(xq)
    Dh = torch.empty((a.nq, k), dtype=torch.float32, pin_memory=True)
    Ih = torch.empty((a.nq, k), dtype=torch.int64, pin_memory=True)

    trace = os.environ.get("VQ_BENCH_TRACE")

    def step_e2e():
        if G == 1:
            t0 = time.perf_counter()
            index.search(xq_host, k, out=(Dh, Ih))    # C-ABI stages H2D / D2H on the stream and synchronises
            if trace:
                print("e2e step wall %.1f ms  phases %s" % ((time.perf_counter() - t0) * 1e3, index.last_timing()),
                      file=sys.stderr)
            return Dh, Ih
        xd = xq_host.to(dev, non_blocking=True)
        Dm, Im = ops.sharded_search(index, xd, k, r0, n_global=a.n)
        Dh.copy_(Dm, non_blocking=True)
        Ih.copy_(Im, non_blocking=True)
        torch.cuda.synchronize()
        return Dh, Ih

    def barrier():
        if G > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, collect=None):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        t0 = time.time()
        e0.record()
        for _ in range(steps):
            fn()
            if collect is not None:
                collect()
        e1.record()
        barrier()
        t1 = time.time()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if G > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), t0, t1

    for _ in range(a.warmup):
        step_device()
    gemm_ms, phase_ms = [], []

    def collect():
        t = index.last_timing()
        gemm_ms.append(t["gemm_topk_ms"])
        phase_ms.append(t)

    sampler = ClockSampler(local) if rank == 0 else None
    del xevents[:]
    total_ms, t0, t1 = timed(step_device, a.steps, collect)
    clocks = sampler.stop(t0, t1) if sampler else None
    xphase = None
    if G > 1 and xevents:   # this rank's exchange phases (CUDA events around pack / all-gather / merge), mean over the steps
        ev = [xevents[i:i + 4] for i in range(0, len(xevents) - 3, 4)]
        xphase = {nm: sum(e[j].elapsed_time(e[j + 1]) for e in ev) / len(ev)
                  for j, nm in enumerate(("pack_ms", "allgather_ms", "merge_ms"))}
        # allgather_ms on one rank includes waiting for the slowest rank's GEMM (rank skew); max - min over ranks shows it
        tt = torch.tensor([xphase["allgather_ms"], -xphase["allgather_ms"]], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        xphase["allgather_ms_max_over_ranks"], xphase["allgather_ms_min_over_ranks"] = float(tt[0].item()), -float(tt[1].item())
    stats = index.last_stats()
    out = step_device()
    checksum = int((out[1].sum() % (1 << 61)).item())
    parity = None
    if rank == 0 and not a.no_parity:
        try:
            parity = parity_subsample(torch, a, out[0], out[1])
        except Exception as e:  # noqa: BLE001 -- reported, never hidden
            parity = {"queries": 0, "ok": False, "error": repr(e)}

    for _ in range(max(1, min(a.warmup, 2))):
        step_e2e()
    # (no nvidia-smi sampler here: the host is on the critical path of an e2e step, and a second process attaching
    # to the driver stalls CUDA API calls by ~100 ms; throttle detection is the device arm's `clocks`)
    e2e_gemm_ms = []
    e2e_ms, _, _ = timed(step_e2e, a.steps, lambda: e2e_gemm_ms.append(index.last_timing()["gemm_topk_ms"]))

    pairs = float(a.nq) * float(a.n)
    ms_per_step = total_ms / a.steps
    value = pairs / (ms_per_step * 1e-3)
    e2e_value = pairs / (e2e_ms / a.steps * 1e-3)

    # roofline of the dominant kernel (gemm_topk_kernel): algorithmic flops 2*Q*N_local*D per launch
    peaks, which = None, "fallback"
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except (OSError, ValueError):
        pass
    g_ms = sum(gemm_ms) / len(gemm_ms)
    flops = 2.0 * a.nq * (r1 - r0) * a.d
    achieved = flops / (g_ms * 1e-3) / 1e12
    if peaks and g_ms > 50.0:
        peak, which = peaks["bf16_tflops_sustained"], "MEASURED_PEAKS.json bf16_tflops_sustained (kernel runs >50 ms per launch)"
    elif peaks:
        peak, which = peaks["bf16_tflops"], "MEASURED_PEAKS.json bf16_tflops (burst)"
    else:
        peak, which = 1590.0, "fallback 1.59 PFLOP/s (B200_PROFILING.md)"
    traffic = None   # ncu dram bytes of ONE launch of this exact workload; measured at N = 1 only (a shard's launch differs)
    if G == 1 and (a.nq, a.n, a.d, a.k) == (100_000, 10_000_000, 512, 100):
        try:
            with open(os.path.join(ROOT, "profiles", "gemm_topk_traffic.json")) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        except (OSError, ValueError):
            pass

    line = {
        "metric": "query-gallery pairs/sec (512-d)", "value": value, "unit": "pairs/s", "n_gpus": G,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(a), "nq": a.nq, "n": a.n, "d": a.d, "k": a.k,
                   "parallelism": "gallery rows sharded %d-way, queries replicated, local top-%d (head of the shard's list), 1 all-gather of [Q,%d] packed (d2,row) keys + merge with completeness proof" % (G, ops.local_k(a.k, G), ops.local_k(a.k, G))
                   if G > 1 else "single GPU",
                   "l2_policy": "inputs larger than L2 (gallery shard %.1f GB per rank)" % ((r1 - r0) * a.d * 2 / 1e9),
                   "tie_break": "d2 ascending, lowest gallery row first", "result": "canonical fp32 d2, verified top-k"},
        "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": a.nq * a.d * 2,
                "d2h_bytes_per_step": a.nq * k * 12, "ms_per_step": e2e_ms / a.steps,
                "gemm_topk_ms": sum(e2e_gemm_ms) / len(e2e_gemm_ms)},
        "gpu_launches": (stats["launches"] + (2 if G > 1 else 0)) * a.steps,   # + topk_pack and topk_merge_keys at N > 1
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                     "traffic": traffic, "kernel": "gemm_topk_pair_kernel<false> (cta_group::2)", "kernel_ms": g_ms,
                     "algorithmic_flops_per_launch": flops, "peak_source": which},
        "phases_ms": {kk: sum(p[kk] for p in phase_ms) / len(phase_ms) for kk in phase_ms[0]},
        "search_stats": stats, "result_checksum": checksum, "clocks": clocks, "parity_subsample": parity,
    }
    if xphase:
        line["phases_ms"].update(xphase)
    if add_info:
        line["index_add"] = add_info

    if rank == 0 and G == 1:
        del index
        torch.cuda.empty_cache()
        if not a.no_aux:
            try:
                line["aux"] = aux_measurements(torch, vq)
            except Exception as e:  # side numbers must not take the headline line down
                line["aux"] = {"error": repr(e)}
        if not a.no_cpu_baseline:
            v, cores, sample, _ = cpu_sample(a, 1, 1)
            line["cpu_baseline"] = {"value": v, "unit": "pairs/s", "cores": cores, "kind": "port", "sample": sample}
    if G > 1 and not a.no_aux:
        # the second half of BASELINE.json's metric at N GPUs: configs[0] mAP evaluation, sharded by query
        del index
        torch.cuda.empty_cache()
        try:
            r = sharded_map_eval(torch, dist, vq, ops, dev, G)
            if rank == 0:
                line["aux"] = {"map_eval_cfg0": r}
        except Exception as e:  # noqa: BLE001
            if rank == 0:
                line["aux"] = {"error": repr(e)}
    if G > 1:
        dist.barrier()
    if rank == 0:
        emit(line)
    if G > 1:
        dist.destroy_process_group()


def emit(obj):
    """The ONE JSON line goes to the process's original stdout; everything else that writes to fd 1 while the
    bench runs (NCCL prints its version banner there) has been pointed at stderr."""
    _REAL_STDOUT.write(json.dumps(obj) + "\n")
    _REAL_STDOUT.flush()


if __name__ == "__main__":
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
