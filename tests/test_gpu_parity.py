"""GPU parity tests: the CUDA path (through the C ABI, via the Python mirror) against the CPU
oracle on the same seeded inputs and against the golden fixtures minted from the reference.
Bar: indices / ranks / mined triplets bit-exact; canonical fp32 distances bit-exact; float64
metrics 1e-9 (north_star: 1e-6); loss / grad bit-exact vs oracle and 1e-6 vs the reference."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import oracle as orc  # noqa: E402


@pytest.fixture(scope="module")
def vq():
    import vqwild_b200
    return vqwild_b200


def _emb(rng, n, d, ncls=0, labels=None, scale=True):
    x = rng.standard_normal((n, d)).astype(np.float32)
    if ncls:
        c = rng.standard_normal((ncls, d)).astype(np.float32)
        x += 1.0 * c[labels]
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    if scale:
        x *= rng.uniform(0.6, 1.0, (n, 1)).astype(np.float32)
    return np.ascontiguousarray(x, np.float32)


def test_l2sq_bit_exact(vq):
    rng = np.random.default_rng(0)
    for nq, n, d in [(5, 77, 32), (70, 130, 512), (3, 9, 100)]:
        q, g = _emb(rng, nq, d), _emb(rng, n, d)
        got = vq.l2sq(q, g)
        assert np.array_equal(got.view(np.uint32), orc.l2sq(q, g).view(np.uint32))


@pytest.mark.parametrize("prec", ["EXACT", "SPLIT3", "BF16"])
def test_search_matches_oracle(vq, prec):
    rng = np.random.default_rng(1)
    nq, n, d, k = 200, 6000, 512, 100
    g = _emb(rng, n, d)
    g[123] = g[45]  # duplicate rows: exact score ties -> lowest index first
    q = _emb(rng, nq, d)
    q[0] = g[45]
    ix = vq.IndexFlatL2(d)
    ix.set_precision(getattr(vq, "PREC_" + prec))
    ix.add(g)
    D, I = ix.search(q, k)
    Do, Io = orc.search(q, g, k)
    st = ix.last_stats()
    print(prec, st)
    assert np.array_equal(I, Io)
    assert np.array_equal(D.view(np.uint32), Do.view(np.uint32))
    assert I[0, 0] == 45 and I[0, 1] == 123 and D[0, 0] == 0.0


@pytest.mark.parametrize("nq,n,d,k,bf16", [(130, 300_000, 64, 10, True), (130, 200_000, 64, 100, False),
                                            (1000, 60_000, 128, 50, True), (20, 2_100, 64, 5, True)])
def test_search_streamk_segments(vq, nq, n, d, k, bf16):
    """Stream-K decomposition: few query groups against a long gallery are cut into many segments
    per group (one candidate list each, merged by the select kernel); duplicates straddle cuts."""
    import torch
    rng = np.random.default_rng(11)
    g = _emb(rng, n, d)
    q = _emb(rng, nq, d)
    if bf16:
        g = torch.from_numpy(g).to(torch.bfloat16).float().numpy()
        q = torch.from_numpy(q).to(torch.bfloat16).float().numpy()
    for a, b in [(5, n - 7), (n // 2 - 1, n // 2 + 300), (1000, 1999)]:
        g[b] = g[a]
    q[1] = g[5]
    q[2] = g[n // 2 - 1]
    ix = vq.IndexFlatL2(d)
    if bf16:
        ix.add(torch.from_numpy(g).to(torch.bfloat16).cuda())
        D, I = ix.search(torch.from_numpy(q).to(torch.bfloat16).cuda(), k)
        D, I = D.cpu().numpy(), I.cpu().numpy()
    else:
        ix.add(g)
        D, I = ix.search(q, k)
    st = ix.last_stats()
    print(st)
    Do, Io = orc.search(q, g, k)
    assert np.array_equal(I, Io)
    assert np.array_equal(D.view(np.uint32), Do.view(np.uint32))
    assert I[1, 0] == 5 and I[1, 1] == n - 7
    if n >= 100_000:
        assert st["segments"] > st["work_units"] // 2 or st["segments"] >= 32  # the groups really were cut
        assert st["fallback_queries"] == 0


def test_search_scores_within_fp64_tolerance(vq):
    rng = np.random.default_rng(11)
    q, g = _emb(rng, 64, 512), _emb(rng, 3000, 512)
    ix = vq.IndexFlatL2(512)
    ix.add(g)
    D, I = ix.search(q, 50)
    ref = ((q[:, None, :].astype(np.float64) - g[I].astype(np.float64)) ** 2).sum(-1)
    assert np.max(np.abs(D - ref) / np.maximum(ref, 1e-30)) < 1e-5  # north_star fp32 tolerance


def test_search_small_and_ragged_shapes(vq):
    rng = np.random.default_rng(2)
    for nq, n, d, k in [(1, 1, 8, 1), (3, 50, 20, 50), (130, 257, 64, 7), (129, 1025, 100, 33)]:
        q, g = _emb(rng, nq, d), _emb(rng, n, d)
        ix = vq.IndexFlatL2(d)
        ix.add(g[: n // 2])
        ix.add(g[n // 2:])  # two adds
        assert ix.ntotal == n
        D, I = ix.search(q, k)
        Do, Io = orc.search(q, g, k)
        assert np.array_equal(I, Io), (nq, n, d, k)
        assert np.array_equal(D.view(np.uint32), Do.view(np.uint32))


def test_search_with_ignore_and_device_tensors(vq):
    import torch
    rng = np.random.default_rng(3)
    nq, n, d, k = 150, 4000, 512, 100
    g, q = _emb(rng, n, d), _emb(rng, nq, d)
    video = (np.arange(n) // 3).astype(np.int32)
    label = rng.integers(0, 20, n).astype(np.int32)
    ign = rng.integers(0, n // 3, (nq, 5)).astype(np.int32)
    ix = vq.IndexFlatL2(d)
    ix.add(torch.from_numpy(g).cuda())
    ix.set_meta(label, video)
    D, I = ix.search(torch.from_numpy(q).cuda(), k, ignore_video=torch.from_numpy(ign).cuda())
    D, I = D.cpu().numpy(), I.cpu().numpy()
    Do, Io = orc.search(q, g, n)
    for a in range(nq):
        keep = ~np.isin(video[Io[a]], ign[a])
        assert np.array_equal(I[a], Io[a][keep][:k])
        assert np.array_equal(D[a].view(np.uint32), Do[a][keep][:k].view(np.uint32))


def test_workspace_query_and_reserve(vq):
    """SURVEY.md 8b convention: the workspace a search needs can be asked for and reserved up front."""
    import torch
    rng = np.random.default_rng(40)
    q, g = _emb(rng, 300, 128), _emb(rng, 5000, 128)
    ix = vq.IndexFlatL2(128)
    ix.add(g)
    pers, trans = ix.workspace_bytes(300, 100)
    assert pers > 0 and trans > 0
    ix.reserve(300, 100)
    torch.cuda.synchronize()
    before = torch.cuda.mem_get_info()[0]
    D, I = ix.search(q, 100)
    torch.cuda.synchronize()
    after = torch.cuda.mem_get_info()[0]
    assert before - after <= trans + (32 << 20)   # nothing persistent was added by the search (pool granularity aside)
    Do, Io = orc.search(q, g, 100)
    assert np.array_equal(I, Io) and np.array_equal(D.view(np.uint32), Do.view(np.uint32))
    assert ix.workspace_bytes(4, 5000)[0] == 0      # k > 512: the full-sort path only uses stream-ordered temporaries


def test_search_forced_fallback_path(vq):
    rng = np.random.default_rng(4)
    q, g = _emb(rng, 40, 128), _emb(rng, 3000, 128)
    from vqwild_b200 import _lib as L
    L.debug_set("FORCE_FALLBACK", 1)
    try:
        ix = vq.IndexFlatL2(128)
        ix.add(g)
        D, I = ix.search(q, 100)
        st = ix.last_stats()
    finally:
        L.debug_set("FORCE_FALLBACK", 0)
    assert st["fallback_queries"] == 40
    Do, Io = orc.search(q, g, 100)
    assert np.array_equal(I, Io) and np.array_equal(D.view(np.uint32), Do.view(np.uint32))


def test_search_bf16_input_mode(vq):
    import torch
    rng = np.random.default_rng(5)
    q, g = _emb(rng, 100, 512, scale=False), _emb(rng, 5000, 512, scale=False)
    qb, gb = torch.from_numpy(q).cuda().bfloat16(), torch.from_numpy(g).cuda().bfloat16()
    ix = vq.IndexFlatL2(512)
    ix.add(gb)
    D, I = ix.search(qb, 100)
    qr, gr = qb.float().cpu().numpy(), gb.float().cpu().numpy()  # the bf16 values, widened
    Do, Io = orc.search(qr, gr, 100)
    assert np.array_equal(I.cpu().numpy(), Io)
    assert np.array_equal(D.cpu().numpy().view(np.uint32), Do.view(np.uint32))
    # bf16-input scores vs the fp32 originals: 2e-3 relative (north_star, stated separately)
    ref = ((q[:, None, :].astype(np.float64) - g[Io].astype(np.float64)) ** 2).sum(-1)
    assert np.max(np.abs(Do - ref) / ref) < 2e-3  # the stated bar; measured ~2e-4 (bf16 rounding of 512 unit-scale components)


@pytest.mark.parametrize("ragged", [False, True])
def test_segmax_topk(vq, ragged):
    rng = np.random.default_rng(6)
    nseg, d, nq, k = 300, 256, 140, 50
    lens = rng.integers(1, 90, nseg) if ragged else np.full(nseg, 64)
    seg_start = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    n = int(seg_start[-1])
    g, q = _emb(rng, n, d), _emb(rng, nq, d)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    ix.set_segments(seg_start)
    V, R, D = ix.search_segmax(q, k)
    Vo, Ro, Do = orc.segmax_topk(q, g, seg_start, k)
    print(ix.last_stats())
    assert np.array_equal(V, Vo) and np.array_equal(R, Ro)
    assert np.array_equal(D.view(np.uint32), Do.view(np.uint32))


@pytest.mark.parametrize("bf16", [False, True])
def test_segmax_best_moment_ties_and_near_ties(vq, bf16):
    """The GEMM marks a (query, video) candidate when its coarse arg-max is not provably the exact best moment; unmarked
    pairs are settled by ONE canonical score.  Adversarial videos: the best moment duplicated (exact tie: the lowest row
    must be reported), duplicated up to one ulp in one coordinate (inside the error band, either order), duplicated
    across the 32-column steps of the epilogue, and more videos than k' so that thresholds skip most steps.  Must equal
    the oracle and the result with the marking switched off, bit for bit."""
    import torch
    from vqwild_b200 import _lib as L
    rng = np.random.default_rng(31)
    nseg, d, nq, k = 2500, 128, 96, 20
    lens = rng.integers(40, 90, nseg)
    seg_start = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    n = int(seg_start[-1])
    g, q = _emb(rng, n, d), _emb(rng, nq, d)
    # every query gets a few videos that hold near copies of it -> those videos are its top candidates
    for qi in range(nq):
        for t, v in enumerate(rng.choice(nseg, 6, replace=False)):
            s0, S = int(seg_start[v]), int(lens[v])
            a, b = s0 + int(rng.integers(0, S)), s0 + int(rng.integers(0, S))
            base = q[qi] + 0.05 * rng.standard_normal(d).astype(np.float32)
            g[a] = base
            g[b] = base                                      # t % 3 == 0: exact duplicate (a != b mostly, any distance apart)
            if t % 3 == 1:
                g[b, 7] = np.nextafter(g[b, 7], np.float32(10))   # one ulp: canonical may or may not differ, coarse cannot tell
            if t % 3 == 2:
                g[b] = base + np.float32(3e-4) * rng.standard_normal(d).astype(np.float32)  # inside / around the band
    if bf16:
        g_t = torch.from_numpy(g).cuda().to(torch.bfloat16)
        q_t = torch.from_numpy(q).cuda().to(torch.bfloat16)
        g, q = g_t.float().cpu().numpy(), q_t.float().cpu().numpy()
    else:
        g_t, q_t = torch.from_numpy(g).cuda(), torch.from_numpy(q).cuda()
    ix = vq.IndexFlatL2(d)
    ix.add(g_t)
    ix.set_segments(seg_start)
    Vo, Ro, Do = orc.segmax_topk(q, g, seg_start, k)
    outs = []
    for off in (0, 1):
        L.debug_set("NO_SEG_UNIQUE", off)
        try:
            V, R, D = ix.search_segmax(q_t, k)
        finally:
            L.debug_set("NO_SEG_UNIQUE", 0)
        V, R, D = V.cpu().numpy(), R.cpu().numpy(), D.cpu().numpy()
        assert np.array_equal(V, Vo) and np.array_equal(R, Ro), "marking %s" % ("off" if off else "on")
        assert np.array_equal(D.view(np.uint32), Do.view(np.uint32))
        outs.append(ix.last_stats()["fallback_queries"])
    assert outs[0] == outs[1]


def _seg_of(video):
    chg = np.flatnonzero(np.diff(video)) + 1
    return np.concatenate([[0], chg, [len(video)]]).astype(np.int64)


def test_eval_moment_golden(vq, golden):
    """ARV_Retrieval_Moment.ranking (forked workers, do_nms, grouped R@N) as run by the reference."""
    rows = golden["mo_q_list_rows"]
    q = np.ascontiguousarray(golden["mo_qfeat"][rows[:, 0]])
    ql = golden["mo_qlabel"][rows[:, 0]]
    ign = golden["mo_qvideo"][rows]
    ix = vq.IndexFlatL2(q.shape[1])
    ix.add(golden["mo_feat"])
    ix.set_meta(golden["mo_hit"], golden["mo_video"])
    ix.set_segments(_seg_of(golden["mo_video"]))
    ix.set_moments(golden["mo_start"], golden["mo_end"], golden["mo_iou"], golden["mo_hit"])
    ap, rec, nk, km = ix.eval_moment(q, ql, ign, want_kept=True)
    np.testing.assert_allclose(ap, golden["mo_ap"], rtol=0, atol=1e-9)
    apo, reco = orc.eval_moment(q, golden["mo_feat"], golden["mo_video"], golden["mo_start"], golden["mo_end"],
                                golden["mo_iou"], golden["mo_hit"], ql, ign)
    np.testing.assert_allclose(ap, apo, rtol=0, atol=1e-12)
    np.testing.assert_allclose(rec, reco, rtol=0, atol=1e-12)
    # aggregated exactly as the reference's get_result
    names = [str(x) for x in golden["label_names"]]
    cls = [names[i] for i in ql]
    typ = ["base" if i < 6 else "novel" for i in ql]  # make_golden: 6 train + 5 test labels
    ev = vq.evaluation_metric(None, [[{"label": c, "retrieval_type": t}] for c, t in zip(cls, typ)])
    ev.update_arrays(cls, typ, ap, rec)
    res = ev.get_result()
    np.testing.assert_allclose([res["ap"], res["base_map"], res["novel_map"]], golden["mo_result"], rtol=0, atol=1e-9)
    got = np.array([[res[k][str(n)] for n in (30, 50, 100)] for k in ("recall", "base_recall", "novel_recall")])
    np.testing.assert_allclose(got, golden["mo_recall"], rtol=0, atol=1e-9)
    for i in (0, 7, 33):
        _, _, kept = orc.eval_moment_kept(q[i], golden["mo_feat"], golden["mo_video"], golden["mo_start"],
                                          golden["mo_end"], golden["mo_iou"], golden["mo_hit"], ql[i], ign[i])
        assert nk[i] == len(kept) and np.array_equal(np.flatnonzero(km[i]), np.sort(kept))


@pytest.mark.parametrize("thresh", [0.1, 0.2, 0.3, 1.0 / 3.0, 0.6, 0.7])
def test_eval_moment_nms_threshold_ties(vq, thresh):
    """Small integer windows: IoU = inter/union hits the threshold's decimal value exactly in real arithmetic all the
    time (1/10, 3/10, 7/10 ... are not doubles), so the kernel's division-free test and its quotient sliver must land
    on the side nms_cpu's float64 `ovr < thresh` does (utils_models.py:165-172)."""
    rng = np.random.default_rng(int(thresh * 1000))
    d, nq = 32, 12
    lens = np.concatenate([np.full(20, 64), rng.integers(1, 64, 20), [65, 130, 256, 300]])
    seg_start = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    n, nv = int(seg_start[-1]), len(lens)
    video = np.repeat(np.arange(nv), lens).astype(np.int32)
    start = rng.integers(0, 20, n).astype(np.float32)
    end = start + rng.integers(0, 10, n).astype(np.float32)      # inclusive lengths 1..10
    hit = rng.integers(-1, 6, n).astype(np.int32)
    iou = np.where(hit >= 0, rng.uniform(0, 1, n), 0).astype(np.float32)
    g, q = _emb(rng, n, d), _emb(rng, nq, d)
    ql = rng.integers(0, 6, nq).astype(np.int32)
    ign = rng.integers(0, nv, (nq, 2)).astype(np.int32)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    ix.set_meta(hit, video)
    ix.set_segments(seg_start)
    ix.set_moments(start, end, iou, hit)
    ap, rec, nk, km = ix.eval_moment(q, ql, ign, nms_thresh=thresh, want_kept=True)
    apo, reco = orc.eval_moment(q, g, video, start, end, iou, hit, ql, ign, nms_thresh=thresh)
    np.testing.assert_allclose(ap, apo, rtol=0, atol=1e-12)
    np.testing.assert_allclose(rec, reco, rtol=0, atol=1e-12)
    for i in range(nq):
        _, _, kept = orc.eval_moment_kept(q[i], g, video, start, end, iou, hit, ql[i], ign[i], nms_thresh=thresh)
        assert nk[i] == len(kept) and np.array_equal(np.flatnonzero(km[i]), np.sort(kept))


@pytest.mark.parametrize("case", ["uniform64", "ragged", "large_videos", "ties"])
def test_eval_moment_random(vq, case):
    rng = np.random.default_rng({"uniform64": 1, "ragged": 2, "large_videos": 3, "ties": 4}[case])
    d, nq, ncls = 64, 40, 12
    if case == "uniform64":
        lens = np.full(150, 64)
    elif case == "large_videos":
        lens = np.array([700, 3, 300, 257, 256, 1, 1200, 40])
    else:
        lens = rng.integers(1, 120, 90)
    seg_start = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    n, nv = int(seg_start[-1]), len(lens)
    video = np.repeat(np.arange(nv), lens).astype(np.int32) + 100
    # sliding windows: start = 5 * s, length in {5, 10, ..., 130}
    start = (5 * rng.integers(0, 60, n)).astype(np.float32)
    end = start + (5 * rng.integers(1, 27, n)).astype(np.float32)
    hit = rng.integers(-1, ncls, n).astype(np.int32)
    iou = np.where(hit >= 0, rng.uniform(0, 1, n), 0).astype(np.float32)
    iou[rng.random(n) < 0.05] = 0.5  # the >= boundary
    g = _emb(rng, n, d)
    if case == "ties":  # exact duplicate rows inside and across videos, duplicate windows
        g[5:40] = g[0]
        g[seg_start[3]:seg_start[3] + 5] = g[0]
        start[5:20], end[5:20] = start[5], end[5]
    q = _emb(rng, nq, d)
    q[0] = g[seg_start[2]]
    ql = rng.integers(0, ncls, nq).astype(np.int32)
    ign = (rng.integers(0, nv, (nq, 5)) + 100).astype(np.int32)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    ix.set_meta(hit, video)
    ix.set_segments(seg_start)
    ix.set_moments(start, end, iou, hit)
    ap, rec, nk, km = ix.eval_moment(q, ql, ign, want_kept=True)
    apo, reco = orc.eval_moment(q, g, video, start, end, iou, hit, ql, ign)
    np.testing.assert_allclose(ap, apo, rtol=0, atol=1e-12)
    np.testing.assert_allclose(rec, reco, rtol=0, atol=1e-12)
    for i in range(0, nq, 9):
        _, _, kept = orc.eval_moment_kept(q[i], g, video, start, end, iou, hit, ql[i], ign[i])
        assert nk[i] == len(kept) and np.array_equal(np.flatnonzero(km[i]), np.sort(kept))
    # no ignore list, device-resident inputs, other NMS threshold
    import torch
    ap2, rec2, _, _ = ix.eval_moment(torch.from_numpy(q).cuda(), torch.from_numpy(ql).cuda(), None, nms_thresh=0.3)
    apo2, reco2 = orc.eval_moment(q, g, video, start, end, iou, hit, ql, np.full((nq, 1), -7, np.int32), nms_thresh=0.3)
    np.testing.assert_allclose(ap2.cpu().numpy(), apo2, rtol=0, atol=1e-12)
    np.testing.assert_allclose(rec2.cpu().numpy(), reco2, rtol=0, atol=1e-12)


@pytest.mark.parametrize("C,device_input", [(512, False), (64, True), (13, False)])
def test_window_means_match_numpy_mean(vq, C, device_input):
    """f1: every moment feature equals the reference's np.mean(feat[:, lo:hi], axis=1) (:1140-1143) to 1e-6."""
    import torch
    from vqwild_b200.retrieval import garner_moments
    rng = np.random.RandomState(11)
    vids = []
    for v, T in enumerate([190, 33, 16, 2400 if C != 512 else 700, 5000 if C == 13 else 60, 455]):
        f = rng.rand(C, T).astype(np.float32)
        f /= np.linalg.norm(f, axis=0, keepdims=True)          # feat_func: unit norm per time step (main.py:232)
        vids.append(dict(feat=f, video_id="v%d" % v, activitynet_duration=T / 3.0 + 0.4,
                         annotations=[dict(segment=[1.0, T / 6.0], label="a")]))
    gv = [dict(g, feat=torch.from_numpy(g["feat"]).cuda()) for g in vids] if device_input else vids
    out = garner_moments(gv, ["a"])
    feat = out["feat"].cpu().numpy() if device_input else out["feat"]
    w = 0
    for v, g in enumerate(vids):
        T = g["feat"].shape[1]
        for cpm in range(1, 27):
            for st in range(0, int(g["activitynet_duration"]) - 5 * cpm, 5):
                a, z = st * 3, (st + 5 * cpm) * 3
                if z > T:
                    continue
                ref = np.mean(g["feat"][:, a:z], axis=1)
                assert out["start_sec"][w] == st and out["end_sec"][w] == st + 5 * cpm and out["video_id"][w] == g["video_id"]
                np.testing.assert_allclose(feat[w], ref, rtol=0, atol=1e-6)
                w += 1
        assert w == out["seg_start"][v + 1]
    assert w == feat.shape[0]


def _golden_trimmed(golden, query_num=1):
    feat, rows = golden["tr_feat"], golden["tr_q_list_rows"]
    acc = np.zeros((rows.shape[0], feat.shape[1]), np.float32)
    for i in range(query_num):
        acc = acc + feat[rows[:, i]]
    q = (acc / np.float32(query_num)).astype(np.float32) if query_num > 1 else acc
    return q, feat, golden["tr_label"], golden["tr_video"], golden["tr_label"][rows[:, 0]], golden["tr_video"][rows]


def test_eval_trimmed_golden(vq, golden):
    for qn, key in [(1, "tr_ap"), (5, "tr5_ap")]:
        q, g, gl, gv, ql, ign = _golden_trimmed(golden, qn)
        ix = vq.IndexFlatL2(g.shape[1])
        ix.add(g)
        ix.set_meta(gl, gv)
        ap, rec, top = ix.eval_trimmed(q, ql, ign, want_top=True)
        np.testing.assert_allclose(ap, golden[key], rtol=0, atol=1e-9)
        apo, reco, topo, _ = orc.eval_trimmed(q, g, gl, gv, ql, ign)
        np.testing.assert_allclose(rec, reco, rtol=0, atol=1e-12)
        assert np.array_equal(top, topo)
        if qn == 1:
            assert (gl[top] == golden["tr_cm_labels_top100"]).all()


def test_arv_retrieval_mirror_golden(vq, golden):
    """The thin caller with the reference's list-of-dict interface reproduces the reference's dict."""
    names = [str(x) for x in golden["label_names"]]
    feat, lab, vid = golden["tr_feat"], golden["tr_label"], golden["tr_video"]
    qrows = set(golden["tr_q_row"].tolist())
    items = []
    for i in range(feat.shape[0]):
        rt = "base" if lab[i] < 6 else ("novel" if lab[i] < 11 else "noise")
        items.append(dict(label=names[lab[i]], video_id="v_%04d" % vid[i], is_query=int(i in qrows),
                          retrieval_type=rt, feat=feat[i], frame_info=i))

    class A:
        query_num = 1
        eval_split = "testing"
        metric_feat_dim = feat.shape[1]

    ev = vq.ARV_Retrieval(A(), query_list=[it for it in items if lab[it["frame_info"]] < 11], gallery_list=items)
    res = ev.ranking()
    # generate_multi_query must reproduce the reference's random.seed(620) draws
    assert [[x["frame_info"] for x in ql] for ql in ev.query_list] == golden["tr_q_list_rows"].tolist()
    np.testing.assert_allclose([res["ap"], res["base_map"], res["novel_map"]], golden["tr_result"], rtol=0, atol=1e-9)
    got = np.array([[res[k][str(n)] for n in (30, 50, 100)] for k in ("recall", "base_recall", "novel_recall")])
    np.testing.assert_allclose(got, golden["tr_recall"], rtol=0, atol=1e-9)


def test_arv_retrieval_moment_mirror_golden(vq, golden):
    """List-of-dict surface of ARV_Retrieval_Moment.ranking -> the reference's map05 dict."""
    names = [str(x) for x in golden["label_names"]]
    mom = []
    for i in range(golden["mo_feat"].shape[0]):
        h = int(golden["mo_hit"][i])
        mom.append(dict(video_id="u_%03d" % golden["mo_video"][i], start_sec=float(golden["mo_start"][i]),
                        end_sec=float(golden["mo_end"][i]), feat=golden["mo_feat"][i],
                        closest_hit=None if h < 0 else dict(label=names[h], iou=float(golden["mo_iou"][i]))))
    mq = [dict(label=names[golden["mo_qlabel"][i]], video_id="u_%03d" % golden["mo_qvideo"][i],
               feat=golden["mo_qfeat"][i], qrow=i, retrieval_type="base" if golden["mo_qlabel"][i] < 6 else "novel")
          for i in range(golden["mo_qfeat"].shape[0])]

    class A:
        query_num = 1
        debug = False

    ev = vq.ARV_Retrieval_Moment(A(), query_list=mq, gallery_list=mom)
    res = ev.ranking()["map05"]
    assert [[x["qrow"] for x in ql] for ql in ev.query_list] == golden["mo_q_list_rows"].tolist()
    np.testing.assert_allclose([res["ap"], res["base_map"], res["novel_map"]], golden["mo_result"], rtol=0, atol=1e-9)
    got = np.array([[res[k][str(n)] for n in (30, 50, 100)] for k in ("recall", "base_recall", "novel_recall")])
    np.testing.assert_allclose(got, golden["mo_recall"], rtol=0, atol=1e-9)


def test_eval_trimmed_random_with_ties(vq):
    rng = np.random.default_rng(7)
    n, d, nq = 3000, 512, 300
    label = rng.integers(0, 25, n).astype(np.int32)
    g = _emb(rng, n, d, 25, label)
    g[1000:1040] = g[10]  # a block of exact duplicates (ties inside AP tie groups and at R@N cut-offs)
    video = (np.arange(n) // 2).astype(np.int32)
    qrow = rng.choice(n, nq, replace=False)
    qrow[:3] = [10, 1005, 1039]
    q = g[qrow]
    ign = np.stack([video[qrow]] + [video[rng.integers(0, n, nq)] for _ in range(4)], 1).astype(np.int32)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    ix.set_meta(label, video)
    ap, rec, top = ix.eval_trimmed(q, label[qrow], ign, want_top=True)
    apo, reco, topo, _ = orc.eval_trimmed(q, g, label, video, label[qrow], ign)
    np.testing.assert_allclose(ap, apo, rtol=0, atol=1e-12)
    np.testing.assert_allclose(rec, reco, rtol=0, atol=1e-12)
    assert np.array_equal(top, topo)


def test_eval_trimmed_many_exact_ties_near_the_top(vq):
    """Galleries with blocks of EXACT duplicate rows: tie groups of 40 and of 300 positives right at the top of the
    list (R@N must count positions inside the tie by gallery row; the 300-group overflows the kernel's tie queue)."""
    rng = np.random.default_rng(21)
    n, d, nq = 3000, 64, 12
    g = _emb(rng, n, d)
    label = rng.integers(0, 6, n).astype(np.int32)
    g[100:140] = g[100]
    label[100:140] = 2
    g[1000:1300] = g[1000]
    label[1000:1300] = 3
    label[1100:1150] = 4                       # non-positives inside the duplicate block
    video = np.arange(n, dtype=np.int32)
    q = _emb(rng, nq, d)
    q[0], q[1], q[2] = g[100], g[1000], g[1000] + 1e-3
    ql = np.array([2, 3, 3] + [int(x) for x in rng.integers(0, 6, nq - 3)], np.int32)
    ign = np.stack([rng.integers(0, n, nq), np.full(nq, 1005)], 1).astype(np.int32)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    ix.set_meta(label, video)
    ap, rec, top = ix.eval_trimmed(q, ql, ign, want_top=True)
    apo, reco, topo, _ = orc.eval_trimmed(q, g, label, video, ql, ign)
    np.testing.assert_allclose(ap, apo, rtol=0, atol=1e-12)
    np.testing.assert_allclose(rec, reco, rtol=0, atol=1e-12)
    assert np.array_equal(top, topo)


@pytest.mark.parametrize("bf16", [False, True])
def test_eval_trimmed_tensor_core_form_equals_canonical(vq, bf16):
    """The trimmed evaluation on tensor-core scores (gemm_topk STORE mode + ap_rows_tc) returns the numbers of the canonical
    all-pairs evaluation bit for bit: class-structured features (few survivors near a positive), duplicate rows that tie
    positives with non-positives, an ignore list, a gallery size that is not a multiple of 4."""
    import torch
    from vqwild_b200 import _lib as L
    rng = np.random.default_rng(31)
    n, d, nq, ncls = 5003, 128, 300, 40
    label = rng.integers(0, ncls, n).astype(np.int32)
    g = _emb(rng, n, d, ncls=ncls, labels=label)
    g[200:230] = g[200]                        # 30 identical rows ...
    label[200:215], label[215:230] = 3, 4      # ... half of them another class: positive / non-positive ties
    video = (np.arange(n) // 2).astype(np.int32)
    qrow = rng.choice(n, nq, replace=False)
    qrow[:3] = [200, 216, 7]
    q = np.ascontiguousarray(g[qrow])
    ql = label[qrow].copy()
    ign = np.stack([video[qrow]] + [rng.integers(0, n // 2, nq).astype(np.int32) for _ in range(4)], 1).astype(np.int32)
    if bf16:
        gt, qt = torch.from_numpy(g).cuda().bfloat16(), torch.from_numpy(q).cuda().bfloat16()
        g, q = gt.float().cpu().numpy(), qt.float().cpu().numpy()
    else:
        gt, qt = g, q
    ix = vq.IndexFlatL2(d)
    ix.add(gt)
    ix.set_meta(label, video)
    ap, rec, _ = ix.eval_trimmed(qt, ql, ign)
    handed_back = ix.last_stats()["segments"]
    L.debug_set("NO_TC_EVAL", 1)
    try:
        ap_c, rec_c, _ = ix.eval_trimmed(qt, ql, ign)
    finally:
        L.debug_set("NO_TC_EVAL", 0)
    to_np = lambda x: x.cpu().numpy() if hasattr(x, "cpu") else x  # noqa: E731
    ap, rec, ap_c, rec_c = map(to_np, (ap, rec, ap_c, rec_c))
    assert handed_back < nq // 4, handed_back            # the tensor-core form did the work
    assert np.array_equal(ap, ap_c) and np.array_equal(rec, rec_c)
    apo, reco, _, _ = orc.eval_trimmed(q, g, label, video, ql, ign)
    np.testing.assert_allclose(ap, apo, rtol=0, atol=1e-12)
    np.testing.assert_allclose(rec, reco, rtol=0, atol=1e-12)


def test_eval_trimmed_tensor_core_form_hands_back_and_switches_off(vq):
    """Labels unrelated to the features: the eps bands around ~n / classes positives spread through the list cover too much
    of the gallery; the queries are handed back to the canonical path (same numbers) and the index stops trying."""
    rng = np.random.default_rng(32)
    n, d, nq = 20000, 64, 64
    g, q = _emb(rng, n, d), _emb(rng, nq, d)
    label = rng.integers(0, 10, n).astype(np.int32)
    video = np.arange(n, dtype=np.int32)
    ql = rng.integers(0, 10, nq).astype(np.int32)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    ix.set_meta(label, video)
    ap, rec, _ = ix.eval_trimmed(q, ql, None)
    assert ix.last_stats()["segments"] == nq             # 2000 positives per query: every list overflows
    apo, reco, _, _ = orc.eval_trimmed(q, g, label, video, ql, np.full((nq, 1), -1, np.int32))
    np.testing.assert_allclose(ap, apo, rtol=0, atol=1e-12)
    np.testing.assert_allclose(rec, reco, rtol=0, atol=1e-12)
    ap2, _, _ = ix.eval_trimmed(q, ql, None)
    assert ix.last_stats()["segments"] == 0 and np.array_equal(ap, ap2)   # canonical path directly


def test_average_precision_lists_golden(vq, golden):
    yts, yss, want = [], [], []
    off = 0
    for c, n in enumerate(golden["hl_n"]):
        s = slice(off, off + n)
        off += n
        keep = golden["hl_ig"][s] == 0
        yt = golden["hl_tp"][s][keep].copy()
        yt[-1] = 1
        yts.append(yt), yss.append(golden["hl_sc"][s][keep]), want.append(golden["hl_ap"][c])
    got = vq.average_precision_batch(yts, yss)
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-12)


def test_evaluation_metric_mirror_golden(vq, golden):
    em = vq.evaluation_metric(None, [])
    off = 0
    for c, n in enumerate(golden["hl_n"]):
        s = slice(off, off + n)
        off += n
        hits = [dict(tp=int(t), ignore=bool(i), score=float(x))
                for t, i, x in zip(golden["hl_tp"][s], golden["hl_ig"][s], golden["hl_sc"][s])]
        r = em.multiprocess_calculate(hits)
        assert abs(r["ap"] - golden["hl_ap"][c]) < 1e-12
        np.testing.assert_allclose(r["recall_list"], golden["hl_rec"][c], rtol=0, atol=1e-12)


def test_nms_golden(vq, golden):
    assert vq.nms_cpu(golden["nms_kat_dets"], 0.5) == [1, 2]
    off, koff = 0, golden["nms_keep_off"]
    offs = np.concatenate([[0], np.cumsum(golden["nms_n"])])
    mask = vq.nms_segmented(golden["nms_dets"], offs, 0.5)
    for c, n in enumerate(golden["nms_n"]):
        want = np.zeros(n, np.uint8)
        want[golden["nms_keep"][koff[c]:koff[c + 1]]] = 1
        assert np.array_equal(mask[off:off + n], want)
        assert vq.nms_cpu(golden["nms_dets"][off:off + n], 0.5) == golden["nms_keep"][koff[c]:koff[c + 1]].tolist()
        off += n


def test_triplet_golden_and_oracle(vq, golden):
    import torch
    emb, labels, margin = golden["tl_emb"], golden["tl_labels"], float(golden["tl_margin"])
    x = torch.from_numpy(emb).cuda()
    d = vq.pdist(x).cpu().numpy()
    assert np.array_equal(d.view(np.uint32), orc.pdist(emb).view(np.uint32))
    np.testing.assert_allclose(d, golden["tl_pdist"], rtol=0, atol=2e-6)
    # reference-sampled (semihard) triplets through the "triplets given" kernel
    from vqwild_b200.loss import _GivenFn, TripletLoss
    xg = x.clone().requires_grad_(True)
    loss = _GivenFn.apply(xg, torch.from_numpy(golden["tl_semihard_trip"].astype(np.int32)), margin)
    loss.backward()
    assert abs(loss.item() - float(golden["tl_semihard_loss"])) < 1e-6
    np.testing.assert_allclose(xg.grad.cpu().numpy(), golden["tl_semihard_grad"], rtol=0, atol=1e-6)
    lo, go, _, _ = orc.triplet_fwd_bwd(emb, golden["tl_semihard_trip"], margin)
    assert loss.item() == lo and np.array_equal(xg.grad.cpu().numpy().view(np.uint32), go.view(np.uint32))
    # batch-hard, mined in the kernel
    crit = TripletLoss(margin=margin, sampling_method="batchhard")
    xb = x.clone().requires_grad_(True)
    loss = crit(xb, labels)
    loss.backward()
    assert abs(loss.item() - float(golden["tl_bh_loss"])) < 1e-6
    np.testing.assert_allclose(xb.grad.cpu().numpy(), golden["tl_bh_grad"], rtol=0, atol=1e-6)
    assert crit.sampler.give(x, labels) == golden["tl_bh_trip"].tolist()


def test_triplet_cfg2_shape_bit_exact_vs_oracle(vq):
    import torch
    from vqwild_b200.loss import _BatchHardFn
    rng = np.random.default_rng(8)
    B, d = 256, 512
    labels = np.repeat(np.arange(64), 4)
    emb = _emb(rng, B, d, 64, labels, scale=False)
    x = torch.from_numpy(emb).cuda().requires_grad_(True)
    loss, trip = _BatchHardFn.apply(x, torch.from_numpy(labels), 1.0)
    loss.backward()
    lo, go, to = orc.batchhard_triplet(emb, labels, 1.0)
    assert np.array_equal(trip.cpu().numpy(), to)
    assert loss.item() == lo
    assert np.array_equal(x.grad.cpu().numpy().view(np.uint32), go.view(np.uint32))
    # grouped variant: G independent batches in one call, incl. a singleton class (skipped anchor)
    G = 3
    embs = np.stack([_emb(rng, 64, 96, 16, np.repeat(np.arange(16), 4)) for _ in range(G)])
    labs = np.stack([np.repeat(np.arange(16), 4) for _ in range(G)])
    labs[1, 0] = 99
    xs = torch.from_numpy(embs).cuda().requires_grad_(True)
    loss, trip = _BatchHardFn.apply(xs, torch.from_numpy(labs), 0.3)
    loss.sum().backward()
    for gi in range(G):
        lo, go, to = orc.batchhard_triplet(embs[gi], labs[gi], 0.3)
        t = trip[gi].cpu().numpy()
        assert np.array_equal(t[t[:, 0] >= 0], to)
        assert loss[gi].item() == lo
        assert np.array_equal(xs.grad[gi].cpu().numpy().view(np.uint32), go.view(np.uint32))


def _check_groups(vq, embs, labs, margin):
    import torch
    from vqwild_b200.loss import _BatchHardFn
    xs = torch.from_numpy(embs).cuda().requires_grad_(True)
    loss, trip = _BatchHardFn.apply(xs, torch.from_numpy(labs), margin)
    loss.sum().backward()
    for gi in range(embs.shape[0]):
        lo, go, to = orc.batchhard_triplet(embs[gi], labs[gi], margin)
        t = trip[gi].cpu().numpy()
        assert np.array_equal(t[t[:, 0] >= 0], to), "group %d mined triplets" % gi
        assert loss[gi].item() == lo, "group %d loss" % gi
        assert np.array_equal(xs.grad[gi].cpu().numpy().view(np.uint32), go.view(np.uint32)), "group %d grad" % gi


@pytest.mark.parametrize("G,B,d,ncls", [(5, 256, 512, 64), (300, 64, 64, 8), (3, 100, 96, 10), (2, 256, 40, 32),
                                        (4, 37, 132, 5), (310, 48, 96, 6)])
def test_triplet_fused_shapes(vq, G, B, d, ncls):
    """Fused tensor-core kernel (B <= 256, d % 4 == 0): ragged B / d, more groups than SMs (a CTA's later groups run
    their pass 1 interleaved with the previous group's pass 3; even and odd k-block counts)."""
    rng = np.random.default_rng(100 + B + d)
    labs = np.stack([rng.permutation(np.arange(B) % ncls) for _ in range(G)]).astype(np.int64)
    embs = np.stack([_emb(rng, B, d, ncls, labs[g], scale=(g % 2 == 0)) for g in range(G)])
    _check_groups(vq, embs, labs, 0.5)


def test_triplet_fused_adversarial(vq):
    """Ties and degenerate groups: duplicate rows (exact distance ties -> lowest index), an anchor
    identical to all its positives (pdist's clamp region), a singleton class, a single-class group
    (no negatives: every anchor skipped), near-duplicates inside the TF32 error band (20 of them: over the
    capacity; 12 of them: within it), and a group whose ambiguity overflows the fused kernel's candidate capacity (handed to the CUDA-core path)."""
    rng = np.random.default_rng(5)
    B, d, ncls = 128, 64, 16
    labs = np.stack([np.arange(B) % ncls for _ in range(7)]).astype(np.int64)
    embs = np.stack([_emb(rng, B, d, ncls, labs[g]) for g in range(7)])
    embs[0, 16] = embs[0, 0]           # same class (0 and 16): d = clamp floor
    embs[0, 33] = embs[0, 1]           # 33 % 16 == 1: same class duplicate
    embs[0, 50] = embs[0, 3]           # 50 % 16 == 2: duplicate in ANOTHER class -> hardest negative ties
    embs[0, 66] = embs[0, 3]           # two identical negatives for the class-3 anchors: lowest index wins
    for j in range(0, B, ncls):        # group 1: class 0 all identical -> every positive in the clamp region
        embs[1, j] = embs[1, 0]
    labs[2, 5] = 77                    # singleton class
    labs[3, :] = 4                     # single class: no negatives at all
    base = embs[4, 7].copy()           # group 4: many rows within 1e-4 of each other (inside the coarse band)
    for j in range(20, 40):
        embs[4, j] = base + rng.standard_normal(d).astype(np.float32) * 1e-5
    embs[5, :] = embs[5, 0] + rng.standard_normal((B, d)).astype(np.float32) * 1e-6  # everything ambiguous
    base = embs[6, 5].copy()           # group 6: 10 such rows, all of different classes and spread over the warps, so
    for j in (5, 38, 71, 104, 9, 42, 75, 108, 13, 46):  # that each has 9 tied negatives (more than 8, at most 16: the
        embs[6, j] = base + rng.standard_normal(d).astype(np.float32) * 1e-5  # second candidate word) and stays fused
    _check_groups(vq, embs, labs, 1.0)
    from vqwild_b200 import _lib as L
    assert L.lib().vq_triplet_last_handed_back() >= 2       # groups 4 and 5 at least
    _check_groups(vq, embs[6:7], labs[6:7], 1.0)
    assert L.lib().vq_triplet_last_handed_back() == 0       # 9-10 negative candidates are within the capacity


def test_triplet_simt_path_matches(vq):
    """The CUDA-core path (B > 256, or d % 4 != 0) produces the same bits as the oracle too."""
    rng = np.random.default_rng(9)
    labs = np.stack([np.arange(300) % 30]).astype(np.int64)
    _check_groups(vq, np.stack([_emb(rng, 300, 48, 30, labs[0])]), labs, 0.7)
    labs = np.stack([np.arange(50) % 5 for _ in range(2)]).astype(np.int64)
    _check_groups(vq, np.stack([_emb(rng, 50, 33, 5, labs[g]) for g in range(2)]), labs, 0.7)


def test_pool_normalize(vq, golden):
    out = vq.pool_normalize(golden["pool_in"])
    np.testing.assert_allclose(out, golden["pool_out"], rtol=0, atol=1e-6)
    assert np.array_equal(out.view(np.uint32), orc.pool_normalize(golden["pool_in"]).view(np.uint32))


def test_topk_merge(vq):
    rng = np.random.default_rng(9)
    G, nq, k = 4, 33, 100
    d2 = np.sort(rng.random((G, nq, k)).astype(np.float32), axis=2)
    idx = rng.permutation(G * nq * k).reshape(G, nq, k).astype(np.int64)
    idx[2, :, 90:] = -1
    d2[1, 0, :5] = d2[0, 0, :5]  # cross-shard ties
    od, oi = vq.topk_merge(d2, idx)
    wd, wi = orc.topk_merge(d2, idx)
    assert np.array_equal(oi, wi) and np.array_equal(od.view(np.uint32), wd.view(np.uint32))


def test_topk_merge_unsorted_lists_take_the_sort_path(vq):
    """The merge ranks sorted lists; a caller's unsorted lists are detected and sorted instead."""
    rng = np.random.default_rng(13)
    G, nq, k = 3, 17, 60
    d2 = rng.random((G, nq, k)).astype(np.float32)            # NOT sorted
    idx = rng.permutation(G * nq * k).reshape(G, nq, k).astype(np.int64)
    idx[1, :, ::7] = -1                                        # holes anywhere
    od, oi = vq.topk_merge(d2, idx)
    wd, wi = orc.topk_merge(d2, idx)
    assert np.array_equal(oi, wi) and np.array_equal(od.view(np.uint32), wd.view(np.uint32))
    # few valid entries: the tail is padded
    idx[:] = -1
    idx[0, :, :5] = np.arange(5)
    d2[0, :, :5] = np.sort(d2[0, :, :5], axis=1)
    od, oi = vq.topk_merge(d2, idx)
    assert (oi[:, :5] == np.arange(5)).all() and (oi[:, 5:] == -1).all() and np.isinf(od[:, 5:]).all()


def test_topk_pack_merge_keys(vq):
    """The packed exchange (one 8-byte key per entry) gives what topk_merge gives on (d2, global row) lists."""
    rng = np.random.default_rng(12)
    G, nq, k = 8, 41, 100
    d2 = np.sort(rng.random((G, nq, k)).astype(np.float32), axis=2)
    local = np.stack([np.stack([np.sort(rng.choice(5000, k, replace=False)) for _ in range(nq)]) for _ in range(G)]).astype(np.int64)
    local[3, :, 80:] = -1
    d2[5, 2, :7] = d2[4, 2, :7]                      # cross-shard ties: the lower global row wins
    offs = np.arange(G) * 5000
    keys = np.stack([vq.topk_pack(d2[g], local[g], offs[g]) for g in range(G)])
    od, oi = vq.topk_merge_keys(keys)
    glob = np.where(local >= 0, local + offs[:, None, None], -1)
    wd, wi = orc.topk_merge(d2, glob)
    assert np.array_equal(oi, wi) and np.array_equal(od.view(np.uint32), wd.view(np.uint32))


def test_head_only_local_search_is_the_head_of_the_full_list_and_merges_to_the_single_index_answer(vq):
    """ops.sharded_search asks a shard only for local_k(k, G) neighbours: that list must be the first entries of the
    shard's top-k list bit for bit (k' 64 instead of 128 inside the GEMM), and the merged heads must be the single-index
    top-k with the completeness proof satisfied.  Four shards on one GPU, no collective."""
    import torch
    from vqwild_b200 import ops
    rng = np.random.default_rng(21)
    n, nq, d, k, G = 40000, 300, 64, 100, 4
    g = rng.standard_normal((n, d)).astype(np.float32)
    g[n - 3] = g[11]                                   # the same row on two shards: lower global row first
    q = rng.standard_normal((nq, d)).astype(np.float32)
    q[0] = g[11]
    m = ops.head_len(k, G)
    assert m == 56
    ops._head_failed.clear()
    assert ops.local_k(k, G) == m and ops.local_k(k, 2) == k
    full = vq.IndexFlatL2(d)
    full.add(g)
    Df, If = full.search(q, k)
    heads = []
    for r in range(G):
        r0, r1 = r * n // G, (r + 1) * n // G
        sh = vq.IndexFlatL2(d)
        sh.add(np.ascontiguousarray(g[r0:r1]))
        Dk, Ik = sh.search(q, k)
        Dm, Im = sh.search(q, m)
        assert sh.last_stats()["kprime"] < 128
        assert np.array_equal(Im, Ik[:, :m]) and np.array_equal(Dm.view(np.uint32), Dk[:, :m].view(np.uint32))
        heads.append(ops.topk_pack(Dm, Im, r0))
    od, oi, inc = ops.topk_merge_heads(np.stack(heads), k)
    assert int(inc[0]) == 0
    assert np.array_equal(oi, If) and np.array_equal(od.view(np.uint32), Df.view(np.uint32))
    ops._head_failed[(k, G)] = True                    # a refused head round is remembered: the next search asks for k
    assert ops.local_k(k, G) == k
    ops._head_failed.clear()


def test_topk_merge_heads_two_round_exchange(vq):
    """Round 1 of the exchange merges only the first m entries of every shard's list: final (and equal to the full merge)
    when every head reaches the merged k-th key, reported incomplete otherwise."""
    from vqwild_b200 import ops
    rng = np.random.default_rng(14)
    G, nq, k, m = 8, 50, 100, 40
    assert ops.head_len(100, 8) == 40 and ops.head_len(100, 4) == 56 and ops.head_len(100, 2) == 100
    d2 = np.sort(rng.random((G, nq, k)).astype(np.float32), axis=2)
    local = np.stack([np.stack([np.sort(rng.choice(5000, k, replace=False)) for _ in range(nq)]) for _ in range(G)]).astype(np.int64)
    local[3, :, 30:] = -1                            # a short list: its head ends inside the list
    offs = np.arange(G) * 5000
    keys = np.stack([vq.topk_pack(d2[g], local[g], offs[g]) for g in range(G)])
    wd, wi = orc.topk_merge(d2, np.where(local >= 0, local + offs[:, None, None], -1))
    od, oi, inc = ops.topk_merge_heads(np.ascontiguousarray(keys[:, :, :m]), k)
    assert int(inc[0]) == 0
    assert np.array_equal(oi, wi) and np.array_equal(od.view(np.uint32), wd.view(np.uint32))
    # one shard holds most of a query's neighbours: its head is too short -> that query is reported
    d2b = d2.copy()
    d2b[5, 7] = np.sort(rng.random(k).astype(np.float32) * 1e-3)
    d2b[2, 9] = np.sort(rng.random(k).astype(np.float32) * 1e-3)
    keysb = np.stack([vq.topk_pack(d2b[g], local[g], offs[g]) for g in range(G)])
    _, _, inc = ops.topk_merge_heads(np.ascontiguousarray(keysb[:, :, :m]), k)
    assert int(inc[0]) == 2
    # fewer than k entries in all
    few = np.full((2, 3, 8), -1, np.int64)
    few[0, :, :5] = np.arange(5)
    kf = np.stack([vq.topk_pack(np.sort(rng.random((3, 8)).astype(np.float32), axis=1), few[g], 100 * g) for g in range(2)])
    od, oi, inc = ops.topk_merge_heads(kf, 10)
    assert int(inc[0]) == 0 and (oi[:, 5:] == -1).all() and (np.sort(oi[:, :5], axis=1) == np.arange(5)).all()


def test_full_size_cfg1_properties(vq):
    """BASELINE configs[0] shape (4096 x 20480 x 512): oracle on a query subsample + properties."""
    rng = np.random.default_rng(10)
    nq, n, d, k = 4096, 20480, 512, 100
    label = rng.integers(0, 180, n).astype(np.int32)
    g = _emb(rng, n, d, 180, label)
    q = g[rng.choice(n, nq, replace=False)] + 0.05 * _emb(rng, nq, d)
    q = np.ascontiguousarray(q, np.float32)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    D, I = ix.search(q, k)
    print(ix.last_stats())
    assert (np.diff(D, axis=1) >= 0).all()                       # sorted
    assert (np.sort(I, axis=1)[:, 1:] != np.sort(I, axis=1)[:, :-1]).all()  # no duplicates
    sub = rng.choice(nq, 48, replace=False)
    Do, Io = orc.search(q[sub], g, k)
    assert np.array_equal(I[sub], Io) and np.array_equal(D[sub].view(np.uint32), Do.view(np.uint32))


def test_large_gallery_cfg3_properties(vq):
    """BASELINE configs[3] shape at a size the box checks in seconds (2 M clips x 512-d bf16, top-100): the properties
    that do not depend on the gallery size -- the tensor-core pre-filter path returns exactly what the canonical
    CUDA-core scan does (scores and rows, bit for bit), lists are sorted with distinct rows, and a row-sharded search
    merged with topk_merge equals the single-index answer (what the N-GPU path relies on)."""
    import torch
    gen = torch.Generator(device="cuda").manual_seed(11)
    n, nq, d, k = 2_000_000, 512, 512, 100
    g = torch.randn(n, d, device="cuda", generator=gen)
    g = (g / g.norm(dim=1, keepdim=True) * (0.7 + 0.3 * torch.rand(n, 1, device="cuda", generator=gen))).to(torch.bfloat16)
    q = torch.randn(nq, d, device="cuda", generator=gen)
    q[:64] = g[torch.randint(0, n, (64,), device="cuda", generator=gen)].float() + 0.02 * q[:64]   # queries with close hits
    q = (q / q.norm(dim=1, keepdim=True)).to(torch.bfloat16)
    q[64:72] = g[5:13]                                                                           # exact hits: d2 = 0
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    D, I = ix.search(q, k)
    st = ix.last_stats()
    assert st["fallback_queries"] == 0, st
    Dc, Ic = D.cpu().numpy(), I.cpu().numpy()
    assert (np.diff(Dc, axis=1) >= 0).all()
    assert (np.sort(Ic, axis=1)[:, 1:] != np.sort(Ic, axis=1)[:, :-1]).all()
    assert (Ic[64:72, 0] == np.arange(5, 13)).all() and (Dc[64:72, 0] == 0).all()
    ix.set_precision(3)                                   # VQ_PREC_EXACT
    De, Ie = ix.search(q, k)
    assert torch.equal(I, Ie) and torch.equal(D.view(torch.int32), De.view(torch.int32))
    # two row shards -> local top-k -> merge == whole gallery
    parts_d, parts_i = [], []
    for lo, hi in ((0, 900_000), (900_000, n)):
        sh = vq.IndexFlatL2(d)
        sh.add(g[lo:hi])
        Ds, Is = sh.search(q, k)
        parts_d.append(Ds)
        parts_i.append(torch.where(Is >= 0, Is + lo, Is))
    Dm, Im = vq.topk_merge(torch.stack(parts_d), torch.stack(parts_i))
    assert torch.equal(Im, I) and torch.equal(Dm.view(torch.int32), D.view(torch.int32))


def test_cfg4_elmo_1024d_k_sweep(vq):
    """BASELINE configs[4]: 1024-d unit-norm embeddings (ELMo word space), 4096 x 20480, top-100 exact + the
    k sweep {5, 30, 50, 100} (R@k as the reference's :475-481; AP@k as defined in metrics.ap_at_k)."""
    rng = np.random.default_rng(31)
    nq, n, d, k = 4096, 20480, 1024, 100
    label = rng.integers(0, 200, n).astype(np.int32)              # 40 + 20 + 140 classes
    g = _emb(rng, n, d, 200, label, scale=False)
    qrow = rng.choice(n, nq, replace=False)
    q = np.ascontiguousarray(g[qrow] + 0.05 * _emb(rng, nq, d, scale=False), np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    video = np.arange(n, dtype=np.int32)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    ix.set_meta(label, video)
    ign = video[qrow][:, None].astype(np.int32)                   # the query's own item leaves the list
    D, I = ix.search(q, k, ignore_video=ign)
    print(ix.last_stats())
    sub = rng.choice(nq, 40, replace=False)
    Do, Io = orc.search(q[sub], g, k + 1)
    for a, s in enumerate(sub):                                   # oracle list with the ignored row removed
        keep = Io[a] != qrow[s]
        assert np.array_equal(I[s], Io[a][keep][:k])
        assert np.array_equal(D[s].view(np.uint32), Do[a][keep][:k].view(np.uint32))
    ref64 = ((q[sub][:, None, :].astype(np.float64) - g[I[sub]].astype(np.float64)) ** 2).sum(-1)
    assert np.max(np.abs(D[sub] - ref64) / np.maximum(ref64, 1e-12)) < 1e-5   # fp32 scores: 1e-5 relative (north_star)
    tp = (label[I] == label[qrow][:, None]).astype(np.int32)
    n_pos = np.array([(label == label[r]).sum() - 1 for r in qrow])
    rk, apk = vq.recall_at_k(tp, n_pos), vq.ap_at_k(tp, n_pos)
    for kk in (5, 30, 50, 100):
        ref_r = np.array([tp[i, :kk].sum() / (n_pos[i] + 1e-10) for i in range(0, nq, 97)])
        np.testing.assert_allclose(rk[kk][::97], ref_r, rtol=0, atol=1e-12)
        ref_ap = []
        for i in range(0, nq, 97):
            hits = np.flatnonzero(tp[i, :kk])
            den = min(n_pos[i], kk)
            ref_ap.append(sum((j2 + 1) / (h + 1) for j2, h in enumerate(hits)) / den if den else 0.0)
        np.testing.assert_allclose(apk[kk][::97], ref_ap, rtol=0, atol=1e-12)
    assert rk[100].mean() >= rk[5].mean()
    # full-list AP + R@30/50/100 with the reference semantics on the same 1024-d index
    ap, rec, _ = ix.eval_trimmed(q[sub], label[qrow][sub], ign[sub])
    apo, reco, _, _ = orc.eval_trimmed(q[sub], g, label, video, label[qrow][sub], ign[sub])
    np.testing.assert_allclose(ap, apo, rtol=0, atol=1e-9)
    np.testing.assert_allclose(rec, reco, rtol=0, atol=1e-9)
    np.testing.assert_allclose(rec[:, 0], rk[30][sub], rtol=0, atol=1e-9)


def test_wordembed_topk_mirror(vq):
    """word_embed/wordembed_retrieval_test.py:9-23 restated with numpy (LA.norm on normalised vectors)."""
    rng = np.random.default_rng(32)
    video_embed = {"act_%d" % i: rng.standard_normal(1024).tolist() for i in range(30)}
    openset = {"k700_%d" % i: rng.standard_normal(1024).tolist() for i in range(700)}
    got = vq.wordembed_topk(video_embed, openset, k=5)
    for key, value in video_embed.items():
        v = np.array(value).reshape(1, -1)
        v = v / np.linalg.norm(v)
        res = []
        for ck, cv in openset.items():
            c = np.array(cv).reshape(1, -1)
            c = c / np.linalg.norm(c)
            res.append((ck, -np.linalg.norm(c - v)))
        res.sort(key=lambda x: x[1], reverse=True)
        np.testing.assert_allclose([x[1] for x in got[key]], [x[1] for x in res[:5]], rtol=0, atol=1e-5)
        # the index works on float32 copies of the float64-normalised vectors: labels agree except where two
        # float64 scores are closer than float32 resolution
        true = dict(res)
        for (lab, _), (_, ref_score) in zip(got[key], res[:5]):
            assert abs(true[lab] - ref_score) < 1e-5
