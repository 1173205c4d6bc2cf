"""GPU parity, round 2: (1) the reference's ACTUAL faiss call -- search(xq[1,d], k = ntotal), the full ascending list --
through the C ABI and through the `sys.modules["faiss"]` shim; (2) the real-metadata goldens minted by running the
reference's evaluators on data_generate/arv_db_100_20_80*.json (tests/golden/make_golden_real.py); (3) BASELINE.json
configs[2] at FULL size against the oracle on query subsamples.  Bars as in test_gpu_parity.py: indices and canonical
distances bit-exact, float64 metrics 1e-9."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import oracle as orc  # noqa: E402

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from synth import class_feats  # noqa: E402


@pytest.fixture(scope="module")
def vq():
    import vqwild_b200
    return vqwild_b200


@pytest.fixture(scope="module")
def real():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_golden_real.npz"),
                   allow_pickle=False)


def _emb(rng, n, d):
    x = rng.standard_normal((n, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    x *= rng.uniform(0.6, 1.0, (n, 1)).astype(np.float32)
    return np.ascontiguousarray(x, np.float32)


def _biteq(a, b):
    return np.array_equal(np.asarray(a).view(np.uint32), np.asarray(b).view(np.uint32))


# ---------------------------------------------------------------------------------------------------------------
# (1) k = ntotal
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [7670, 20480])
def test_full_sort_reference_call_pattern(vq, n):
    """dataloader_baseline.py:1592-1597: `index.search(q.reshape(1, 512), k=faiss_xb.shape[0])` in a loop over the
    queries, plus the same queries in one batched call.  D and I bit-equal to the oracle's full sort; a duplicate
    gallery row makes an exact tie (lowest row first); a query equal to a gallery row gives d2 = 0 first."""
    rng = np.random.default_rng(n)
    d = 512
    g = _emb(rng, n, d)
    g[4321] = g[77]
    g[n - 1] = g[77]
    q = _emb(rng, 24, d)
    q[0] = g[77]
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    assert ix.ntotal == n
    Do, Io = orc.search(q, g, n)
    for i in range(6):                                   # the reference's loop, one query per call
        D, I = ix.search(q[i].reshape(1, d), k=n)
        assert D.shape == (1, n) and I.shape == (1, n) and D.dtype == np.float32 and I.dtype == np.int64
        assert np.array_equal(np.squeeze(I), Io[i]) and _biteq(np.squeeze(D), Do[i])
    D, I = ix.search(q, n)                               # batched
    assert np.array_equal(I, Io) and _biteq(D, Do)
    assert I[0, :3].tolist() == [77, 4321, n - 1] and D[0, 0] == 0.0 and D[0, 2] == 0.0
    assert (np.diff(D, axis=1) >= 0).all() and (np.sort(I, axis=1) == np.arange(n)).all()
    # 512 < k < ntotal takes the same path and returns the head of the list
    D2, I2 = ix.search(q, 1000)
    assert np.array_equal(I2, Io[:, :1000]) and _biteq(D2, Do[:, :1000])
    # device tensors in, device tensors out
    import torch
    Dt, It = ix.search(torch.from_numpy(q).cuda(), n)
    assert np.array_equal(It.cpu().numpy(), Io) and _biteq(Dt.cpu().numpy(), Do)


def test_full_sort_ragged_sizes_and_ignore(vq):
    """Sizes around the sort's 4096-element blocks, odd dims, bf16 rows, and the ignore rule on the full-sort path
    (ignored rows leave the list; the tail is padded with +inf / -1)."""
    import torch
    rng = np.random.default_rng(3)
    for n, d, nq in [(513, 32, 5), (4096, 64, 3), (4097, 100, 4), (12289, 24, 7)]:
        g, q = _emb(rng, n, d), _emb(rng, nq, d)
        g[n // 2] = g[1]
        ix = vq.IndexFlatL2(d)
        ix.add(g)
        D, I = ix.search(q, n)
        Do, Io = orc.search(q, g, n)
        assert np.array_equal(I, Io) and _biteq(D, Do), (n, d)
    n, d, nq = 9000, 64, 9
    g, q = _emb(rng, n, d), _emb(rng, nq, d)
    video = (np.arange(n) // 3).astype(np.int32)
    ign = rng.integers(0, n // 3, (nq, 5)).astype(np.int32)
    ix = vq.IndexFlatL2(d)
    ix.add(g)
    ix.set_meta(np.zeros(n, np.int32), video)
    D, I = ix.search(q, n, ignore_video=ign)
    Do, Io = orc.search(q, g, n)
    for a in range(nq):
        keep = ~np.isin(video[Io[a]], ign[a])
        m = int(keep.sum())
        assert np.array_equal(I[a, :m], Io[a][keep]) and _biteq(D[a, :m], Do[a][keep])
        assert (I[a, m:] == -1).all() and np.isinf(D[a, m:]).all()
    gb, qb = torch.from_numpy(g).cuda().bfloat16(), torch.from_numpy(q).cuda().bfloat16()
    ixb = vq.IndexFlatL2(d)
    ixb.add(gb)
    Db, Ib = ixb.search(qb, n)
    Do, Io = orc.search(qb.float().cpu().numpy(), gb.float().cpu().numpy(), n)
    assert np.array_equal(Ib.cpu().numpy(), Io) and _biteq(Db.cpu().numpy(), Do)


def test_faiss_shim_zero_diff_route_golden(vq, golden):
    """INTEGRATION.md route 1: `sys.modules["faiss"] = faiss_module()` and NOTHING else replaced -- the per-query
    body of ARV_Retrieval.ranking (dataloader_baseline.py:1526-1610: labelling loop, search(k = ntotal), reorder loop,
    add2dict) written against the faiss surface, with the host-side evaluation_metric mirror.  Equals the reference's
    own result on the golden gallery."""
    faiss = vq.faiss_module()
    names = [str(x) for x in golden["label_names"]]
    feat, lab, vid = golden["tr_feat"], golden["tr_label"], golden["tr_video"]
    index = faiss.IndexFlatL2(feat.shape[1])
    index.add(feat)
    assert index.ntotal == feat.shape[0]
    rows = golden["tr_q_list_rows"]
    qlists = [[dict(label=names[lab[r]], video_id=int(vid[r]), feat=feat[r],
                    retrieval_type="base" if lab[r] < 6 else "novel") for r in rr] for rr in rows]
    em = vq.evaluation_metric(None, qlists)
    for queries in qlists:
        query = queries[0]
        ignore_videoid_list = [x["video_id"] for x in queries]
        single_query_hit = [dict(gt_label=query["label"], label=names[lab[j]], tp=int(names[lab[j]] == query["label"]),
                                 ignore=int(vid[j]) in ignore_videoid_list) for j in range(feat.shape[0])]
        D, I = index.search(query["feat"].reshape(1, -1), k=feat.shape[0])
        D, I = np.squeeze(D), np.squeeze(I)
        ranked = []
        for j in range(I.shape[0]):
            single_query_hit[I[j]]["score"] = -float(D[j])
            ranked.append(single_query_hit[I[j]])
        em.add2dict(cls_name=query["label"], retrieval_type=query["retrieval_type"], single_query_hit=ranked)
    np.testing.assert_allclose(em.class_agnostic_ap, golden["tr_ap"], rtol=0, atol=1e-9)
    res = em.get_result()
    np.testing.assert_allclose([res["ap"], res["base_map"], res["novel_map"]], golden["tr_result"], rtol=0, atol=1e-9)
    got = np.array([[res[k][str(n)] for n in (30, 50, 100)] for k in ("recall", "base_recall", "novel_recall")])
    np.testing.assert_allclose(got, golden["tr_recall"], rtol=0, atol=1e-9)
    assert np.array_equal(np.array([names.index(x) for x in res["cm_dict"]["label"]]).reshape(len(qlists), -1),
                          golden["tr_cm_labels_top100"])


# ---------------------------------------------------------------------------------------------------------------
# (2) real metadata
# ---------------------------------------------------------------------------------------------------------------
class _Args(dict):
    __getattr__ = dict.__getitem__


def _check_result(res, z, pre):
    np.testing.assert_allclose([res["ap"], res["base_map"], res["novel_map"]], z[pre + "_result"], rtol=0, atol=1e-9)
    got = np.array([[res[k][str(n)] for n in (30, 50, 100)] for k in ("recall", "base_recall", "novel_recall")])
    np.testing.assert_allclose(got, z[pre + "_recall"], rtol=0, atol=1e-9)


def test_real_trimmed_metadata_golden(vq, real):
    """ARV_Retrieval mirror on the reference's real test split (7 670 items, 6 338 queries, the real is_query /
    retrieval_type / shared-video_id distribution) == the reference's own ranking() on the same seeded features."""
    names = [str(x) for x in real["rt_label_names"]]
    lab, vid = real["rt_label"].astype(np.int64), real["rt_video"]
    feat = class_feats(11, lab, 512)
    rtn = ("base", "novel", "noise")
    items = [dict(label=names[lab[i]], video_id="v%d" % vid[i], is_query=int(real["rt_is_query"][i]),
                  retrieval_type=rtn[real["rt_rtype"][i]], feat=feat[i], frame_info=i) for i in range(len(lab))]
    args = _Args(query_num=1, eval_split="testing", metric_feat_dim=512)
    ev = vq.ARV_Retrieval(args, query_list=[it for i, it in enumerate(items) if real["rt_in_query_list"][i]],
                          gallery_list=items)
    res = ev.ranking()
    assert np.array_equal(np.array([[x["frame_info"] for x in ql] for ql in ev.query_list]), real["rt_q_list_rows"])
    np.testing.assert_allclose(ev.class_map_evaluation.class_agnostic_ap, real["rt_ap"], rtol=0, atol=1e-9)
    _check_result(res, real, "rt")
    # the zero-diff route on the same index: the reference's call for a handful of queries
    rows = real["rt_q_list_rows"][:4, 0]
    Do, Io = orc.search(feat[rows], feat, len(lab))
    for a, r in enumerate(rows):
        D, I = ev.faiss_index.search(feat[r].reshape(1, 512), k=len(lab))
        assert np.array_equal(np.squeeze(I), Io[a]) and _biteq(np.squeeze(D), Do[a])


def test_real_clip_mirror_golden(vq, real):
    """ARV_Retrieval_Clip mirror (dataloader_baseline.py:883-955) == the reference on clips its own garner_feat /
    cal_label cut from real untrimmed-video metadata."""
    lab, vid = real["cl_label"], real["cl_video"]
    cfeat, qfeat = class_feats(21, lab, 64), class_feats(22, real["cl_qlabel"], 64)
    clips = [dict(clip_label="unknown" if lab[i] == 0 else "c%d" % lab[i], video_id="v%d" % vid[i], feat=cfeat[i])
             for i in range(len(lab))]
    rtn = ("base", "novel")
    ql = real["cl_qlabel"]
    queries = [dict(label="c%d" % ql[i], video_id="v%d" % real["cl_qvideo"][i], feat=qfeat[i], qrow=i,
                    retrieval_type=rtn[real["cl_qrtype"][i]]) for i in range(len(ql))]
    ev = vq.ARV_Retrieval_Clip(_Args(query_num=1, debug=False), query_list=queries, gallery_list=clips)
    res = ev.ranking()
    assert np.array_equal(np.array([[x["qrow"] for x in q] for q in ev.query_list]), real["cl_q_list_rows"])
    np.testing.assert_allclose(ev.class_map_evaluation.class_agnostic_ap, real["cl_ap"], rtol=0, atol=1e-9)
    _check_result(res, real, "cl")


def _real_moments(real):
    hit = real["mr_hit"]
    mfeat, qfeat = class_feats(31, hit + 1, 64, spread=0.6), class_feats(32, real["mr_qlabel"] + 1, 64, spread=0.6)
    return mfeat, qfeat


def test_real_moment_metadata_golden(vq, real):
    """ARV_Retrieval_Moment mirror on moments the reference's own garner_feat cut from real metadata -- ragged videos,
    incl. the 3 523-moment one (the kernel's block-per-video class) -- == the reference's ranking() (forked worker, NMS,
    float64 IoU >= 0.5, grouped R@N)."""
    mfeat, qfeat = _real_moments(real)
    hit, iou = real["mr_hit"], real["mr_iou"]
    assert np.bincount(real["mr_video"]).max() == 3523
    mom = [dict(video_id="u%d" % real["mr_video"][i], start_sec=int(real["mr_start"][i]), end_sec=int(real["mr_end"][i]),
                feat=mfeat[i], closest_hit=None if hit[i] < 0 else dict(label="c%d" % hit[i], iou=float(iou[i])))
           for i in range(len(hit))]
    rtn = ("base", "novel")
    ql = real["mr_qlabel"]
    queries = [dict(label="c%d" % ql[i], video_id="u%d" % real["mr_qvideo"][i], feat=qfeat[i], qrow=i,
                    retrieval_type=rtn[real["mr_qrtype"][i]]) for i in range(len(ql))]
    ev = vq.ARV_Retrieval_Moment(_Args(query_num=1, debug=False), query_list=queries, gallery_list=mom)
    res = ev.ranking()["map05"]
    assert np.array_equal(np.array([[x["qrow"] for x in q] for q in ev.query_list]), real["mr_q_list_rows"])
    np.testing.assert_allclose(ev.class_map_evaluation05.class_agnostic_ap, real["mr_ap"], rtol=0, atol=1e-9)
    _check_result(res, real, "mr")


def test_moment_iou_float64_boundary(vq):
    """a float64 IoU one ulp below 0.5 rounds to 0.5f; the reference compares the float64 value
    (dataloader_baseline.py:1315-1319), so that moment is NOT a true positive."""
    rng = np.random.default_rng(8)
    nv, S, d = 6, 4, 32
    n = nv * S
    g = _emb(rng, n, d)
    below = np.nextafter(0.5, 0.0)
    assert np.float32(below) == np.float32(0.5)
    iou = np.full(n, below, np.float64)
    iou[::3] = 0.5
    mom = [dict(video_id="u%d" % (i // S), start_sec=5 * (i % S), end_sec=5 * (i % S) + 5, feat=g[i],
                closest_hit=dict(label="a", iou=float(iou[i]))) for i in range(n)]
    q = _emb(rng, 8, d)
    queries = [dict(label="a", video_id="x%d" % i, feat=q[i], retrieval_type="base") for i in range(8)]
    ev = vq.ARV_Retrieval_Moment(_Args(query_num=1, debug=True), query_list=queries, gallery_list=mom)
    ev.ranking()
    hit = np.zeros(n, np.int32)
    video = (np.arange(n) // S).astype(np.int32)
    st = (5.0 * (np.arange(n) % S)).astype(np.float32)
    flag = (iou >= 0.5).astype(np.float32)
    apo, _ = orc.eval_moment(q, g, video, st, st + 5, flag, hit, np.zeros(8, np.int32), np.full((8, 1), -7, np.int32))
    np.testing.assert_allclose(ev.class_map_evaluation05.class_agnostic_ap, apo, rtol=0, atol=1e-12)
    wrong, _ = orc.eval_moment(q, g, video, st, st + 5, iou.astype(np.float32), hit, np.zeros(8, np.int32),
                               np.full((8, 1), -7, np.int32))
    assert np.abs(wrong - apo).max() > 1e-3   # the float32 narrowing would have changed the answer


# ---------------------------------------------------------------------------------------------------------------
# (3) BASELINE.json configs[2] at full size: 4 096 queries x 20 480 videos x 64 moments x 512-d
# ---------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def cfg2():
    import torch
    V, S, d, nq = 20480, 64, 512, 4096
    gen = torch.Generator(device="cuda").manual_seed(2)
    g = torch.randn((V * S, d), generator=gen, device="cuda")
    g /= g.norm(dim=1, keepdim=True)
    g *= torch.empty((V * S, 1), device="cuda").uniform_(0.6, 1.0, generator=gen)
    q = torch.randn((nq, d), generator=gen, device="cuda")
    q /= q.norm(dim=1, keepdim=True)
    return dict(V=V, S=S, d=d, nq=nq, g=g.contiguous(), q=q.contiguous())


@pytest.mark.parametrize("dtype", ["f32", "bf16"])
def test_cfg2_full_size_segmax_vs_oracle(vq, cfg2, dtype):
    import torch
    V, S, nq = cfg2["V"], cfg2["S"], cfg2["nq"]
    g, q = cfg2["g"], cfg2["q"]
    if dtype == "bf16":
        g, q = g.bfloat16(), q.bfloat16()
    ix = vq.IndexFlatL2(cfg2["d"])
    ix.add(g)
    seg = np.arange(0, V * S + 1, S, dtype=np.int64)
    ix.set_segments(seg)
    vid, row, D = ix.search_segmax(q, 100)
    st = ix.last_stats()
    vid, row, D = vid.cpu().numpy(), row.cpu().numpy(), D.cpu().numpy()
    sub = np.random.default_rng(0).choice(nq, 32, replace=False)
    gh = g.float().cpu().numpy()
    qh = q.float().cpu().numpy()[sub]
    vo, ro, do = orc.segmax_topk(qh, gh, seg, 100)
    assert np.array_equal(vid[sub], vo) and np.array_equal(row[sub], ro) and _biteq(D[sub], do), st
    # size-independent properties over ALL queries: sorted, distinct videos, the row belongs to the video
    assert (np.diff(D, axis=1) >= 0).all()
    assert (row // S == vid).all()
    assert (np.sort(vid, axis=1)[:, 1:] != np.sort(vid, axis=1)[:, :-1]).all()


def test_cfg2_full_size_eval_moment_vs_oracle(vq, cfg2):
    import torch
    V, S, nq = cfg2["V"], cfg2["S"], cfg2["nq"]
    n = V * S
    rng = np.random.default_rng(1)
    ix = vq.IndexFlatL2(cfg2["d"])
    ix.add(cfg2["g"])
    video = (np.arange(n) // S).astype(np.int32)
    # the configs[2] grid: 8 start positions x 8 lengths per video
    start = (5 * (np.arange(n) % 8)).astype(np.float32)
    end = start + 5 * (1 + (np.arange(n) % S) // 8).astype(np.float32)
    hit = rng.integers(-1, 180, n).astype(np.int32)
    iou = np.where(hit >= 0, rng.uniform(0, 1, n), 0).astype(np.float32)
    ix.set_meta(hit, video)
    ix.set_segments(np.arange(0, n + 1, S, dtype=np.int64))
    ix.set_moments(start, end, iou, hit)
    sub = rng.choice(nq, 16, replace=False)
    q = cfg2["q"][torch.from_numpy(sub).cuda()].contiguous()
    ql = rng.integers(0, 180, 16).astype(np.int32)
    ign = rng.integers(0, V, (16, 5)).astype(np.int32)
    ap, rec, nk, _ = ix.eval_moment(q, ql, ign)
    apo, reco = orc.eval_moment(q.cpu().numpy(), cfg2["g"].cpu().numpy(), video, start, end, iou, hit, ql, ign)
    np.testing.assert_allclose(ap.cpu().numpy(), apo, rtol=0, atol=1e-12)
    np.testing.assert_allclose(rec.cpu().numpy(), reco, rtol=0, atol=1e-12)
