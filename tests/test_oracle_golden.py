"""CPU: the oracle (oracle/vq_oracle.c) against the fixtures minted from the reference's own
code (tests/golden/make_golden.py).  Tolerances: indices exact; float64 metrics 1e-12;
fp32 loss/grad 1e-6 absolute (north_star)."""
import numpy as np

from oracle import oracle as orc


def test_nms_known_answer(golden):
    # unittests/dataset_tests.py:32-37 input; the reference's nms_cpu returns [1, 2]
    assert golden["nms_kat_keep"].tolist() == [1, 2]
    assert orc.nms1d(golden["nms_kat_dets"], 0.5) == [1, 2]


def test_nms_random_cases(golden):
    off, koff = 0, golden["nms_keep_off"]
    for c, n in enumerate(golden["nms_n"]):
        dets = golden["nms_dets"][off:off + n]
        off += n
        assert orc.nms1d(dets, 0.5) == golden["nms_keep"][koff[c]:koff[c + 1]].tolist()


def test_hit_list_ap_recall(golden):
    off = 0
    for c, n in enumerate(golden["hl_n"]):
        s = slice(off, off + n)
        off += n
        ap, rec = orc.eval_hit_list(golden["hl_tp"][s], golden["hl_ig"][s], golden["hl_sc"][s])
        assert abs(ap - golden["hl_ap"][c]) < 1e-12
        np.testing.assert_allclose(rec, golden["hl_rec"][c], rtol=0, atol=1e-12)


def _trimmed_inputs(golden, query_num):
    feat = golden["tr_feat"]
    rows = golden["tr_q_list_rows"]
    q = feat[rows[:, :query_num]].astype(np.float32)
    # dataloader_baseline.py:1588-1591: running sum then divide, in fp32
    acc = np.zeros_like(q[:, 0])
    for i in range(query_num):
        acc = acc + q[:, i]
    q = (acc / np.float32(query_num)).astype(np.float32) if query_num > 1 else q[:, 0]
    ign = golden["tr_video"][rows]
    return q, feat, golden["tr_label"], golden["tr_video"], golden["tr_label"][rows[:, 0]], ign


def _names_types(golden, qlabel):
    names = [str(x) for x in golden["label_names"]]
    cls = [names[i] for i in qlabel]
    typ = ["base" if i < 6 else "novel" for i in qlabel]  # make_golden: 6 train + 5 test labels
    return cls, typ


def test_trimmed_ranking_matches_reference(golden):
    q, g, gl, gv, ql, ign = _trimmed_inputs(golden, 1)
    ap, rec, top_idx, _ = orc.eval_trimmed(q, g, gl, gv, ql, ign)
    np.testing.assert_allclose(ap, golden["tr_ap"], rtol=0, atol=1e-12)
    cls, typ = _names_types(golden, ql)
    agg = orc.aggregate(ap, rec, cls, typ)
    np.testing.assert_allclose([agg["ap"], agg["base_map"], agg["novel_map"]], golden["tr_result"],
                               rtol=0, atol=1e-12)
    got = np.array([[agg[k][str(n)] for n in (30, 50, 100)]
                    for k in ("recall", "base_recall", "novel_recall")])
    np.testing.assert_allclose(got, golden["tr_recall"], rtol=0, atol=1e-12)
    # confusion-matrix side data (:441-442): labels of the first 100 survivors, in rank order
    assert (gl[top_idx] == golden["tr_cm_labels_top100"]).all()


def test_trimmed_ranking_query_num5(golden):
    q, g, gl, gv, ql, ign = _trimmed_inputs(golden, 5)
    ap, rec, _, _ = orc.eval_trimmed(q, g, gl, gv, ql, ign)
    np.testing.assert_allclose(ap, golden["tr5_ap"], rtol=0, atol=1e-12)
    cls, typ = _names_types(golden, ql)
    agg = orc.aggregate(ap, rec, cls, typ)
    np.testing.assert_allclose([agg["ap"], agg["base_map"], agg["novel_map"]], golden["tr5_result"],
                               rtol=0, atol=1e-12)


def test_moment_ranking_matches_reference(golden):
    rows = golden["mo_q_list_rows"]
    q = golden["mo_qfeat"][rows[:, 0]]
    ql = golden["mo_qlabel"][rows[:, 0]]
    ign = golden["mo_qvideo"][rows]
    ap, rec = orc.eval_moment(q, golden["mo_feat"], golden["mo_video"], golden["mo_start"],
                              golden["mo_end"], golden["mo_iou"], golden["mo_hit"], ql, ign)
    np.testing.assert_allclose(ap, golden["mo_ap"], rtol=0, atol=1e-12)
    cls, typ = _names_types(golden, ql)
    agg = orc.aggregate(ap, rec, cls, typ)
    np.testing.assert_allclose([agg["ap"], agg["base_map"], agg["novel_map"]], golden["mo_result"],
                               rtol=0, atol=1e-12)
    got = np.array([[agg[k][str(n)] for n in (30, 50, 100)]
                    for k in ("recall", "base_recall", "novel_recall")])
    np.testing.assert_allclose(got, golden["mo_recall"], rtol=0, atol=1e-12)


def test_pdist_and_triplet(golden):
    emb, labels, margin = golden["tl_emb"], golden["tl_labels"], float(golden["tl_margin"])
    np.testing.assert_allclose(orc.pdist(emb), golden["tl_pdist"], rtol=0, atol=2e-6)
    loss, grad, _, _ = orc.triplet_fwd_bwd(emb, golden["tl_semihard_trip"], margin)
    assert abs(loss - float(golden["tl_semihard_loss"])) < 1e-6
    np.testing.assert_allclose(grad, golden["tl_semihard_grad"], rtol=0, atol=1e-6)
    loss, grad, trip = orc.batchhard_triplet(emb, labels, margin)
    assert (trip == golden["tl_bh_trip"]).all()
    assert abs(loss - float(golden["tl_bh_loss"])) < 1e-6
    np.testing.assert_allclose(grad, golden["tl_bh_grad"], rtol=0, atol=1e-6)


def test_pool_normalize(golden):
    np.testing.assert_allclose(orc.pool_normalize(golden["pool_in"]), golden["pool_out"], rtol=0,
                               atol=1e-6)


def test_search_is_sorted_with_index_tiebreak():
    rng = np.random.RandomState(3)
    g = rng.randn(200, 16).astype(np.float32)
    g[50] = g[10]
    q = g[[10, 77]]
    D, I = orc.search(q, g, 200)
    assert (np.diff(D, axis=1) >= 0).all()
    assert I[0, 0] == 10 and I[0, 1] == 50 and D[0, 0] == 0 and D[0, 1] == 0
    assert sorted(I[1].tolist()) == list(range(200))
