"""CPU: the oracle and the host-side mirrors against the REAL-METADATA goldens (tests/golden/make_golden_real.py:
the reference's own evaluators run on data_generate/arv_db_100_20_80*.json with seeded features).  SURVEY.md 8c
case (1)."""
import os
import sys

import numpy as np
import pytest

from oracle import oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from synth import class_feats  # noqa: E402


@pytest.fixture(scope="module")
def real():
    return np.load(os.path.join(HERE, "golden", "reference_golden_real.npz"), allow_pickle=False)


def _agg(real, pre, ap, rec, qlabel, qrtype):
    cls = ["c%d" % x for x in qlabel]
    typ = [("base", "novel")[t] for t in qrtype]
    agg = orc.aggregate(ap, rec, cls, typ)
    np.testing.assert_allclose([agg["ap"], agg["base_map"], agg["novel_map"]], real[pre + "_result"], rtol=0, atol=1e-12)


def test_real_trimmed_oracle(real):
    """6 338 queries x 7 670 items x 512-d, the real label / video_id / is_query structure."""
    lab, vid = real["rt_label"].astype(np.int32), real["rt_video"].astype(np.int32)
    feat = class_feats(11, lab, 512)
    rows = real["rt_q_list_rows"].astype(np.int64)
    assert rows.shape == (6338, 5) and len(lab) == 7670
    assert (real["rt_is_query"][rows[:, 0]] == 1).all()
    q = np.ascontiguousarray(feat[rows[:, 0]])
    ap, rec, _, _ = orc.eval_trimmed(q, feat, lab, vid, lab[rows[:, 0]], vid[rows])
    np.testing.assert_allclose(ap, real["rt_ap"], rtol=0, atol=1e-12)
    _agg(real, "rt", ap, rec, lab[rows[:, 0]], real["rt_rtype"][rows[:, 0]])


def test_real_clip_oracle(real):
    lab, vid = real["cl_label"].astype(np.int32), real["cl_video"].astype(np.int32)
    cfeat, qfeat = class_feats(21, lab, 64), class_feats(22, real["cl_qlabel"], 64)
    rows = real["cl_q_list_rows"]
    ql, qv = real["cl_qlabel"].astype(np.int32), real["cl_qvideo"].astype(np.int32)
    ap, rec, _, _ = orc.eval_trimmed(np.ascontiguousarray(qfeat[rows[:, 0]]), cfeat, lab, vid, ql[rows[:, 0]], qv[rows])
    np.testing.assert_allclose(ap, real["cl_ap"], rtol=0, atol=1e-12)
    _agg(real, "cl", ap, rec, ql[rows[:, 0]], real["cl_qrtype"])


def test_real_moment_oracle(real):
    """Ragged real videos incl. the 3 523-moment one; tp from the float64 IoU (the flag is taken before narrowing)."""
    hit, vid = real["mr_hit"].astype(np.int32), real["mr_video"].astype(np.int32)
    mfeat, qfeat = class_feats(31, hit + 1, 64, spread=0.6), class_feats(32, real["mr_qlabel"] + 1, 64, spread=0.6)
    rows = real["mr_q_list_rows"]
    ql, qv = real["mr_qlabel"].astype(np.int32), real["mr_qvideo"].astype(np.int32)
    flag = (real["mr_iou"] >= 0.5).astype(np.float32)
    ap, rec = orc.eval_moment(np.ascontiguousarray(qfeat[rows[:, 0]]), mfeat, vid, real["mr_start"].astype(np.float32),
                              real["mr_end"].astype(np.float32), flag, hit, ql[rows[:, 0]], qv[rows])
    np.testing.assert_allclose(ap, real["mr_ap"], rtol=0, atol=1e-12)
    _agg(real, "mr", ap, rec, ql[rows[:, 0]], real["mr_qrtype"])


def _annotations(real, pre, v, label_of=lambda x: "c%d" % x):
    m = real[pre + "_ann_video"] == v
    return [dict(segment=[float(s[0]), float(s[1])], label=label_of(l) if l >= 0 else "other")
            for s, l in zip(real[pre + "_ann_seg"][m], real[pre + "_ann_label"][m])]


def test_moment_grid_and_closest_hits_match_reference_garner_feat(real):
    """retrieval.moment_grid / closest_hits (the host half of f1) re-derive, video by video, exactly the moments the
    reference's garner_feat produced (:1091-1151): same grid, same order, same closest-hit label, IoU bit-equal."""
    from vqwild_b200.retrieval import closest_hits, moment_grid
    possible = {"c%d" % i for i in range(1000)}
    vid = real["mr_video"]
    off = 0
    for v in range(len(real["mr_vid_duration"])):
        s, e, lo, hi = moment_grid(real["mr_vid_duration"][v], int(real["mr_vid_T"][v]))
        n = len(s)
        sl = slice(off, off + n)
        assert (vid[sl] == v).all() and (off + n == len(vid) or vid[off + n] != v)
        assert np.array_equal(s, real["mr_start"][sl]) and np.array_equal(e, real["mr_end"][sl])
        labs, iou = closest_hits(s, e, _annotations(real, "mr", v), possible)
        want = real["mr_hit"][sl]
        assert [(-1 if x is None else int(x[1:])) for x in labs] == want.tolist()
        assert np.array_equal(iou.view(np.uint64), real["mr_iou"][sl].view(np.uint64))
        off += n
    assert off == len(vid)


def test_clip_grid_and_cal_label_match_reference_garner_feat(real):
    """retrieval.clip_grid / cal_label against the reference's clip garner_feat (:790-834)."""
    from vqwild_b200.retrieval import cal_label, clip_grid
    possible = {"c%d" % i for i in range(1, 1000)}
    vid, loc, lab = real["cl_video"], real["cl_loc_sec"], real["cl_label"]
    for v in np.unique(vid):
        m = vid == v
        dur = real["cl_vid_duration"][m][0]
        s, e, lo, hi = clip_grid(dur, 10 ** 9)
        # the reference dropped the clips whose frame window ran past the features; the rest is a prefix of the grid
        k = int(m.sum())
        assert np.array_equal(s[:k], loc[m][:, 0]) and np.array_equal(e[:k], loc[m][:, 1])
        ann = _annotations(real, "cl", v)
        got = [cal_label((int(a), int(b)), ann, possible) for a, b in zip(s[:k], e[:k])]
        assert [0 if x == "unknown" else int(x[1:]) for x in got] == lab[m].tolist()
