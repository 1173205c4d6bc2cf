"""ctypes front-end of the CPU oracle (oracle/vq_oracle.c) plus the Python-level restatement
of the reference's metric aggregation.

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.

Citations are file:line into /root/reference.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libvqoracle.so")
_lib = None

R_AT_N = (30, 50, 100)  # dataloader_baseline.py:332


def build(force=False):
    """Compile the C restatement in-tree (gcc only)."""
    src = os.path.join(_HERE, "vq_oracle.c")
    if not force and os.path.exists(_SO) and os.path.getmtime(_SO) >= os.path.getmtime(src):
        return _SO
    base = ["-O2", "-fPIC", "-shared", "-ffp-contract=off"]
    last = None
    for cc in ("/usr/bin/gcc", "gcc", "cc"):
        for omp in (["-fopenmp"], []):
            try:
                subprocess.run([cc] + base + omp + ["-o", _SO, src, "-lm"], check=True,
                               capture_output=True)
                return _SO
            except (subprocess.CalledProcessError, FileNotFoundError) as e:  # try next
                last = e
    raise RuntimeError("cannot build oracle: %r" % (last,))


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_ap.restype = C.c_double
        _lib.orc_ap_sorted.restype = C.c_double
        _lib.orc_nms1d.restype = C.c_int64
        _lib.orc_batchhard_mine.restype = C.c_int64
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def num_threads():
    return lib().orc_num_threads()


def set_num_threads(n):
    lib().orc_set_num_threads(C.c_int(int(n)))


def l2sq(q, g):
    q, g = _f32(q), _f32(g)
    out = np.empty((q.shape[0], g.shape[0]), np.float32)
    lib().orc_l2sq(_p(q), _p(g), C.c_int64(q.shape[0]), C.c_int64(g.shape[0]),
                   C.c_int(q.shape[1]), _p(out))
    return out


def bf16_round(x):
    x = _f32(x)
    out = np.empty_like(x)
    lib().orc_bf16_round_array(_p(x), C.c_int64(x.size), _p(out))
    return out


def search(q, g, k):
    """faiss.IndexFlatL2.search stand-in (dataloader_baseline.py:1592): D ascending, I."""
    q, g = _f32(q), _f32(g)
    nq, n = q.shape[0], g.shape[0]
    k = int(k)
    D = np.full((nq, k), np.inf, np.float32)
    I = np.full((nq, k), -1, np.int64)
    lib().orc_search(_p(q), _p(g), C.c_int64(nq), C.c_int64(n), C.c_int(q.shape[1]),
                     C.c_int64(k), _p(D), _p(I))
    return D, I


def rank_row(d2row):
    d2row = _f32(d2row)
    order = np.empty(d2row.shape[0], np.int64)
    lib().orc_rank_row(_p(d2row), C.c_int64(d2row.shape[0]), _p(order))
    return order


def average_precision(y_true, y_score):
    """sklearn.metrics.average_precision_score restatement (dataloader_baseline.py:461)."""
    yt = _i32(y_true)
    ys = np.ascontiguousarray(y_score, dtype=np.float64)
    return float(lib().orc_ap(_p(yt), _p(ys), C.c_int64(yt.shape[0])))


def eval_hit_list(tp, ignore, score):
    """add2dict / multiprocess_calculate on one ranked hit list (:431-481 / :386-402)."""
    tp, ig = _i32(tp), _i32(ignore)
    sc = np.ascontiguousarray(score, dtype=np.float64)
    r = _i32(R_AT_N)
    ap = C.c_double()
    rec = np.empty(3, np.float64)
    lib().orc_eval_hit_list(_p(tp), _p(ig), _p(sc), C.c_int64(tp.shape[0]), _p(r), C.c_int(3),
                            C.byref(ap), _p(rec))
    return ap.value, rec


def eval_trimmed(q, g, g_label, g_video, q_label, q_ignore_vids, topk=100):
    """ARV_Retrieval.ranking per-query numbers (dataloader_baseline.py:1526-1610)."""
    q, g = _f32(q), _f32(g)
    nq, n = q.shape[0], g.shape[0]
    gl, gv, ql = _i32(g_label), _i32(g_video), _i32(q_label)
    qi = _i32(q_ignore_vids).reshape(nq, -1)
    ap = np.empty(nq, np.float64)
    rec = np.empty((nq, 3), np.float64)
    ti = np.empty((nq, topk), np.int64)
    td = np.empty((nq, topk), np.float32)
    lib().orc_eval_trimmed(_p(q), _p(g), C.c_int64(nq), C.c_int64(n), C.c_int(q.shape[1]),
                           _p(gl), _p(gv), _p(ql), _p(qi), C.c_int(qi.shape[1]), _p(ap),
                           _p(rec), C.c_int64(topk), _p(ti), _p(td))
    return ap, rec, ti, td


def nms1d(dets, thresh):
    """nms_cpu (misc_utils/utils_models.py:153-174)."""
    dets = np.ascontiguousarray(dets, dtype=np.float64).reshape(-1, 3)
    keep = np.empty(max(dets.shape[0], 1), np.int64)
    nk = lib().orc_nms1d(_p(dets), C.c_int64(dets.shape[0]), C.c_double(thresh), _p(keep))
    return keep[:nk].tolist()


def eval_moment(q, g, g_video, g_start, g_end, g_iou, g_hit_label, q_label, q_ignore_vids,
                nms_thresh=0.5):
    """ARV_Retrieval_Moment.ranking per-query numbers (dataloader_baseline.py:1225-1330)."""
    q, g = _f32(q), _f32(g)
    nq, n = q.shape[0], g.shape[0]
    qi = _i32(q_ignore_vids).reshape(nq, -1)
    gv, gs, ge = _i32(g_video), _f32(g_start), _f32(g_end)
    gi, gh, ql = _f32(g_iou), _i32(g_hit_label), _i32(q_label)
    ap = np.empty(nq, np.float64)
    rec = np.empty((nq, 3), np.float64)
    lib().orc_eval_moment(_p(q), _p(g), C.c_int64(nq), C.c_int64(n), C.c_int(q.shape[1]),
                          _p(gv), _p(gs), _p(ge), _p(gi), _p(gh), _p(ql), _p(qi),
                          C.c_int(qi.shape[1]), C.c_double(nms_thresh), _p(ap), _p(rec))
    return ap, rec


def eval_moment_kept(q1, g, g_video, g_start, g_end, g_iou, g_hit_label, q_label, ignore_vids, nms_thresh=0.5):
    """One query: (ap, recall[3], kept) with kept = the grouped, NMS-kept gallery rows in list order."""
    q1, g = _f32(q1).reshape(1, -1), _f32(g)
    n = g.shape[0]
    D, I = search(q1, g, n)
    order, d2 = np.ascontiguousarray(I[0]), np.ascontiguousarray(D[0])
    gv, gs, ge = _i32(g_video), _f32(g_start), _f32(g_end)
    gi, gh = _f32(g_iou), _i32(g_hit_label)
    ig = _i32(ignore_vids).reshape(-1)
    ap = C.c_double()
    rec = np.empty(3, np.float64)
    kept = np.empty(n, np.int64)
    nk = C.c_int64()
    lib().orc_eval_moment_ranked(_p(order), _p(d2), C.c_int64(n), _p(gv), _p(gs), _p(ge), _p(gi), _p(gh),
                                 C.c_int32(int(q_label)), _p(ig), C.c_int(ig.shape[0]), C.c_double(nms_thresh),
                                 C.byref(ap), _p(rec), _p(kept), C.byref(nk))
    return ap.value, rec, kept[:nk.value]


def segmax_topk(q, g, seg_start, k):
    q, g = _f32(q), _f32(g)
    ss = _i64(seg_start)
    nq, nseg = q.shape[0], ss.shape[0] - 1
    tv = np.empty((nq, k), np.int64)
    tr = np.empty((nq, k), np.int64)
    td = np.empty((nq, k), np.float32)
    lib().orc_segmax_topk(_p(q), _p(g), C.c_int64(nq), C.c_int64(g.shape[0]),
                          C.c_int(q.shape[1]), _p(ss), C.c_int64(nseg), C.c_int64(k), _p(tv),
                          _p(tr), _p(td))
    return tv, tr, td


def pdist(A, eps=1e-4):
    A = _f32(A)
    out = np.empty((A.shape[0], A.shape[0]), np.float32)
    lib().orc_pdist(_p(A), C.c_int64(A.shape[0]), C.c_int(A.shape[1]), C.c_float(eps), _p(out))
    return out


def batchhard_mine(dist, labels):
    dist, labels = _f32(dist), _i64(labels)
    B = dist.shape[0]
    trip = np.empty((B, 3), np.int32)
    T = lib().orc_batchhard_mine(_p(dist), _p(labels), C.c_int64(B), _p(trip))
    return trip[:T].copy()


def triplet_fwd_bwd(A, trip, margin):
    A = _f32(A)
    trip = _i32(trip).reshape(-1, 3)
    T = trip.shape[0]
    loss = C.c_float()
    grad = np.empty_like(A)
    per = np.empty(max(T, 1), np.float32)
    act = np.empty(max(T, 1), np.int32)
    lib().orc_triplet_fwd_bwd(_p(A), C.c_int64(A.shape[0]), C.c_int(A.shape[1]), _p(trip),
                              C.c_int64(T), C.c_float(margin), C.byref(loss), _p(grad), _p(per),
                              _p(act))
    return loss.value, grad, per[:T], act[:T]


def batchhard_triplet(A, labels, margin=1.0):
    trip = batchhard_mine(pdist(A), labels)
    loss, grad, _, _ = triplet_fwd_bwd(A, trip, margin)
    return loss, grad, trip


def pool_normalize(x):
    x = _f32(x)
    B, Cc, T = x.shape
    out = np.empty((B, Cc), np.float32)
    lib().orc_pool_normalize(_p(x), C.c_int64(B), C.c_int(Cc), C.c_int(T), _p(out))
    return out


def topk_merge(d2, idx):
    d2, idx = _f32(d2), _i64(idx)
    G, nq, k = d2.shape
    od = np.empty((nq, k), np.float32)
    oi = np.empty((nq, k), np.int64)
    lib().orc_topk_merge(_p(d2), _p(idx), C.c_int(G), C.c_int64(nq), C.c_int64(k), _p(od),
                         _p(oi))
    return od, oi


# ---------------------------------------------------------------------------------------
# evaluation_metric aggregation (multiprocess_update :404-429 + get_result :498-658),
# restated on per-query vectors.  Python floats, sequential sum(), exactly as the reference.
# ---------------------------------------------------------------------------------------
def _average(lst):  # dataloader_baseline.py:292-293
    return sum(lst) / (len(lst) + 1e-10)


def _hmean2(a, b):  # scipy.stats.hmean of two positives (:576, :590): n / sum(1/x)
    return 2.0 / (1.0 / a + 1.0 / b)


def aggregate(ap, recall, cls_names, retrieval_types):
    """ap[Q], recall[Q,3], class name and 'base'/'novel' per query (query order) ->
    the numeric part of get_result's dict (:650-658)."""
    ap = [float(x) for x in ap]
    class_dict, base_classes, novel_classes = {}, [], []
    for a, c, t in zip(ap, cls_names, retrieval_types):
        class_dict.setdefault(c, []).append(a)
        (base_classes if t == "base" else novel_classes).append(c)  # one entry per QUERY (:380)
    out_recall, out_base_recall, out_novel_recall = {}, {}, {}
    for j, n in enumerate(R_AT_N):
        col = [float(r[j]) for r in recall]
        out_recall[str(n)] = _average(col)
        out_base_recall[str(n)] = _average([x for x, t in zip(col, retrieval_types) if t == "base"])
        out_novel_recall[str(n)] = _average([x for x, t in zip(col, retrieval_types) if t == "novel"])
    class_mean = {c: _average(v) for c, v in class_dict.items()}
    base_map = _average([class_mean[c] for c in base_classes])
    novel_map = _average([class_mean[c] for c in novel_classes])
    return dict(ap=_hmean2(base_map + 1e-10, novel_map + 1e-10), base_map=base_map,
                novel_map=novel_map, recall=out_recall, base_recall=out_base_recall,
                novel_recall=out_novel_recall, class_map_dict=class_mean)
