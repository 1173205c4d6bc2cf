"""Host-side mirror of the reference's metric utilities: same names,
arguments and return values; the per-list arithmetic runs in the CUDA extension.

    Average, generate_multi_query, evaluation_metric   dataloader_baseline.py:292-658
    average_precision_score                            sklearn, as called at :391 / :461
    nms_cpu                                            misc_utils/utils_models.py:153-174
"""
import ctypes as C
import random

import numpy as np

from . import _lib as L

R_AT_N = (30, 50, 100)  # dataloader_baseline.py:332


def Average(lst):  # :292-293
    return sum(lst) / (len(lst) + 1e-10)


def _hmean2(a, b):  # scipy.stats.hmean([a, b]) for positive a, b (:576, :590)
    return 2.0 / (1.0 / a + 1.0 / b)


def generate_multi_query(query_list):
    """:296-322 -- each query followed by 4 random same-class items from other videos."""
    random.seed(620)
    cls_dict = {}
    for q in query_list:
        cls_dict.setdefault(q["label"], []).append(q)
    out = []
    for q in query_list:
        same = [x for x in cls_dict[q["label"]] if x["video_id"] != q["video_id"]]
        out.append([q] + random.choices(same, k=4))
    return out


def average_precision_batch(y_true_lists, y_score_lists):
    """average_precision_score for many (y_true, y_score) lists in one launch."""
    n = len(y_true_lists)
    off = np.zeros(n + 1, np.int64)
    for i, y in enumerate(y_true_lists):
        off[i + 1] = off[i] + len(y)
    yt = np.ascontiguousarray(np.concatenate([np.asarray(y) for y in y_true_lists]) if n else [], dtype=np.int32)
    ys = np.ascontiguousarray(np.concatenate([np.asarray(s) for s in y_score_lists]) if n else [], dtype=np.float64)
    ap = np.empty(n, np.float64)
    L.check(L.lib().vq_average_precision(L.ptr(yt), L.ptr(ys), L.ptr(off), C.c_int64(n), L.ptr(ap), C.c_void_p(0)))
    return ap


def average_precision_score(y_true, y_score):
    return float(average_precision_batch([y_true], [y_score])[0])


def nms_segmented(dets, offsets, thresh):
    """nms_cpu on many segments at once; returns a uint8 keep mask over the rows of dets."""
    dets = np.ascontiguousarray(dets, dtype=np.float64).reshape(-1, 3)
    off = np.ascontiguousarray(offsets, dtype=np.int64)
    keep = np.zeros(dets.shape[0], np.uint8)
    L.check(L.lib().vq_nms1d_segmented(L.ptr(dets), L.ptr(off), C.c_int64(off.shape[0] - 1), C.c_double(thresh),
                                       L.ptr(keep), C.c_void_p(0)))
    return keep


NMS1D_STATIC_MAX = 1048576  # VQ_NMS1D_STATIC_MAX (include/vqwild.h)


def nms_segmented_static(dets, offsets, thresh, out=None):
    """nms_segmented without any allocation inside the library (vq_nms1d_segmented_static: the scratch is a static
    device array of NMS1D_STATIC_MAX bytes).  dets float64[total,3] and offsets int64[nseg+1] are CUDA tensors; `out`
    (uint8[total], CUDA) is allocated here when the caller does not pass one.  Runs on the current stream, no sync;
    calls that overlap on the GPU race on the shared scratch.  No counterpart in the reference."""
    import torch
    assert dets.is_cuda and offsets.is_cuda and dets.dtype == torch.float64 and offsets.dtype == torch.int64
    total = dets.shape[0]
    if out is None:
        out = torch.empty(total, dtype=torch.uint8, device=dets.device)
    L.check(L.lib().vq_nms1d_segmented_static(L.ptr(dets), L.ptr(offsets), C.c_int64(offsets.shape[0] - 1),
                                              C.c_int64(total), C.c_double(thresh), L.ptr(out), L.stream_of(dets)))
    return out


def nms_cpu(dets, thresh):
    """Same contract as the reference's nms_cpu: indices into dets in descending-score order."""
    dets = np.ascontiguousarray(dets, dtype=np.float64).reshape(-1, 3)
    keep = nms_segmented(dets, [0, dets.shape[0]], thresh)
    idx = np.nonzero(keep)[0]
    order = np.lexsort((idx, -dets[idx, 2]))  # score desc, row asc
    return idx[order].tolist()


class evaluation_metric:
    """Accumulator with the reference's interface (:325-658).  `add2dict` /
    `multiprocess_calculate` accept the reference's list-of-dict hit lists; `update_arrays` takes
    the per-query vectors the fused kernels produce."""

    def __init__(self, args=None, query_list=()):
        self.args = args
        self.r_at_n = list(R_AT_N)
        self.class_dict = {}
        self.class_agnostic_ap = []
        self.base_classes, self.novel_classes = [], []
        self.full_retrieval_top = {str(n): [] for n in self.r_at_n}
        self.base_retrieval_top = {str(n): [] for n in self.r_at_n}
        self.novel_retrieval_top = {str(n): [] for n in self.r_at_n}
        self.avg2_full_retrieval_top = {str(n): {} for n in self.r_at_n}
        self.cm_gt_labels, self.cm_labels = [], []
        self.set_class_info(query_list)

    def set_class_info(self, query_list):  # :371-384 -- one entry per QUERY, not per class
        for q in query_list:
            q = q[0]
            if q["retrieval_type"] == "base":
                self.base_classes.append(q["label"])
            elif q["retrieval_type"] == "novel":
                self.novel_classes.append(q["label"])
            else:
                raise ValueError("query of retrieval_type %r" % (q["retrieval_type"],))

    def multiprocess_calculate(self, single_query_hit):  # :386-402
        hits = [d for d in single_query_hit if not d["ignore"]]
        y_true = np.array([d["tp"] for d in hits])
        recall_list = []
        npos = float(y_true.sum()) + 1e-10
        for n in self.r_at_n:
            recall_list.append(float(y_true[:n].sum()) * 1.0 / npos)
        y_true = y_true.copy()
        y_true[-1] = 1  # for robust mAP (:389)
        ap = average_precision_score(y_true, np.array([d["score"] for d in hits], np.float64))
        return dict(ap=ap, recall_list=recall_list)

    def multiprocess_update(self, cls_name, retrieval_type, r_dict):  # :404-429
        ap, recall_list = float(r_dict["ap"]), r_dict["recall_list"]
        self.class_agnostic_ap.append(ap)
        self.class_dict.setdefault(cls_name, []).append(ap)
        for j, n in enumerate(self.r_at_n):
            r = float(recall_list[j])
            self.full_retrieval_top[str(n)].append(r)
            self.avg2_full_retrieval_top[str(n)].setdefault(cls_name, []).append(r)
            if retrieval_type == "base":
                self.base_retrieval_top[str(n)].append(r)
            elif retrieval_type == "novel":
                self.novel_retrieval_top[str(n)].append(r)
            else:
                raise ValueError(retrieval_type)

    def add2dict(self, cls_name, retrieval_type, single_query_hit):  # :431-496
        hits = [d for d in single_query_hit if not d["ignore"]]
        if hits and "gt_label" in hits[0]:
            self.cm_gt_labels.extend(i["gt_label"] for i in hits[:100])
        if hits and "label" in hits[0]:
            self.cm_labels.extend(i["label"] for i in hits[:100])
        self.multiprocess_update(cls_name, retrieval_type, self.multiprocess_calculate(single_query_hit))

    def update_arrays(self, cls_names, retrieval_types, ap, recall):
        for c, t, a, r in zip(cls_names, retrieval_types, np.asarray(ap), np.asarray(recall)):
            self.multiprocess_update(c, t, dict(ap=a, recall_list=r))

    def get_result(self, pass_content=()):  # :498-658 (call once: it folds the lists, as the reference does)
        for n in self.r_at_n:
            k = str(n)
            self.full_retrieval_top[k] = Average(self.full_retrieval_top[k])
            self.base_retrieval_top[k] = Average(self.base_retrieval_top[k])
            self.novel_retrieval_top[k] = Average(self.novel_retrieval_top[k])
            for c, lst in self.avg2_full_retrieval_top[k].items():
                self.avg2_full_retrieval_top[k][c] = Average(lst)
        for c, lst in self.class_dict.items():
            self.class_dict[c] = Average(lst)
        base_map = Average([self.class_dict[c] for c in self.base_classes])
        novel_map = Average([self.class_dict[c] for c in self.novel_classes])
        o2_hmean = _hmean2(base_map + 1e-10, novel_map + 1e-10)
        cm_dict = dict(gt_labels=self.cm_gt_labels, label=self.cm_labels, base_classes=self.base_classes,
                       novel_classes=self.novel_classes, class_map_dict=self.class_dict,
                       pass_content=list(pass_content))
        return dict(ap=o2_hmean, base_map=base_map, novel_map=novel_map, recall=self.full_retrieval_top,
                    base_recall=self.base_retrieval_top, novel_recall=self.novel_retrieval_top, cm_dict=cm_dict)


# ---------------------------------------------------------------------------------------------
# BASELINE.json configs[4]: k-sweep on top-k search results (ELMo 1024-d split 40_20_140)
# ---------------------------------------------------------------------------------------------
def recall_at_k(tp_ranked, n_pos, ks=(5, 30, 50, 100)):
    """R@k as the reference defines it (dataloader_baseline.py:475-481): sum_{i<k} tp_i / (n_pos + 1e-10), for every
    k of the sweep (5 from word_embed/wordembed_retrieval_test.py:20; 30/50/100 from r_at_n :332).
    tp_ranked [nq, K] 0/1 in rank order (survivors only), n_pos [nq] = positives in the whole list."""
    tp = np.asarray(tp_ranked, np.float64)
    den = np.asarray(n_pos, np.float64) + 1e-10
    return {int(k): tp[:, :k].sum(1) / den for k in ks}


def ap_at_k(tp_ranked, n_pos, ks=(5, 30, 50, 100)):
    """AP@k.  The reference has no mAP@k (its AP is over the full list, :461); BASELINE.json configs[4] asks for a
    sweep, so it is DEFINED here: AP@k = sum_{i<=k} P(i) * rel_i / min(n_pos, k), P(i) = precision
    at rank i; 0 for queries without positives."""
    tp = np.asarray(tp_ranked, np.float64)
    prec = np.cumsum(tp, 1) / np.arange(1, tp.shape[1] + 1)
    npos = np.asarray(n_pos, np.float64)
    out = {}
    for k in ks:
        den = np.minimum(npos, k)
        out[int(k)] = np.where(den > 0, (prec[:, :k] * tp[:, :k]).sum(1) / np.maximum(den, 1), 0.0)
    return out
