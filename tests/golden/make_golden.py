"""Mint golden vectors by running the REFERENCE's own Python code (authoring container only).

    python tests/golden/make_golden.py        # needs /root/reference; writes reference_golden.npz

The reference holds no golden values for this path, so these fixtures are the
pin: inputs are seeded synthetic features (the reference's own --memory_leak_debug idea,
dataloader_baseline.py:1471-1474) on label names taken from the reference's dataset_config;
outputs come from the unmodified reference modules:
  * misc_utils/utils_models.py:nms_cpu
  * dataloader_baseline.py:evaluation_metric (+ sklearn average_precision_score, scipy hmean)
  * dataloader_baseline.py:ARV_Retrieval.ranking, ARV_Retrieval_Moment.ranking
  * misc_utils/metric_learning_loss.py:Sampler.pdist, TripletLoss (semihard), and a batch-hard
    loss composed from pdist + triplet_distance
  * main.py:feat_func's F.normalize + the evaluator's np.mean pooling
Two shims are injected before import: `termcolor` (cosmetic) and `faiss` (not installable
here).  The faiss stand-in returns squared L2 ascending over all N using the oracle's
canonical fp32 arithmetic, ties -> lowest index -- that boundary stays "parity unpinned".
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

from oracle import oracle as orc  # noqa: E402


def install_shims():
    tc = types.ModuleType("termcolor")
    tc.colored = lambda s, *a, **k: s
    sys.modules["termcolor"] = tc

    class IndexFlatL2:
        def __init__(self, d):
            self.d, self.xb, self.ntotal = d, np.zeros((0, d), np.float32), 0

        def add(self, xb):
            self.xb = np.concatenate([self.xb, np.ascontiguousarray(xb, np.float32)], 0)
            self.ntotal = self.xb.shape[0]

        def search(self, xq, k):
            return orc.search(xq, self.xb, k)

    fa = types.ModuleType("faiss")
    fa.IndexFlatL2 = IndexFlatL2
    sys.modules["faiss"] = fa


class Args(dict):
    """argparse.Namespace look-alike that also supports `"x" in args` (:1559)."""
    __getattr__ = dict.__getitem__


def main():
    install_shims()
    os.chdir(REF)
    import torch
    import dataloader_baseline as dl
    from misc_utils.utils_models import nms_cpu
    from misc_utils import metric_learning_loss as mll

    out = {}
    rng = np.random.RandomState(0)

    # ---- nms_cpu ------------------------------------------------------------------
    kat = np.array([[0, 10, 0.6], [3, 13, 0.7], [12.8, 50, 0.1]])  # unittests/dataset_tests.py:32-37
    out["nms_kat_dets"], out["nms_kat_keep"] = kat, np.array(nms_cpu(kat, 0.5))
    nms_in, nms_keep, nms_off = [], [], [0]
    for case in range(24):
        n = int(rng.randint(1, 40))
        st = rng.randint(0, 40, n) * 5.0
        ln = rng.randint(1, 27, n) * 5.0
        sc = -rng.rand(n)  # distinct scores (ties unpinned in numpy's argsort)
        dets = np.stack([st, st + ln, sc], 1)
        nms_in.append(dets)
        nms_keep.append(np.array(nms_cpu(dets, 0.5), np.int64))
        nms_off.append(nms_off[-1] + len(nms_keep[-1]))
    out["nms_dets"] = np.concatenate(nms_in, 0)
    out["nms_n"] = np.array([d.shape[0] for d in nms_in])
    out["nms_keep"] = np.concatenate(nms_keep)
    out["nms_keep_off"] = np.array(nms_off)

    # ---- label universe from the reference's dataset_config ------------------------
    cfg = dl.dataset_config["100_20_80"]
    train_l, test_l, val_l = cfg["arv_train_label"], cfg["arv_test_label"], cfg["arv_val_label"]
    names = list(train_l[:6]) + list(test_l[:5]) + [val_l[0], dl.NOISE_LABEL]
    out["label_names"] = np.array(names)
    args = Args(meta_split="100_20_80", eval_split="testing", query_num=1, debug=False,
                memory_leak_debug=True, metric_feat_dim=32)

    # ---- evaluation_metric on random hit lists (incl. score ties) -------------------
    qstub = [[dict(label=names[0], retrieval_type="base")]]
    em = dl.evaluation_metric(args, qstub)
    hl_tp, hl_ig, hl_sc, hl_n, hl_ap, hl_rec = [], [], [], [], [], []
    for case in range(40):
        n = int(rng.randint(2, 400))
        tp = (rng.rand(n) < 0.1).astype(np.int64)
        ig = (rng.rand(n) < 0.05)
        ig[rng.randint(n)] = False
        sc = -rng.rand(n).astype(np.float32)
        if case % 3 == 0:  # ties
            sc = np.round(sc, 1)
        order = np.argsort(-sc, kind="stable")
        tp, ig, sc = tp[order], ig[order], sc[order]
        hits = [dict(tp=int(t), ignore=bool(i), score=float(s)) for t, i, s in zip(tp, ig, sc)]
        r = em.multiprocess_calculate(hits)
        hl_tp.append(tp), hl_ig.append(ig.astype(np.int64)), hl_sc.append(sc.astype(np.float64))
        hl_n.append(n), hl_ap.append(r["ap"]), hl_rec.append(r["recall_list"])
    out["hl_tp"], out["hl_ig"] = np.concatenate(hl_tp), np.concatenate(hl_ig)
    out["hl_sc"], out["hl_n"] = np.concatenate(hl_sc), np.array(hl_n)
    out["hl_ap"], out["hl_rec"] = np.array(hl_ap), np.array(hl_rec)

    # ---- ARV_Retrieval.ranking (trimmed) --------------------------------------------
    D = 32
    N = 360
    lab = rng.randint(0, len(names), N)
    lab[:len(names) * 6] = np.repeat(np.arange(len(names)), 6)  # >= 6 per class
    vid = np.arange(N) // 2 * 2 + (rng.rand(N) < 0.7) * (np.arange(N) % 2)  # some shared video ids
    centre = rng.randn(len(names), D).astype(np.float32)
    feat = rng.randn(N, D).astype(np.float32) + 0.9 * centre[lab]
    feat /= np.linalg.norm(feat, axis=1, keepdims=True)
    feat *= rng.uniform(0.6, 1.0, (N, 1)).astype(np.float32)
    feat[7] = feat[3]  # exact duplicate rows -> score ties
    feat = np.ascontiguousarray(feat, np.float32)
    is_query = (rng.rand(N) < 0.6).astype(np.int64)
    items = []
    for i in range(N):
        nm = names[lab[i]]
        rt = "base" if nm in train_l else ("novel" if nm in test_l else "noise")
        items.append(dict(label=nm, video_id="v_%04d" % vid[i], is_query=int(is_query[i]),
                          retrieval_type=rt, segment=[0.0, float(i % 7 + 1)], frame_info=i,
                          feat=feat[i].copy()))
    possible = list(train_l) + list(test_l)
    ev = object.__new__(dl.ARV_Retrieval)
    ev.args, ev.eval_split = args, "testing"
    ev.gallery_list = items
    ev.query_list = [it for it in items if it["label"] in possible]
    import faiss
    ev.faiss_index = faiss.IndexFlatL2(D)
    ev.faiss_index.add(feat)
    ev.faiss_xb = feat
    res = ev.ranking()
    qlists = ev.query_list  # multi-query lists, filtered to is_query == 1
    fi = lambda it: it["frame_info"]  # noqa: E731  (gallery row of an item)
    out["tr_feat"], out["tr_label"], out["tr_video"] = feat, lab, vid
    out["tr_q_row"] = np.array([fi(ql[0]) for ql in qlists])
    out["tr_q_list_rows"] = np.array([[fi(x) for x in ql] for ql in qlists])
    out["tr_ap"] = np.array(ev.class_map_evaluation.class_agnostic_ap)
    out["tr_result"] = np.array([res["ap"], res["base_map"], res["novel_map"]])
    out["tr_recall"] = np.array([[res[k][str(n)] for n in (30, 50, 100)]
                                 for k in ("recall", "base_recall", "novel_recall")])
    out["tr_cm_labels_top100"] = np.array(
        [names.index(x) for x in res["cm_dict"]["label"]]).reshape(len(qlists), -1)[:, :100]
    # same data with query_num = 5 (averaged multi-query feature, :1588-1591)
    ev2 = object.__new__(dl.ARV_Retrieval)
    ev2.args, ev2.eval_split = Args(args, query_num=5), "testing"
    ev2.gallery_list, ev2.query_list = items, [it for it in items if it["label"] in possible]
    ev2.faiss_index, ev2.faiss_xb = ev.faiss_index, feat
    res5 = ev2.ranking()
    out["tr5_ap"] = np.array(ev2.class_map_evaluation.class_agnostic_ap)
    out["tr5_result"] = np.array([res5["ap"], res5["base_map"], res5["novel_map"]])

    # ---- ARV_Retrieval_Moment.ranking ------------------------------------------------
    import multiprocessing
    multiprocessing.cpu_count = lambda: 4  # -> 1 forked worker (:1334) => results in job order
    orc.set_num_threads(1)  # libgomp's thread pool does not survive fork(); 1 thread needs none
    V, M = 40, 0
    mom = []
    mv, ms, me, miou, mhit = [], [], [], [], []
    for v in range(V):
        nclip = int(rng.randint(3, 9))
        for cpm in range(1, 4):
            for st in range(0, nclip - cpm + 1):
                hit = None
                if rng.rand() < 0.5:
                    hit = dict(label=names[int(rng.randint(0, 11))], iou=float(rng.rand()))
                mv.append(v), ms.append(st * 5.0), me.append((st + cpm) * 5.0)
                miou.append(hit["iou"] if hit else 0.0)
                mhit.append(names.index(hit["label"]) if hit else -1)
                mom.append(dict(video_id="u_%03d" % v, start_sec=st * 5, end_sec=(st + cpm) * 5,
                                closest_hit=hit))
                M += 1
    mfeat = rng.randn(M, D).astype(np.float32)
    mfeat /= np.linalg.norm(mfeat, axis=1, keepdims=True)
    mfeat = np.ascontiguousarray(mfeat * rng.uniform(0.6, 1.0, (M, 1)), np.float32)
    for i, m_ in enumerate(mom):
        m_["feat"] = mfeat[i].copy()
    NQ = 60
    qlab = rng.randint(0, 11, NQ)
    qlab[:22] = np.repeat(np.arange(11), 2)
    qvid = rng.randint(0, V, NQ)
    qfeat = rng.randn(NQ, D).astype(np.float32)
    qfeat /= np.linalg.norm(qfeat, axis=1, keepdims=True)
    qfeat = np.ascontiguousarray(qfeat, np.float32)
    mq = [dict(label=names[qlab[i]], video_id="u_%03d" % qvid[i], feat=qfeat[i].copy(), qrow=i,
               retrieval_type="base" if names[qlab[i]] in train_l else "novel") for i in range(NQ)]
    evm = object.__new__(dl.ARV_Retrieval_Moment)
    evm.args = args
    evm.gallery_list, evm.query_list = mom, mq
    evm.faiss_index = faiss.IndexFlatL2(D)
    evm.faiss_index.add(mfeat)
    evm.faiss_xb = mfeat
    resm = evm.ranking()["map05"]
    orc.set_num_threads(os.cpu_count())
    out["mo_feat"], out["mo_video"] = mfeat, np.array(mv)
    out["mo_start"], out["mo_end"] = np.array(ms, np.float32), np.array(me, np.float32)
    out["mo_iou"], out["mo_hit"] = np.array(miou, np.float32), np.array(mhit)
    out["mo_qfeat"], out["mo_qlabel"] = qfeat, qlab
    out["mo_q_list_rows"] = np.array([[x["qrow"] for x in ql] for ql in evm.query_list])
    out["mo_qvideo"] = qvid
    out["mo_ap"] = np.array(evm.class_map_evaluation05.class_agnostic_ap)
    out["mo_result"] = np.array([resm["ap"], resm["base_map"], resm["novel_map"]])
    out["mo_recall"] = np.array([[resm[k][str(n)] for n in (30, 50, 100)]
                                 for k in ("recall", "base_recall", "novel_recall")])

    # ---- metric-learning loss ---------------------------------------------------------
    torch.manual_seed(0)
    B, DL = 48, 64
    emb = torch.nn.functional.normalize(torch.randn(B, DL), dim=1)
    labels = np.repeat(np.arange(12), 4)
    samp = mll.Sampler("semihard")
    out["tl_emb"], out["tl_labels"] = emb.numpy().copy(), labels
    out["tl_pdist"] = samp.pdist(emb).numpy()
    np.random.seed(1)
    trip = np.array(samp.give(emb, labels))
    out["tl_semihard_trip"] = trip
    crit = mll.TripletLoss(margin=0.2, sampling_method="semihard")
    crit.sampler.give = lambda b, l: trip.tolist()
    x = emb.clone().requires_grad_(True)
    loss = crit(x, labels)
    loss.backward()
    out["tl_semihard_loss"], out["tl_semihard_grad"] = loss.item(), x.grad.numpy().copy()
    # batch-hard composed from reference pieces (SURVEY.md 8c "Batch-hard oracle")
    d = samp.pdist(emb.detach()).numpy()
    bh = []
    for i in range(B):
        pos = labels == labels[i]
        pos[i] = False
        neg = labels != labels[i]
        pi = np.where(pos)[0][np.argmax(d[i][pos])]
        ni = np.where(neg)[0][np.argmin(d[i][neg])]
        bh.append([i, pi, ni])
    bh = np.array(bh)
    crit.sampler.give = lambda b, l: bh.tolist()
    x = emb.clone().requires_grad_(True)
    loss = crit(x, labels)
    loss.backward()
    out["tl_bh_trip"], out["tl_bh_loss"], out["tl_bh_grad"] = bh, loss.item(), x.grad.numpy().copy()
    out["tl_margin"] = 0.2

    # ---- pre-match pooling ---------------------------------------------------------------
    f = torch.from_numpy(rng.rand(6, 32, 8).astype(np.float32))
    out["pool_in"] = f.numpy().copy()
    out["pool_out"] = np.mean(torch.nn.functional.normalize(f, p=2, dim=1).numpy(), axis=-1)

    np.savez_compressed(os.path.join(HERE, "reference_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "reference_golden.npz"),
          {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    main()
