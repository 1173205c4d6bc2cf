"""Mint the REAL-METADATA goldens (SURVEY.md 8c case 1) by running the reference's own evaluators on its own
metadata files with seeded synthetic features (authoring container only: needs /root/reference).

    python tests/golden/make_golden_real.py [trimmed] [clip] [moment]   # -> reference_golden_real.npz

  trimmed  ARV_Retrieval.ranking() over data_generate/arv_db_100_20_80.json["testing"]: 7 670 gallery items,
           6 338 queries (is_query == 1), 512-d features; the real retrieval_type / is_query / shared-video_id
           distribution.  ~10 minutes of the reference's Python loop.
  clip     ARV_Retrieval_Clip: the reference's OWN extract_item_feature (garner_feat :790-834, cal_label
           :795-807) driven through its --memory_leak_debug path on a subsample of
           arv_db_100_20_80_untrimmed.json's gallery videos, then ranking() :883-955.
  moment   ARV_Retrieval_Moment: the reference's OWN garner_feat (:1091-1151: moment grid, closest_hit in
           float64) on a subsample that contains the longest video (the 3 5xx-moment one), then ranking()
           :1206-1371 through a forked worker (NMS, grouped R@N).
Features are regenerated from seeds by the tests (tests/golden/synth.py), so the fixture holds metadata and the
reference's outputs only.  Shims: termcolor, faiss (oracle-backed, as make_golden.py), the video loader
(LongVideoDataset / DataLoader -> a list of (None, meta) batches; no frames are read in --memory_leak_debug).
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402  (install_shims, Args; also puts ROOT and REF on sys.path)
from synth import class_feats  # noqa: E402

orc = mg.orc
OUT = os.path.join(HERE, "reference_golden_real.npz")


def ids_of(values, table):
    return np.array([table.setdefault(v, len(table)) for v in values], np.int32)


def trimmed(dl, out):
    import json
    args = mg.Args(meta_split="100_20_80", eval_split="testing", query_num=1, debug=False, memory_leak_debug=True,
                   metric_feat_dim=512)
    cfg = dl.dataset_config["100_20_80"]
    data = json.load(open(cfg["json_path"]))
    items = []
    for _, v in data["testing"].items():                      # :1444-1445
        items.extend(v)
    ltab, vtab = {}, {}
    lab = ids_of([it["label"] for it in items], ltab)
    vid = ids_of([it["video_id"] for it in items], vtab)
    feat = class_feats(11, lab, 512)
    for i, it in enumerate(items):
        it["feat"], it["frame_info"] = feat[i].copy(), i
    possible = cfg["arv_train_label"] + cfg["arv_test_label"]
    ev = object.__new__(dl.ARV_Retrieval)
    ev.args, ev.eval_split = args, "testing"
    ev.gallery_list = items
    ev.query_list = [it for it in items if it["label"] in possible]      # :1486-1489
    import faiss
    ev.faiss_index = faiss.IndexFlatL2(512)
    ev.faiss_index.add(feat)
    ev.faiss_xb = feat
    res = ev.ranking()
    ql = ev.query_list
    names = [None] * len(ltab)
    for k, v in ltab.items():
        names[v] = k
    rt = {"base": 0, "novel": 1, "noise": 2}
    out["rt_label_names"] = np.array(names)
    out["rt_label"], out["rt_video"] = lab.astype(np.int16), vid
    out["rt_is_query"] = np.array([it["is_query"] for it in items], np.int8)
    out["rt_rtype"] = np.array([rt[it["retrieval_type"]] for it in items], np.int8)
    out["rt_in_query_list"] = np.array([it["label"] in possible for it in items], np.int8)
    out["rt_q_list_rows"] = np.array([[x["frame_info"] for x in q] for q in ql], np.int16)
    out["rt_ap"] = np.array(ev.class_map_evaluation.class_agnostic_ap)
    out["rt_result"] = np.array([res["ap"], res["base_map"], res["novel_map"]])
    out["rt_recall"] = np.array([[res[k][str(n)] for n in (30, 50, 100)]
                                 for k in ("recall", "base_recall", "novel_recall")])
    print("trimmed real:", len(items), "items", len(ql), "queries", out["rt_result"])


def untrimmed_subsample(dl):
    """Queries: 16 classes x 4 queries from distinct videos; gallery videos: the queries' own long videos (ignore rule
    hits), the longest video of the file, and a seeded random remainder."""
    import json
    cfg = dl.dataset_config["100_20_80"]
    data = json.load(open(cfg["moment_eval_json_path"]))
    rng = np.random.RandomState(5)
    by_cls = {}
    for q in data["query"]:
        if q["retrieval_type"] != dl.RETRIEVAL_TYPE_NOISE:
            by_cls.setdefault(q["label"], []).append(q)
    classes = sorted(c for c, v in by_cls.items() if len({x["video_id"] for x in v}) >= 6)
    rng.shuffle(classes)
    queries = []
    for c in classes[:16]:
        seen = set()
        for q in by_cls[c]:
            if q["video_id"] not in seen and len(seen) < 4:
                seen.add(q["video_id"])
                queries.append(dict(q))
    gal = data["gallery"]
    pos = {g["video_id"]: i for i, g in enumerate(gal)}
    pick = [pos[q["video_id"]] for q in queries[:40] if q["video_id"] in pos]
    pick.append(int(np.argmax([g["activitynet_duration"] for g in gal])))
    rest = rng.permutation(len(gal))[:24].tolist()
    order = []
    for i in pick + rest:
        if i not in order:
            order.append(i)
    return queries, [dict(gal[i]) for i in order]


class StubLoader(list):
    pass


def run_extract(dl, ev, T_frames):
    """The reference's own extract_item_feature with a loader that yields only (None, meta): in --memory_leak_debug
    the features are np.random.rand and no frame is read (:1068-1073, :763-768)."""
    import torch

    def fake_dataset(gallery_list, input_size, test_frame_num):
        return gallery_list

    def fake_loader(gallery_list, batch_size, **kw):
        pairs = []
        for i, g in enumerate(gallery_list):
            nseg = int(np.ceil(g["activitynet_duration"] * dl.fps / T_frames)) + 1   # enough frames for every window
            pairs.extend((i, s) for s in range(nseg))
        batches = StubLoader()
        for b in range(0, len(pairs), batch_size):
            chunk = pairs[b:b + batch_size]
            batches.append((None, dict(long_video_id=torch.tensor([c[0] for c in chunk]),
                                       seg_id=torch.tensor([c[1] for c in chunk]))))
        return batches

    saved = dl.LongVideoDataset, torch.utils.data.DataLoader
    dl.LongVideoDataset, torch.utils.data.DataLoader = fake_dataset, fake_loader
    dl.logger.LOG_DIR = tempfile.mkdtemp(prefix="vq_golden_")
    try:
        ev.extract_item_feature()
    except AssertionError as e:   # the moment evaluator's class-completeness check (:1186-1194) on a SUBSAMPLE
        assert "doesn't exist in gallery" in str(e), e
        print("  (class-completeness assert on the subsample, expected:", str(e)[:60], ")")
    finally:
        dl.LongVideoDataset, torch.utils.data.DataLoader = saved


def base_args(**kw):
    a = dict(meta_split="100_20_80", eval_split="testing", query_num=1, debug=False, memory_leak_debug=True,
             metric_feat_dim=64, temporal_stride=1, test_batch_size=8, test_frame_num=16, input_size=112, workers=0,
             read_cache_feat=False, clip_sec=6)
    a.update(kw)
    return mg.Args(a)


def clip(dl, out):
    import faiss
    np.random.seed(3)
    queries, gallery = untrimmed_subsample(dl)
    gallery = gallery[:40]
    ev = object.__new__(dl.ARV_Retrieval_Clip)
    ev.args = base_args()
    for k in ("temporal_stride", "test_batch_size", "test_frame_num", "input_size", "clip_sec"):
        setattr(ev, k, ev.args[k])
    ev.feat_dim = 64
    cfg = dl.dataset_config["100_20_80"]
    ev.possible_classes = cfg["arv_train_label"] + cfg["arv_test_label"]
    ev.query_list, ev.gallery_list = queries, gallery
    run_extract(dl, ev, 16)
    for i, q in enumerate(ev.query_list):
        q["qrow"] = i
    clips = ev.gallery_list
    ltab, vtab = {"unknown": 0}, {}
    clab = ids_of([c["clip_label"] for c in clips], ltab)
    cvid = ids_of([c["video_id"] for c in clips], vtab)
    qlab = ids_of([q["label"] for q in ev.query_list], ltab)
    qvid = ids_of([q["video_id"] for q in ev.query_list], vtab)
    cfeat, qfeat = class_feats(21, clab, 64), class_feats(22, qlab, 64)
    for i, c in enumerate(clips):
        c["feat"] = cfeat[i].copy()
    for i, q in enumerate(ev.query_list):
        q["feat"] = qfeat[i].copy()
    ev.faiss_index = faiss.IndexFlatL2(64)
    ev.faiss_index.add(cfeat)
    ev.faiss_xb = cfeat
    res = ev.ranking()
    rt = {"base": 0, "novel": 1}
    out["cl_label"], out["cl_video"] = clab, cvid
    out["cl_loc_sec"] = np.array([c["loc_sec"] for c in clips], np.int32)
    out["cl_vid_duration"] = np.array([c["activitynet_duration"] for c in clips], np.float64)
    out["cl_qlabel"], out["cl_qvideo"] = qlab, qvid
    out["cl_qrtype"] = np.array([rt[q[0]["retrieval_type"]] for q in ev.query_list], np.int8)
    out["cl_q_list_rows"] = np.array([[x["qrow"] for x in q] for q in ev.query_list], np.int32)
    out["cl_ap"] = np.array(ev.class_map_evaluation.class_agnostic_ap)
    out["cl_result"] = np.array([res["ap"], res["base_map"], res["novel_map"]])
    out["cl_recall"] = np.array([[res[k][str(n)] for n in (30, 50, 100)]
                                 for k in ("recall", "base_recall", "novel_recall")])
    # the annotations, so that the test can re-derive clip_label with the mirror's cal_label
    ann_v, ann_s, ann_l = [], [], []
    for g in ev.old_gallery:
        for a in g["annotations"]:
            ann_v.append(vtab[g["video_id"]]), ann_s.append(a["segment"])
            ann_l.append(ltab.get(a["label"], -1) if a["label"] in ev.possible_classes else -1)
    out["cl_ann_video"], out["cl_ann_seg"], out["cl_ann_label"] = np.array(ann_v), np.array(ann_s, np.float64), np.array(ann_l)
    print("clip:", len(clips), "clips", len(ev.query_list), "queries", out["cl_result"])


def moment(dl, out):
    import faiss
    import multiprocessing
    multiprocessing.cpu_count = lambda: 4     # -> 1 forked worker (:1334): results arrive in job order
    orc.set_num_threads(1)                    # libgomp's pool does not survive fork()
    np.random.seed(4)
    queries, gallery = untrimmed_subsample(dl)
    ev = object.__new__(dl.ARV_Retrieval_Moment)
    ev.args = base_args()
    for k in ("temporal_stride", "test_batch_size", "test_frame_num", "input_size"):
        setattr(ev, k, ev.args[k])
    ev.feat_dim = 64
    cfg = dl.dataset_config["100_20_80"]
    ev.possible_classes = cfg["arv_train_label"] + cfg["arv_test_label"]
    ev.query_list, ev.gallery_list = queries, gallery
    run_extract(dl, ev, 16)
    for i, q in enumerate(ev.query_list):
        q["qrow"] = i
    mom = ev.gallery_list
    ltab, vtab = {}, {}
    hit = np.array([ltab.setdefault(m["closest_hit"]["label"], len(ltab)) if m["closest_hit"] else -1 for m in mom], np.int32)
    iou = np.array([m["closest_hit"]["iou"] if m["closest_hit"] else 0.0 for m in mom], np.float64)
    mvid = ids_of([m["video_id"] for m in mom], vtab)
    qlab = ids_of([q["label"] for q in ev.query_list], ltab)
    qvid = ids_of([q["video_id"] for q in ev.query_list], vtab)
    mfeat, qfeat = class_feats(31, hit + 1, 64, spread=0.6), class_feats(32, qlab + 1, 64, spread=0.6)
    for i, m in enumerate(mom):
        m["feat"] = mfeat[i].copy()
    for i, q in enumerate(ev.query_list):
        q["feat"] = qfeat[i].copy()
    ev.faiss_index = faiss.IndexFlatL2(64)
    ev.faiss_index.add(mfeat)
    ev.faiss_xb = mfeat
    res = ev.ranking()["map05"]
    orc.set_num_threads(os.cpu_count())
    rt = {"base": 0, "novel": 1}
    out["mr_video"], out["mr_hit"], out["mr_iou"] = mvid, hit, iou
    out["mr_start"] = np.array([m["start_sec"] for m in mom], np.int32)
    out["mr_end"] = np.array([m["end_sec"] for m in mom], np.int32)
    out["mr_qlabel"], out["mr_qvideo"] = qlab, qvid
    out["mr_qrtype"] = np.array([rt[q[0]["retrieval_type"]] for q in ev.query_list], np.int8)
    out["mr_q_list_rows"] = np.array([[x["qrow"] for x in q] for q in ev.query_list], np.int32)
    out["mr_ap"] = np.array(ev.class_map_evaluation05.class_agnostic_ap)
    out["mr_result"] = np.array([res["ap"], res["base_map"], res["novel_map"]])
    out["mr_recall"] = np.array([[res[k][str(n)] for n in (30, 50, 100)]
                                 for k in ("recall", "base_recall", "novel_recall")])
    # the videos behind the grid, so that the test can re-derive it with the mirror's moment_grid / closest_hits
    out["mr_vid_duration"] = np.array([g["activitynet_duration"] for g in ev.old_gallery], np.float64)
    out["mr_vid_T"] = np.array([g["feat"].shape[1] for g in ev.old_gallery], np.int64)
    ann_v, ann_s, ann_l = [], [], []
    for vi, g in enumerate(ev.old_gallery):
        for a in g["annotations"]:
            ann_v.append(vi), ann_s.append(a["segment"])
            ann_l.append(ltab.setdefault(a["label"], len(ltab)) if a["label"] in ev.possible_classes else -1)
    out["mr_ann_video"], out["mr_ann_seg"], out["mr_ann_label"] = np.array(ann_v), np.array(ann_s, np.float64), np.array(ann_l)
    cnt = np.bincount(mvid)
    print("moment real:", len(mom), "moments", len(ev.old_gallery), "videos, max per video", cnt.max(), len(ev.query_list),
          "queries", out["mr_result"])


def main():
    which = sys.argv[1:] or ["trimmed", "clip", "moment"]
    mg.install_shims()
    os.chdir(mg.REF)
    import dataloader_baseline as dl
    out = dict(np.load(OUT)) if os.path.exists(OUT) else {}
    for w in which:
        {"trimmed": trimmed, "clip": clip, "moment": moment}[w](dl, out)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
