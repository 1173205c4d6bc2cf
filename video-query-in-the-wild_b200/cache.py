"""columnar feature cache.

The reference caches extracted features as a pickle of two lists of dicts, one ``np.float32[D]`` per item
(``feat_cache.pkl`` = ``dict(query_list=..., gallery_list=...)``, dataloader_baseline.py:1447-1500; same scheme
at :699-707, :846-854, :1004-1012, :1176-1184), and rebuilds the gallery matrix with a Python loop + concat on
every evaluation (:1503-1506).  Here the same content is a handful of arrays -- ``feat [N, D]`` plus integer id
columns with their name tables -- in one ``.npz``: the matrix goes to ``IndexFlatL2.add`` as it is, a gallery
shard for one GPU is a slice, and ``to_items`` gives back the reference's list-of-dicts (rows are views).

    convert_feat_cache("feat_cache.pkl", "feat_cache.npz")           # one-shot converter
    query_list, gallery_list = load_feat_cache("feat_cache.npz")      # the reference's hand-off lists
    cols = load_columns("feat_cache.npz", "gallery")                   # arrays, e.g. for sharding
"""
import json
import pickle

import numpy as np

_ID_KEYS = ("label", "video_id", "retrieval_type")          # strings (or any hashable) -> int32 id + name table
_NUM_KEYS = ("is_query", "start_sec", "end_sec", "duration_sec", "activitynet_duration")


def _table(values):
    names, ids, out = [], {}, np.empty(len(values), np.int32)
    for i, v in enumerate(values):
        k = json.dumps(v, sort_keys=True, default=str)
        if k not in ids:
            ids[k] = len(names)
            names.append(k)
        out[i] = ids[k]
    return out, np.array(names)


def to_columns(items):
    """list of item dicts -> dict of arrays.  ``feat`` rows are stacked; the id keys become int32 columns +
    ``<key>__names`` (JSON-encoded originals, so ints / tuples / None survive); ``closest_hit`` (moments,
    :1144) becomes ``hit_label`` (-1 = None) + ``hit_iou``; every other key goes to ``extra`` as JSON per item."""
    n = len(items)
    cols = {"feat": np.ascontiguousarray(np.stack([np.asarray(it["feat"], np.float32).reshape(-1) for it in items]))
            if n else np.zeros((0, 0), np.float32)}
    keys = set().union(*[it.keys() for it in items]) if n else set()
    for k in _ID_KEYS:
        if k in keys:
            cols[k], cols[k + "__names"] = _table([it.get(k) for it in items])
    for k in _NUM_KEYS:
        if k in keys:
            cols[k] = np.array([it.get(k, np.nan) for it in items], np.float64)
    if "closest_hit" in keys:
        lab, cols["hit_label__names"] = _table([(it["closest_hit"] or {}).get("label") for it in items])
        none_id = {nm: i for i, nm in enumerate(cols["hit_label__names"].tolist())}.get("null", -2)
        cols["hit_label"] = np.where(lab == none_id, -1, lab).astype(np.int32)
        cols["hit_iou"] = np.array([(it["closest_hit"] or {}).get("iou", 0.0) for it in items], np.float64)
        cols["hit_extra"] = np.array([json.dumps({k: v for k, v in (it["closest_hit"] or {}).items() if k not in ("label", "iou")},
                                                 default=_np_default) for it in items])
    rest = keys - set(_ID_KEYS) - set(_NUM_KEYS) - {"feat", "closest_hit"}
    if rest:
        cols["extra"] = np.array([json.dumps({k: it[k] for k in sorted(rest) if k in it}, default=_np_default) for it in items])
    return cols


def _np_default(o):
    if isinstance(o, np.ndarray):
        return o.tolist()
    if isinstance(o, np.generic):
        return o.item()
    return str(o)


def to_items(cols):
    """Inverse of ``to_columns``: the reference's list of dicts (``feat`` entries are row views of the matrix)."""
    n = cols["feat"].shape[0]
    names = {k: [json.loads(s) for s in cols[k + "__names"].tolist()] for k in _ID_KEYS + ("hit_label",) if k + "__names" in cols}
    items = []
    for i in range(n):
        it = {"feat": cols["feat"][i]}
        for k in _ID_KEYS:
            if k in cols:
                it[k] = names[k][int(cols[k][i])]
        for k in _NUM_KEYS:
            if k in cols and not np.isnan(cols[k][i]):
                v = float(cols[k][i])
                it[k] = int(v) if v.is_integer() else v
        if "hit_label" in cols:
            h = int(cols["hit_label"][i])
            if h < 0:
                it["closest_hit"] = None
            else:
                it["closest_hit"] = dict(label=names["hit_label"][h], iou=float(cols["hit_iou"][i]),
                                         **json.loads(str(cols["hit_extra"][i])))
        if "extra" in cols:
            it.update(json.loads(str(cols["extra"][i])))
        items.append(it)
    return items


def save_feat_cache(path, query_list, gallery_list):
    """Write both hand-off lists.  Queries that ARE gallery items (the trimmed path: the query list is a filter of
    the gallery list, :1486-1491) are stored as row numbers, not copies."""
    pos = {id(g): i for i, g in enumerate(gallery_list)}
    rows = [pos.get(id(q), -1) for q in query_list]
    out = {"gallery__" + k: v for k, v in to_columns(gallery_list).items()}
    if query_list and all(r >= 0 for r in rows):
        out["query_rows"] = np.array(rows, np.int64)
    else:
        out.update({"query__" + k: v for k, v in to_columns(query_list).items()})
    np.savez(path, **out)


def load_columns(path, which="gallery"):
    """The arrays of one list (``which`` = "gallery" or "query") without building any dict."""
    z = np.load(path, allow_pickle=False)
    pre = which + "__"
    if which == "query" and "query_rows" in z.files:
        g = load_columns(path, "gallery")
        rows = z["query_rows"]
        return {k: (v if k.endswith("__names") else v[rows]) for k, v in g.items()}
    return {k[len(pre):]: z[k] for k in z.files if k.startswith(pre)}


def load_feat_cache(path):
    """-> (query_list, gallery_list) as the reference's extract_item_feature leaves them (:1450-1456)."""
    z = np.load(path, allow_pickle=False)
    gallery = to_items(load_columns(path, "gallery"))
    if "query_rows" in z.files:
        return [gallery[int(r)] for r in z["query_rows"]], gallery
    return to_items(load_columns(path, "query")), gallery


def convert_feat_cache(pkl_path, npz_path):
    """One-shot converter from the reference's ``feat_cache.pkl`` (:1447-1500).  Unpickle only files you trust."""
    with open(pkl_path, "rb") as f:
        obj = pickle.load(f)
    save_feat_cache(npz_path, obj["query_list"], obj["gallery_list"])
    return npz_path
