"""Stateless entry points: pre-match pooling (a1), canonical all-pairs scores, the multi-GPU
top-k merge (8e) and the sharded search built on it."""
import ctypes as C

import numpy as np

from . import _lib as L


def pool_normalize(x):
    """F.normalize(p=2, dim=1) per time step (main.py:232) + temporal mean
    (dataloader_baseline.py:1481): [B,C,T] -> [B,C]."""
    x, _ = L.emb_dtype(x)
    B, Cc, T = x.shape
    out = L.empty_like_kind(x, (B, Cc), np.float32)
    L.check(L.lib().vq_pool_normalize(L.ptr(x), C.c_int64(B), C.c_int(Cc), C.c_int(T), L.ptr(out), L.stream_of(x)))
    return out


def l2sq(q, g):
    q, _ = L.emb_dtype(q)
    g, _ = L.emb_dtype(g)
    out = L.empty_like_kind(q, (q.shape[0], g.shape[0]), np.float32)
    L.check(L.lib().vq_l2sq(L.ptr(q), L.ptr(g), C.c_int64(q.shape[0]), C.c_int64(g.shape[0]), C.c_int(q.shape[1]),
                            L.ptr(out), L.stream_of(q)))
    return out


def topk_merge(d2, idx):
    """[G,nq,k] per-shard lists (global row ids, -1 = empty) -> [nq,k] by (d2 asc, idx asc)."""
    G, nq, k = d2.shape
    od = L.empty_like_kind(d2, (nq, k), np.float32)
    oi = L.empty_like_kind(d2, (nq, k), np.int64)
    L.check(L.lib().vq_topk_merge(L.ptr(d2), L.ptr(idx), C.c_int(G), C.c_int64(nq), C.c_int64(k), L.ptr(od),
                                  L.ptr(oi), L.stream_of(d2)))
    return od, oi


def topk_pack(d2, idx, row_offset=0):
    """Local list -> packed keys int64[...] (bit pattern of u64: ordered d2 << 32 | row_offset + idx; -1 = empty)."""
    keys = L.empty_like_kind(d2, tuple(d2.shape), np.int64)
    L.check(L.lib().vq_topk_pack(L.ptr(d2), L.ptr(idx), C.c_int64(int(np.prod(d2.shape))), C.c_int64(int(row_offset)),
                                 L.ptr(keys), L.stream_of(d2)))
    return keys


def topk_merge_keys(keys):
    """[G,nq,k] packed keys -> ([nq,k] d2, [nq,k] global rows), the answer of ``topk_merge``."""
    G, nq, k = keys.shape
    od = L.empty_like_kind(keys, (nq, k), np.float32)
    oi = L.empty_like_kind(keys, (nq, k), np.int64)
    L.check(L.lib().vq_topk_merge_keys(L.ptr(keys), C.c_int(G), C.c_int64(nq), C.c_int64(k), L.ptr(od), L.ptr(oi),
                                       L.stream_of(keys)))
    return od, oi


def topk_merge_heads(keys, k):
    """[G,nq,m] heads of the per-shard packed lists -> ([nq,k] d2, [nq,k] global rows, incomplete int32[1]); the result
    is final iff ``incomplete`` is 0 (vq_topk_merge_heads)."""
    G, nq, m = keys.shape
    od = L.empty_like_kind(keys, (nq, k), np.float32)
    oi = L.empty_like_kind(keys, (nq, k), np.int64)
    inc = L.empty_like_kind(keys, (1,), np.int32)
    L.check(L.lib().vq_topk_merge_heads(L.ptr(keys), C.c_int(G), C.c_int64(nq), C.c_int64(m), C.c_int64(k), L.ptr(od),
                                        L.ptr(oi), L.ptr(inc), L.stream_of(keys)))
    return od, oi, inc


def head_len(k, ws):
    """Entries of each shard's list sent in the first round: a query's global top-k takes Binomial(k, 1/ws) entries from a
    shard of a randomly ordered gallery -- mean + 6 sigma + 4, rounded up to 8 (k = 100: 40 of 100 at 8 shards, 64 at 4);
    k itself (single round) when that saves less than a third."""
    mu = k / ws
    m = int(-(-(mu + 6.0 * (k * (1.0 / ws) * (1.0 - 1.0 / ws)) ** 0.5 + 4.0) // 8) * 8)
    return k if 3 * m > 2 * k else m


_head_failed = {}   # (k, world size) -> the head round came back incomplete last time: skip it (gallery ordered by class, ...)


def exchange_merge(D, I, row_offset, group=None, events=None, head=None):
    """The exchange step of the row-sharded search on CUDA tensors.  Round 1: pack the first ``m`` entries of the local
    list (LOCAL rows I, this rank's first global row ``row_offset``) into 8-byte keys, ONE all-gather, merge the heads
    (``head_len``: 40 % of the bytes at 8 shards).  Every rank sees the same gathered heads, hence takes the same
    decision: if some query's merge is not provably final (a shard's head ended before the merged k-th key), round 2
    repeats the exchange with the full lists -- and later calls skip round 1.  Same result as
    ``merge_gathered(D, I + row_offset)``.  ``events``: a list that receives four CUDA events (start, packed,
    gathered, merged) for phase timing."""
    import torch
    import torch.distributed as dist
    ws = dist.get_world_size(group)
    nq, k = D.shape

    def mark():
        if events is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            events.append(e)

    m = head_len(k, ws) if head is None else int(head)
    mark()
    if m < k and not _head_failed.get((k, ws), False):
        hk = topk_pack(D[:, :m].contiguous(), I[:, :m].contiguous(), row_offset)
        mark()
        gk = torch.empty((ws * nq, m), dtype=torch.int64, device=D.device)
        dist.all_gather_into_tensor(gk, hk, group=group)
        mark()
        od, oi, inc = topk_merge_heads(gk.view(ws, nq, m), k)
        mark()
        if int(inc.item()) == 0:      # (one 4-byte read-back; identical on every rank)
            return od, oi
        _head_failed[(k, ws)] = True
        if events is not None:
            del events[-4:]
            mark()
    keys = topk_pack(D.contiguous(), I.contiguous(), row_offset)
    mark()
    gk = torch.empty((ws * nq, k), dtype=torch.int64, device=D.device)
    dist.all_gather_into_tensor(gk, keys, group=group)
    mark()
    out = topk_merge_keys(gk.view(ws, nq, k))
    mark()
    return out


def local_k(k, ws):
    """How many entries a rank asks of its own shard: only the head of its list (``head_len``) while the head exchange
    works for this (k, world size) -- the merged top-k takes about k / ws entries from each shard, and a search for 40
    instead of 100 neighbours keeps half as many candidates per query (k' 64 instead of 128: thresholds tighten sooner,
    9 % less time on a 1.25 M-row shard) -- else k."""
    m = head_len(k, ws)
    return m if (m < k and not _head_failed.get((k, ws), False)) else k


def exchange_heads(Dh, Ih, row_offset, k, group=None, events=None):
    """One round on lists that ARE only heads (the local search asked for ``local_k`` entries): pack, ONE all-gather,
    merge.  Returns (d2 [nq,k], global rows [nq,k]) when the merge is provably the global top-k on every query, else
    None -- the caller searches again for k entries and runs ``exchange_merge`` (every rank sees the same gathered
    heads, hence takes the same branch).  ``events``: receives four CUDA events (start, packed, gathered, merged)."""
    import torch
    import torch.distributed as dist
    ws = dist.get_world_size(group)
    nq, m = Dh.shape

    def mark():
        if events is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            events.append(e)

    mark()
    hk = topk_pack(Dh.contiguous(), Ih.contiguous(), row_offset)
    mark()
    gk = torch.empty((ws * nq, m), dtype=torch.int64, device=Dh.device)
    dist.all_gather_into_tensor(gk, hk, group=group)
    mark()
    od, oi, inc = topk_merge_heads(gk.view(ws, nq, m), k)
    mark()
    if int(inc.item()) == 0:      # (one 4-byte read-back; identical on every rank)
        return od, oi
    if events is not None:
        del events[-4:]
    return None


MAX_GLOBAL_ROWS = 1 << 32   # merge keys carry the global row in 32 bits


def check_global_rows(n_global):
    """The exchange (packed or not) merges on 8-byte (d2, global row) keys: galleries of more than 2^32 rows are
    refused instead of silently truncated.  The decision uses the GLOBAL row count, so every rank takes the same
    branch (a per-rank test could issue different collectives on either side of the limit and hang)."""
    if int(n_global) > MAX_GLOBAL_ROWS:
        raise ValueError("row-sharded search supports at most 2^32 gallery rows in total, got %d" % int(n_global))


def sharded_search(index, xq, k, row_offset, group=None, n_global=None, merge=None, events=None):
    """Row-sharded gallery: every rank searches its shard (`index` holds rows
    [row_offset, row_offset + ntotal) of the global gallery), the per-rank lists are exchanged with ONE all-gather of
    packed keys and merged on every rank.  On CUDA a rank first asks its shard only for the HEAD of its list
    (``local_k``: 40 of 100 at 8 ranks); if the merged heads are not provably final for some query (a gallery ordered
    by class, ...) the search is repeated for k entries, and later calls go there directly.
    xq: CUDA tensor (NCCL) or CPU tensor / numpy (gloo tests call `merge_gathered` with an oracle-made local list).
    ``n_global``: total rows over all ranks; when omitted it is all-reduced (max of row_offset + ntotal)."""
    import torch
    import torch.distributed as dist
    if n_global is None:
        t = torch.tensor([int(row_offset) + int(index.ntotal)], dtype=torch.int64,
                         device=xq.device if getattr(xq, "is_cuda", False) else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        n_global = int(t.item())
    check_global_rows(n_global)
    if getattr(xq, "is_cuda", False):
        ws = dist.get_world_size(group)
        m = local_k(k, ws)
        if m < k:
            Dh, Ih = index.search(xq, m)
            res = exchange_heads(Dh, Ih, row_offset, k, group, events)
            if res is not None:
                return res
            _head_failed[(k, ws)] = True
    D, I = index.search(xq, k)
    D = torch.as_tensor(D)
    I = torch.as_tensor(I)
    if D.is_cuda:
        return exchange_merge(D, I, row_offset, group, events)
    I = torch.where(I >= 0, I + int(row_offset), I)
    return merge_gathered(D, I, group, merge)


def merge_gathered(D, I, group=None, merge=None):
    """All-gather [nq,k] lists over the process group and merge them (merge defaults to the CUDA
    kernel; tests on CPU pass the oracle's merge)."""
    import torch
    import torch.distributed as dist
    ws = dist.get_world_size(group)
    nq, k = D.shape
    # rank-major concatenation along dim 0 (the layout both NCCL and gloo accept), viewed [ws, nq, k]
    gD = torch.empty((ws * nq, k), dtype=D.dtype, device=D.device)
    gI = torch.empty((ws * nq, k), dtype=I.dtype, device=I.device)
    dist.all_gather_into_tensor(gD, D.contiguous(), group=group)
    dist.all_gather_into_tensor(gI, I.contiguous(), group=group)
    return (merge or topk_merge)(gD.view(ws, nq, k), gI.view(ws, nq, k))


def window_means(feat, vid_t0, win_start, lo, hi, out=None):
    """every window mean ``np.mean(video_feat[:, lo:hi], axis=1)`` of garner_feat
    (dataloader_baseline.py:1125-1147, :809-834) in one launch.  feat [C, Ttot] (per-frame features of the
    videos concatenated along time; numpy or CUDA tensor), vid_t0 / win_start int64[V+1], lo / hi int32 per
    window relative to the video's first frame.  Returns [n_windows, C] float32 where ``feat`` lives (``out``: a
    buffer of that shape to fill instead of allocating one)."""
    t0 = np.ascontiguousarray(vid_t0, dtype=np.int64)
    ws = np.ascontiguousarray(win_start, dtype=np.int64)
    feat = L.as_f32(feat)
    nch, Ttot = int(feat.shape[0]), int(feat.shape[1])
    lo_, hi_ = L.as_i32(lo), L.as_i32(hi)
    if L._is_torch(feat) and feat.is_cuda and not L._is_torch(lo_):
        import torch
        lo_, hi_ = torch.from_numpy(lo_).to(feat.device), torch.from_numpy(hi_).to(feat.device)
    if out is None:
        out = L.empty_like_kind(feat, (int(ws[-1]), nch), np.float32)
    L.check(L.lib().vq_window_means(L.ptr(feat), C.c_int(nch), C.c_int64(Ttot), L.ptr(t0), L.ptr(ws),
                                    C.c_int64(len(t0) - 1), L.ptr(lo_), L.ptr(hi_), L.ptr(out), L.stream_of(feat)))
    return out


def query_slice(nq, group=None):
    """Queries [q0, q1) of this rank when an evaluation is sharded by query."""
    import torch.distributed as dist
    ws, r = dist.get_world_size(group), dist.get_rank(group)
    return r * nq // ws, (r + 1) * nq // ws


def sharded_eval(eval_fn, nq, group=None, device=None):
    """mAP evaluation over several GPUs.  The evaluators rank the FULL gallery for every query (k = ntotal,
    dataloader_baseline.py:1594), and a trimmed / moment gallery (tens of MB to a few GB) fits every GPU, so the
    evaluation shards by QUERY: rank r runs ``eval_fn(q0, q1) -> (ap[q1-q0], recall[q1-q0,3])`` on its slice
    against its full local index and ONE all-gather of the per-query [ap, R@30, R@50, R@100] rows (32 bytes per
    query) puts the complete vectors on every rank.  No score ever crosses NVLink."""
    import torch
    import torch.distributed as dist
    ws = dist.get_world_size(group)
    q0, q1 = query_slice(nq, group)
    ap, rec = eval_fn(q0, q1)
    ap, rec = torch.as_tensor(ap), torch.as_tensor(rec)
    dev = device if device is not None else ap.device
    qmax = (nq + ws - 1) // ws + 1
    mine = torch.zeros((qmax, 4), dtype=torch.float64, device=dev)
    mine[:q1 - q0, 0] = ap.to(dev)
    mine[:q1 - q0, 1:] = rec.to(dev)
    allr = torch.empty((ws * qmax, 4), dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(allr, mine, group=group)
    allr = allr.view(ws, qmax, 4)
    parts = [allr[r, :(r + 1) * nq // ws - r * nq // ws] for r in range(ws)]
    out = torch.cat(parts, 0)
    return out[:, 0].contiguous(), out[:, 1:].contiguous()
