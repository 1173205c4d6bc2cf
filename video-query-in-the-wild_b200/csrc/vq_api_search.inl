// vq_api_search.inl -- search / evaluation orchestration (included by vq_api.cu).
#include <stdlib.h>

namespace vq {



static int resolve_prec(const vq_index *ix) {
    if (ix->prec != VQ_PREC_AUTO) return ix->prec;
    return ix->dtype == VQ_F32 ? VQ_PREC_SPLIT3 : VQ_PREC_BF16;
}

static int choose_kprime(int k, int prec, int dtype) {
    int base = k + std::max(16, k / 4);
    if (dtype == VQ_F32 && prec == VQ_PREC_BF16) base = k * 5 / 2 + 32;  // single bf16 pass on fp32 data: wide band
    base = (base + 31) / 32 * 32;
    return std::min(kMaxKPrime, std::max(64, base));
}

// worst-case |coarse - (|q|^2 - d2_canonical)| <= eps_rel*|q|*|g|max + eps_abs*(|q|^2+|g|max^2)
static void error_model(int prec, int dtype, int dp, float *eps_rel, float *eps_abs) {
    const float u = 5.9604645e-8f;                          // 2^-24
    const int npass = prec == VQ_PREC_SPLIT3 ? 3 : 1;
    float rel = 2.0f * npass * 2.0f * dp * u;               // fp32 accumulation inside the tensor core, x2 (score = 2*dot)
    if (dtype == VQ_F32) rel += prec == VQ_PREC_SPLIT3 ? 2.4e-5f : 8.0e-3f;  // operand rounding to bf16 (split / single)
    *eps_rel = rel;
    *eps_abs = 6.0f * dp * u;                               // |g|^2 reduction and its 3-term bf16 split + the canonical chain's own rounding
    if (g_hooks.force_fallback) *eps_abs = 1e30f;  // test hook: no query can be proven
}

// Work plan (vq_kernels.h, GemmTopkParams).  The gallery is cut into `n_chunks` row ranges so that
// (query groups x chunks) fills the persistent CTAs evenly; segment (chunk, group) goes to CTA
// (chunk * n_groups + group) mod n_cta, i.e. the CTAs sweep a chunk TOGETHER (every gallery tile is
// fetched from HBM once per wave and shared through L2).  Segment (chunk, g) names (chunk-1, g) as
// its predecessor: its threshold starts from the predecessor's final one when that is already
// known, so only the first chunk of a group pays the threshold warm-up.
struct SegPlan {
    std::vector<int64_t> rows;  // seg_begin[nseg] | seg_end[nseg]
    std::vector<int32_t> ints;  // seg_group[nseg] | cta_seg_start[n_cta+1] | seg_pred[nseg] | group_segs[n_groups*n_chunks]
    int nseg = 0, n_cta = 0, n_chunks = 0;
};

static void plan_segments(const vq_index *ix, int64_t n_groups, int slots, int kprime, bool segmax, bool split, SegPlan *pl) {
    const int64_t n = ix->n;
    const int64_t ntiles = (n + kTileN - 1) / kTileN;
    int64_t max_chunks = std::min<int64_t>(ntiles, (int64_t)(128 * 64) / kprime);  // select's gather capacity
    max_chunks = std::min<int64_t>(max_chunks, std::max<int64_t>(1, ntiles / 8));   // >= 8 tiles per chunk
    max_chunks = std::max<int64_t>(1, std::min<int64_t>(max_chunks, 64));
    // Few query groups (fewer than half the CTA slots): "walkers".  The lock-step plan below would put a segment's
    // predecessor (same group, previous chunk) into the SAME round on another CTA -- running concurrently, its
    // threshold is not published yet, and nearly every segment would warm up from an empty list (configs[2]: 16 pair
    // groups on 74 CTA pairs, 460 of 592 segments cold).  Instead the gallery is cut into W = slots / n_groups parts
    // and CTA (w, g) walks part w for group g as ONE long segment: W cold starts per group, each amortised over
    // 1/W of the gallery, and the CTAs of one part still sweep the same rows together (L2 sharing).
    const bool walkers = n_groups * 2 <= slots && !g_hooks.no_chain && ntiles >= 8 * (slots / n_groups);
    int64_t best = 1;
    double best_eff = -1.0;
    if (walkers) best = std::min<int64_t>(slots / n_groups, max_chunks);
    if (walkers && split) {
        // Three MMA passes per tile: a part's cold start weighs a third as much against its sweep, and filling the
        // last round of CTAs pays -- configs[2] f32: 32 groups x 4 parts on 148 CTAs (86 %) 13.65 ms, x 9 parts (two
        // full rounds, 97 %) 12.29 ms; with one pass (CTA pairs) the extra cold starts cancel the gain (6.93 / 6.91 ms).
        const int64_t w0 = slots / n_groups;
        double e_best = -1.0;
        // (only where a part still sweeps >= 64 tiles: on a 20 480-row gallery nine parts of nine tiles each lose to
        // four of twenty, 1.51 against 1.12 ms, scripts/ab_walkers_small.py)
        for (int64_t w = w0; w <= std::min<int64_t>(max_chunks, 3 * w0) && (w == w0 || ntiles >= 64 * w); ++w) {
            const int64_t units = n_groups * w;
            const double eff = (double)units / (double)(((units + slots - 1) / slots) * slots);
            if (eff > e_best + 0.02) {
                e_best = eff;
                best = w;
            }
        }
    }
    if (walkers && g_hooks.walkers > 0) best = std::min<int64_t>(g_hooks.walkers, max_chunks);  // A/B hook: parts per group
    for (int64_t c = 1; c <= max_chunks && !walkers; ++c) {
        const int64_t units = n_groups * c;
        const double eff = (double)units / (double)(((units + slots - 1) / slots) * slots);
        if (eff > best_eff + 0.01) {
            best_eff = eff;
            best = c;
        }
    }
    std::vector<int64_t> cs{0};
    for (int64_t c = 1; c < best; ++c) {
        int64_t target = (ntiles * c / best) * kTileN;
        if (segmax) {  // a video's moments stay in one chunk
            auto it = std::lower_bound(ix->seg_start.begin(), ix->seg_start.end(), target);
            target = it == ix->seg_start.end() ? n : *it;
        }
        if (target > cs.back() && target < n) cs.push_back(target);
    }
    cs.push_back(n);
    const int64_t n_chunks = (int64_t)cs.size() - 1;
    const int64_t units = n_groups * n_chunks;
    const int64_t P = std::max<int64_t>(1, std::min<int64_t>(slots, units));
    // segment ids in CTA-major order
    std::vector<int32_t> s_of_u((size_t)units), cta((size_t)P + 1);
    std::vector<int64_t> sb((size_t)units), se((size_t)units);
    std::vector<int32_t> sg((size_t)units), pred((size_t)units), gsegs((size_t)units);
    int32_t sid = 0;
    for (int64_t c = 0; c < P; ++c) {
        cta[(size_t)c] = sid;
        for (int64_t u = c; u < units; u += P) s_of_u[(size_t)u] = sid++;
    }
    cta[(size_t)P] = sid;
    for (int64_t u = 0; u < units; ++u) {
        const int64_t chunk = u / n_groups, g = u % n_groups;
        const int32_t s_ = s_of_u[(size_t)u];
        sb[(size_t)s_] = cs[(size_t)chunk];
        se[(size_t)s_] = cs[(size_t)chunk + 1];
        sg[(size_t)s_] = (int32_t)g;
        // -1: first chunk of the gallery; -2: no predecessor to continue, but earlier chunks exist (walker parts)
        pred[(size_t)s_] = chunk == 0 ? -1 : (walkers || g_hooks.no_chain) ? -2 : s_of_u[(size_t)(u - n_groups)];
        gsegs[(size_t)(g * n_chunks + chunk)] = s_;
    }
    pl->nseg = (int)units;
    pl->n_cta = (int)P;
    pl->n_chunks = (int)n_chunks;
    pl->rows = sb;
    pl->rows.insert(pl->rows.end(), se.begin(), se.end());
    pl->ints = sg;
    pl->ints.insert(pl->ints.end(), cta.begin(), cta.end());
    pl->ints.insert(pl->ints.end(), pred.begin(), pred.end());
    pl->ints.insert(pl->ints.end(), gsegs.begin(), gsegs.end());
}

// Search workspace of an index (candidate lists, plan, ...): owned by the index, grown on demand on the CALLER's
// stream from the library's pool -- stream-ordered (cudaFreeAsync of the old block, cudaMallocFromPoolAsync of the new
// one), so growing never synchronises the device.  vq_index_reserve() sizes it ahead of time.
template <typename T>
static int ensure(T **p, size_t *cap, size_t need, cudaStream_t st) {
    if (*cap >= need) return VQ_OK;
    if (*p) {
        VQ_CUDA(cudaFreeAsync(*p, st));
        *p = nullptr;
        *cap = 0;
    }
    cudaMemPool_t pool = nullptr;
    VQ_TRY(scratch_pool(&pool));
    if (cudaMallocFromPoolAsync((void **)p, need * sizeof(T), pool, st) != cudaSuccess) {
        cudaGetLastError();
        set_error("out of device memory for a %zu-byte workspace", need * sizeof(T));
        return VQ_ERR_NOMEM;
    }
    *cap = need;
    return VQ_OK;
}

// Host-side planning of one tensor-core search: everything search_core needs to size its buffers.
struct SearchPlan {
    int prec = 0, kprime = 0, n_qblocks = 0, group_size = 1, n_groups = 0, slots = 0;
    bool pair = false, split = false;
    SegPlan pl;
    size_t units = 0;
    // bytes: persistent workspace of the index | stream-ordered temporaries of the call
    size_t persistent_bytes() const {
        return pl.rows.size() * 8 + pl.ints.size() * 4 + units * kTileM * kCandCap * 8 + units * kTileM * 4 * 3;
    }
};

static void make_search_plan(const vq_index *ix, int64_t nq, int dtype, int k, bool segmax, SearchPlan *sp) {
    sp->prec = resolve_prec(ix);
    sp->kprime = choose_kprime(k, sp->prec, dtype);
    if (g_hooks.kprime >= k && g_hooks.kprime <= kMaxKPrime) sp->kprime = (g_hooks.kprime + 7) / 8 * 8;  // A/B hook
    sp->split = sp->prec == VQ_PREC_SPLIT3;
    // bf16 single pass with dim <= 512 runs on CTA pairs (resident queries, split gallery tiles)
    sp->pair = !sp->split && ix->dp <= 512 && !g_hooks.no_pair;
    sp->n_qblocks = (int)((nq + kTileM - 1) / kTileM);
    if (sp->pair) sp->n_qblocks = (sp->n_qblocks + 1) / 2 * 2;  // the pad block is all out of range
    sp->group_size = sp->pair ? 2 : 1;
    sp->n_groups = sp->n_qblocks / sp->group_size;
    sp->slots = sp->pair ? gemm_topk_pair_slots() : sm_count();
    plan_segments(ix, sp->n_groups, sp->slots, sp->kprime, segmax, sp->split, &sp->pl);
    sp->units = (size_t)sp->pl.nseg * sp->group_size;
}

static int ensure_workspace(vq_index *ix, const SearchPlan &sp, int64_t nq, cudaStream_t st) {
    VQ_TRY(ensure(&ix->plan_rows_dev, &ix->plan_rows_cap, sp.pl.rows.size(), st));
    VQ_TRY(ensure(&ix->plan_ints_dev, &ix->plan_ints_cap, sp.pl.ints.size(), st));
    VQ_TRY(ensure(&ix->cand, &ix->cand_bytes, sp.units * kTileM * kCandCap, st));
    VQ_TRY(ensure(&ix->cand_cnt, &ix->cand_cnt_cap, sp.units * kTileM, st));
    VQ_TRY(ensure(&ix->cand_tau, &ix->cand_tau_cap, sp.units * kTileM, st));
    VQ_TRY(ensure(&ix->cand_cum, &ix->cand_cum_cap, sp.units * kTileM, st));
    VQ_TRY(ensure(&ix->unproven, &ix->unproven_cap, (size_t)nq, st));
    return VQ_OK;
}

// All pointers are device pointers here.
static int search_core(vq_index *ix, const void *xq, int64_t nq, int dtype, int k, bool segmax, const int32_t *ign,
                       int n_ign, float *D, int64_t *I, int64_t *seg_out, cudaStream_t st) {
    const int prec = resolve_prec(ix);
    int launches = 0;
    VQ_CUDA(cudaMemsetAsync(ix->unproven_count, 0, 4, st));
    if (prec == VQ_PREC_EXACT) {
        FallbackParams f{};
        f.nq = nq; f.n = ix->n; f.d = ix->d; f.dtype = dtype; f.q = xq; f.g = ix->raw;
        f.col_video = ix->video; f.q_ignore = ign; f.n_ign = n_ign;
        f.seg_of_row = segmax ? ix->seg_of_row : nullptr;
        f.seg_start = segmax ? ix->seg_start_dev : nullptr;
        f.nseg = segmax ? (int64_t)ix->seg_start.size() - 1 : 0;
        f.k = k; f.out_d2 = D; f.out_idx = I; f.out_seg = seg_out;
        VQ_TRY(launch_exact_topk(f, (int)std::min<int64_t>(nq, 4 * sm_count()), st));
        ix->stats[1] = 0; ix->stats[2] = 0; ix->stats[3] = 0; ix->stats[4] = 0; ix->stats[5] = 1;
        return VQ_OK;
    }
    SearchPlan spn;
    make_search_plan(ix, nq, dtype, k, segmax, &spn);
    const int kprime = spn.kprime;
    VQ_CHECK(k <= kprime, VQ_ERR_CAPACITY, "k=%d exceeds the streaming top-k limit %d", k, kMaxKPrime);
    const bool pair = spn.pair, split = spn.split;
    const SegPlan &pl = spn.pl;
    const int n_qblocks = spn.n_qblocks, group_size = spn.group_size;
    const size_t units = spn.units;
    // query operand planes
    Tmp qhi, qlo;
    ix->ev_valid = false;
    VQ_CUDA(cudaEventRecord(ix->ev[0], st));
    VQ_TRY(qhi.alloc((size_t)nq * ix->dpa * 2, st));
    if (split) VQ_TRY(qlo.alloc((size_t)nq * ix->dpa * 2, st));
    VQ_TRY(launch_prepare_rows(xq, dtype, nq, ix->d, ix->dp, true, (__nv_bfloat16 *)qhi.p,
                               split ? (__nv_bfloat16 *)qlo.p : nullptr, nullptr, nullptr, st));
    ++launches;
    CUtensorMap maps[4];
    if (pair) {
        VQ_TRY(make_map(&maps[0], qhi.p, nq, ix->dpa, kTileM));
        VQ_TRY(make_map(&maps[1], ix->hi, ix->n, ix->dpa, kTileN / 2));
    } else {
        VQ_TRY(make_map(&maps[0], qhi.p, nq, ix->dpa, kTileM));
        VQ_TRY(make_map(&maps[1], split ? qlo.p : qhi.p, nq, ix->dpa, kTileM));
        VQ_TRY(make_map(&maps[2], ix->hi, ix->n, ix->dpa, kTileN));
        VQ_TRY(make_map(&maps[3], split ? (void *)ix->lo : (void *)ix->hi, ix->n, ix->dpa, kTileN));
    }
    VQ_TRY(ensure_workspace(ix, spn, nq, st));
    VQ_CUDA(cudaMemcpyAsync(ix->plan_rows_dev, pl.rows.data(), pl.rows.size() * 8, cudaMemcpyHostToDevice, st));
    VQ_CUDA(cudaMemcpyAsync(ix->plan_ints_dev, pl.ints.data(), pl.ints.size() * 4, cudaMemcpyHostToDevice, st));

    GemmTopkParams gp{};
    gp.nq = nq; gp.n = ix->n; gp.kb_count = ix->dp / 64; gp.n_qblocks = n_qblocks; gp.n_seg = pl.nseg;
    gp.seg_begin = ix->plan_rows_dev; gp.seg_end = ix->plan_rows_dev + pl.nseg;
    gp.seg_group = ix->plan_ints_dev; gp.cta_seg_start = ix->plan_ints_dev + pl.nseg;
    gp.seg_pred = ix->plan_ints_dev + pl.nseg + pl.n_cta + 1;
    // 0xff.. = NaN = "threshold of this unit not final yet" (read by successor segments)
    VQ_CUDA(cudaMemsetAsync(ix->cand_tau, 0xff, units * kTileM * 4, st));
    gp.col_video = n_ign > 0 ? ix->video : nullptr;
    gp.col_seg = segmax ? ix->seg_of_row : nullptr;
    gp.q_ignore = n_ign > 0 ? ign : nullptr; gp.n_ign = n_ign; gp.kprime = kprime;
    gp.dbg_no_insert = g_hooks.no_insert;
    gp.trigger = g_hooks.trigger;
    gp.dense_rows = 4;  // measured: scripts/ab_gemm.py, profiles/r01_ab_gemm_chain_dense.txt
    if (g_hooks.dense_rows >= 0) gp.dense_rows = g_hooks.dense_rows;  // A/B hook (33 = never)
    gp.cand = ix->cand; gp.cand_cnt = ix->cand_cnt; gp.cand_tau = ix->cand_tau; gp.cand_cum = ix->cand_cum;

    SelectParams sp{};
    sp.nq = nq; sp.n = ix->n; sp.d = ix->d; sp.dtype = dtype; sp.q = xq; sp.g = ix->raw;
    error_model(prec, ix->dtype, ix->dp, &sp.eps_rel, &sp.eps_abs);
    sp.gmax_norm2_bits = ix->gmax_bits;
    // SEGMAX: the GEMM marks the (query, video) candidates whose coarse arg-max is not provably the exact best moment;
    // all others are settled by one canonical score instead of the whole video (vq_select.cu)
    Tmp amb, reps;
    // (one bit per (query, video): not worth more than 1 GiB -- beyond that every pair takes the video-major re-scoring)
    const bool seg_unique = segmax && !g_hooks.no_vmajor && !g_hooks.no_seg_unique && ix->n < (1ll << 31) &&
                            (double)nq * (double)(((int64_t)ix->seg_start.size() + 30) / 32) * 4.0 <= 1073741824.0;
    if (seg_unique) {
        const int64_t nseg = (int64_t)ix->seg_start.size() - 1;
        const int words = (int)((nseg + 31) / 32);
        VQ_TRY(amb.alloc((size_t)nq * words * 4, st));
        VQ_TRY(reps.alloc((size_t)nq * 4, st));
        VQ_CUDA(cudaMemsetAsync(amb.p, 0, (size_t)nq * words * 4, st));
        VQ_TRY(launch_segmax_row_eps(sp, (float *)reps.p, st));
        ++launches;
        gp.seg_amb = (uint32_t *)amb.p; gp.amb_words = words; gp.row_eps = (const float *)reps.p;
        sp.seg_amb = (const uint32_t *)amb.p; sp.amb_words = words;
    }
    VQ_CUDA(cudaEventRecord(ix->ev[1], st));
    if (pair) {
        VQ_TRY(launch_gemm_topk_pair(maps, gp, segmax, pl.n_cta, st));
    } else {
        VQ_TRY(launch_gemm_topk(maps, gp, split ? 3 : 1, segmax, pl.n_cta, st));
    }
    VQ_CUDA(cudaEventRecord(ix->ev[2], st));
    ++launches;

    sp.group_size = group_size; sp.group_segs = ix->plan_ints_dev + 2 * pl.nseg + pl.n_cta + 1;
    sp.segs_per_group = pl.n_chunks; sp.cand = ix->cand; sp.cand_cnt = ix->cand_cnt;
    sp.cand_tau = ix->cand_tau; sp.cand_cum = ix->cand_cum; sp.kprime = kprime; sp.k = k;
    sp.seg_of_row = segmax ? ix->seg_of_row : nullptr;
    sp.seg_start = segmax ? ix->seg_start_dev : nullptr;
    sp.out_d2 = D; sp.out_idx = I; sp.out_seg = seg_out;
    sp.unproven = ix->unproven; sp.unproven_count = ix->unproven_count;
    Tmp vm;
    if (segmax && !g_hooks.no_vmajor) {
        const int64_t nseg = (int64_t)ix->seg_start.size() - 1;
        VQ_TRY(vm.alloc(select_vmajor_scratch_bytes(nq, kprime, nseg), st));
        VQ_TRY(launch_select_segmax_vmajor(sp, nseg, vm.p, st));
        launches += seg_unique ? 5 : 4;
    } else {
        VQ_TRY(launch_select(sp, st));
    }
    VQ_CUDA(cudaEventRecord(ix->ev[3], st));
    ++launches;

    FallbackParams f{};
    f.nq = nq; f.n = ix->n; f.d = ix->d; f.dtype = dtype; f.q = xq; f.g = ix->raw;
    f.col_video = ix->video; f.q_ignore = n_ign > 0 ? ign : nullptr; f.n_ign = n_ign;
    f.seg_of_row = segmax ? ix->seg_of_row : nullptr;
    f.seg_start = segmax ? ix->seg_start_dev : nullptr;
    f.nseg = segmax ? (int64_t)ix->seg_start.size() - 1 : 0;
    f.k = k; f.out_d2 = D; f.out_idx = I; f.out_seg = seg_out;
    f.unproven = ix->unproven; f.unproven_count = ix->unproven_count;
    VQ_TRY(launch_exact_topk(f, (int)std::min<int64_t>(nq, 2 * sm_count()), st));
    VQ_CUDA(cudaEventRecord(ix->ev[4], st));
    ix->ev_valid = true;
    ++launches;

    int64_t tiles = 0;
    for (int c = 0; c < pl.nseg; ++c) tiles += (pl.rows[(size_t)pl.nseg + c] - pl.rows[(size_t)c] + kTileN - 1) / kTileN;
    ix->stats[1] = tiles * group_size;
    ix->stats[2] = kprime;
    ix->stats[3] = (int64_t)units;
    ix->stats[4] = pl.nseg;
    ix->stats[5] = launches;
    return VQ_OK;
}

// Coarse scores of every (query, gallery row) pair on the tensor cores: out[q, j] = 2<q,g_j> - |g_j|^2 (gemm_topk STORE
// mode: same operand planes, same pipeline, no candidate handling).  |out - (|q|^2 - d2)| <= error_model().  Device pointers.
static int coarse_core(vq_index *ix, const void *xq, int64_t nq, int dtype, float *out, int64_t ld, cudaStream_t st) {
    SearchPlan spn;
    make_search_plan(ix, nq, dtype, 100, false, &spn);
    const bool pair = spn.pair, split = spn.split;
    const SegPlan &pl = spn.pl;
    Tmp qhi, qlo;
    VQ_TRY(qhi.alloc((size_t)nq * ix->dpa * 2, st));
    if (split) VQ_TRY(qlo.alloc((size_t)nq * ix->dpa * 2, st));
    VQ_TRY(launch_prepare_rows(xq, dtype, nq, ix->d, ix->dp, true, (__nv_bfloat16 *)qhi.p,
                               split ? (__nv_bfloat16 *)qlo.p : nullptr, nullptr, nullptr, st));
    CUtensorMap maps[4];
    if (pair) {
        VQ_TRY(make_map(&maps[0], qhi.p, nq, ix->dpa, kTileM));
        VQ_TRY(make_map(&maps[1], ix->hi, ix->n, ix->dpa, kTileN / 2));
    } else {
        VQ_TRY(make_map(&maps[0], qhi.p, nq, ix->dpa, kTileM));
        VQ_TRY(make_map(&maps[1], split ? qlo.p : qhi.p, nq, ix->dpa, kTileM));
        VQ_TRY(make_map(&maps[2], ix->hi, ix->n, ix->dpa, kTileN));
        VQ_TRY(make_map(&maps[3], split ? (void *)ix->lo : (void *)ix->hi, ix->n, ix->dpa, kTileN));
    }
    VQ_TRY(ensure_workspace(ix, spn, nq, st));
    VQ_CUDA(cudaMemcpyAsync(ix->plan_rows_dev, pl.rows.data(), pl.rows.size() * 8, cudaMemcpyHostToDevice, st));
    VQ_CUDA(cudaMemcpyAsync(ix->plan_ints_dev, pl.ints.data(), pl.ints.size() * 4, cudaMemcpyHostToDevice, st));
    GemmTopkParams gp{};
    gp.nq = nq; gp.n = ix->n; gp.kb_count = ix->dp / 64; gp.n_qblocks = spn.n_qblocks; gp.n_seg = pl.nseg;
    gp.seg_begin = ix->plan_rows_dev; gp.seg_end = ix->plan_rows_dev + pl.nseg;
    gp.seg_group = ix->plan_ints_dev; gp.cta_seg_start = ix->plan_ints_dev + pl.nseg;
    gp.seg_pred = ix->plan_ints_dev + pl.nseg + pl.n_cta + 1;
    VQ_CUDA(cudaMemsetAsync(ix->cand_tau, 0xff, spn.units * kTileM * 4, st));
    gp.kprime = spn.kprime; gp.dense_rows = 4;
    gp.cand = ix->cand; gp.cand_cnt = ix->cand_cnt; gp.cand_tau = ix->cand_tau; gp.cand_cum = ix->cand_cum;
    gp.store = out; gp.store_ld = ld;
    if (pair) {
        VQ_TRY(launch_gemm_topk_pair(maps, gp, false, pl.n_cta, st));
    } else {
        VQ_TRY(launch_gemm_topk(maps, gp, split ? 3 : 1, false, pl.n_cta, st));
    }
    return VQ_OK;
}

// Bytes of the [tile_q, n] canonical score matrix the evaluators work on at a time.  One block per query ranks a row,
// so the tile has to hold several queries per SM to keep the memory system busy: a quarter of the free HBM, between
// 256 MiB and 8 GiB (kept mask and per-video keys add ~ 1/3 of that on the moment path).
static int64_t score_tile_bytes() {
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return (int64_t)1 << 31;
    const int64_t quarter = (int64_t)(free_b / 4);
    return std::max<int64_t>((int64_t)256 << 20, std::min<int64_t>((int64_t)8 << 30, quarter));
}

// k > kMaxKPrime: the reference's own call, search(xq, k = ntotal) (dataloader_baseline.py:1592-1597, :935-940,
// :1268-1273) -- canonical scores of a tile of queries against every row, then a stable per-row radix sort
// (vq_sort.cu).  The first k entries of each sorted row are written.  All pointers are device pointers.
static int full_sort_core(vq_index *ix, const void *xq, int64_t nq, int dtype, int64_t k, const int32_t *ign, int n_ign,
                          float *D, int64_t *I, cudaStream_t st) {
    // per query: 4 B score + 2 x 8 B sort buffers per row (+ histograms)
    int64_t tile_q = score_tile_bytes() / std::max<int64_t>(1, ix->n * 20);
    tile_q = std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(nq, tile_q), 32768));
    Tmp d2, a, b, h;
    VQ_TRY(d2.alloc((size_t)tile_q * ix->n * 4, st));
    VQ_TRY(a.alloc((size_t)tile_q * ix->n * 8, st));
    VQ_TRY(b.alloc((size_t)tile_q * ix->n * 8, st));
    VQ_TRY(h.alloc(rowsort_hist_ints(tile_q, ix->n) * 4, st));
    int launches = 0;
    for (int64_t q0 = 0; q0 < nq; q0 += tile_q) {
        const int64_t tq = std::min(tile_q, nq - q0);
        const char *qt = (const char *)xq + (size_t)q0 * ix->d * esize(dtype);
        VQ_TRY(launch_l2sq_all(qt, ix->raw, dtype, tq, ix->n, ix->d, (float *)d2.p, st));
        RowSortParams r{};
        r.nq_tile = tq; r.n = ix->n; r.k = k; r.d2 = (const float *)d2.p;
        r.col_video = n_ign > 0 ? ix->video : nullptr;
        r.q_ignore = n_ign > 0 ? ign + q0 * n_ign : nullptr;
        r.n_ign = n_ign; r.hist = (int32_t *)h.p;
        r.out_d2 = D + q0 * k; r.out_idx = I + q0 * k;
        VQ_TRY(launch_row_sort(r, (uint64_t *)a.p, (uint64_t *)b.p, st));
        launches += 13;
    }
    ix->ev_valid = false;
    ix->stats[0] = 0; ix->stats[1] = 0; ix->stats[2] = 0; ix->stats[3] = 0; ix->stats[4] = 0; ix->stats[5] = launches;
    VQ_CUDA(cudaMemsetAsync(ix->unproven_count, 0, 4, st));
    return VQ_OK;
}

static int check_search_args(const vq_index *ix, const void *xq, int64_t nq, int dtype, int64_t k) {
    VQ_CHECK(ix && xq, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(ix->n > 0, VQ_ERR_STATE, "search on an empty index");
    VQ_CHECK(dtype == ix->dtype, VQ_ERR_INVALID, "query dtype %d differs from the index dtype %d", dtype, ix->dtype);
    VQ_CHECK(nq >= 0 && nq < (1ll << 31), VQ_ERR_INVALID, "nq=%lld", (long long)nq);
    VQ_CHECK(k >= 1, VQ_ERR_INVALID, "k=%lld", (long long)k);
    return VQ_OK;
}

}  // namespace vq

extern "C" {

int vq_index_search(vq_index *ix, const void *xq, int64_t nq, int dtype, int64_t k, const int32_t *ignore_video,
                    int n_ignore, float *D, int64_t *I, void *stream) {
    VQ_TRY(check_search_args(ix, xq, nq, dtype, k));
    VQ_CHECK(D && I, VQ_ERR_INVALID, "NULL output");
    VQ_CHECK(k <= ix->n, VQ_ERR_INVALID, "k=%lld exceeds ntotal=%lld", (long long)k, (long long)ix->n);
    VQ_CHECK(n_ignore >= 0 && n_ignore <= kMaxIgnore, VQ_ERR_INVALID, "n_ignore=%d (max %d)", n_ignore, kMaxIgnore);
    VQ_CHECK(!(ignore_video && n_ignore > 0) || ix->video, VQ_ERR_STATE, "ignore list given but no video ids set");
    if (nq == 0) return VQ_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int n_ign = ignore_video ? n_ignore : 0;
    In q, ig;
    Out oD, oI;
    VQ_TRY(q.bind(xq, (size_t)nq * ix->d * esize(dtype), st));
    VQ_TRY(ig.bind(ignore_video, (size_t)nq * n_ign * 4, st));
    VQ_TRY(oD.bind(D, (size_t)nq * k * 4, st));
    VQ_TRY(oI.bind(I, (size_t)nq * k * 8, st));
    if (k > kMaxKPrime)
        VQ_TRY(full_sort_core(ix, q.dev, nq, dtype, k, ig.as<int32_t>(), n_ign, oD.as<float>(), oI.as<int64_t>(), st));
    else
        VQ_TRY(search_core(ix, q.dev, nq, dtype, (int)k, false, ig.as<int32_t>(), n_ign, oD.as<float>(), oI.as<int64_t>(),
                           nullptr, st));
    VQ_TRY(oD.finish(st));
    VQ_TRY(oI.finish(st));
    if (oD.host || oI.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_index_search_workspace_bytes(vq_index *ix, int64_t nq, int64_t k, int segmax, int64_t *persistent_bytes,
                                    int64_t *transient_bytes) {
    VQ_CHECK(ix && persistent_bytes && transient_bytes, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(ix->n > 0, VQ_ERR_STATE, "empty index");
    VQ_CHECK(nq >= 0 && k >= 1, VQ_ERR_INVALID, "bad sizes");
    if (k > kMaxKPrime || resolve_prec(ix) == VQ_PREC_EXACT) {  // full-sort tiles / exact scan: stream-ordered temporaries only
        int64_t tile_q = score_tile_bytes() / std::max<int64_t>(1, ix->n * 20);
        tile_q = std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(std::max<int64_t>(nq, 1), tile_q), 32768));
        *persistent_bytes = 0;
        *transient_bytes = k > kMaxKPrime ? tile_q * ix->n * 20 + (int64_t)rowsort_hist_ints(tile_q, ix->n) * 4 : 0;
        return VQ_OK;
    }
    SearchPlan sp;
    make_search_plan(ix, nq, ix->dtype, (int)k, segmax != 0, &sp);
    *persistent_bytes = (int64_t)sp.persistent_bytes() + nq * 4;
    *transient_bytes = nq * ix->dpa * 2 * (sp.split ? 2 : 1);
    if (segmax && !ix->seg_start.empty()) {  // video-major re-scoring scratch + the (query, video) ambiguity bits and row bounds
        const int64_t nseg = (int64_t)ix->seg_start.size() - 1;
        *transient_bytes += (int64_t)select_vmajor_scratch_bytes(nq, sp.kprime, nseg) + nq * ((nseg + 31) / 32) * 4 + nq * 4;
    }
    return VQ_OK;
}

int vq_index_reserve(vq_index *ix, int64_t nq, int64_t k, int segmax, void *stream) {
    VQ_CHECK(ix, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(ix->n > 0, VQ_ERR_STATE, "empty index");
    VQ_CHECK(nq >= 1 && k >= 1, VQ_ERR_INVALID, "bad sizes");
    if (k > kMaxKPrime || resolve_prec(ix) == VQ_PREC_EXACT) return VQ_OK;
    SearchPlan sp;
    make_search_plan(ix, nq, ix->dtype, (int)k, segmax != 0, &sp);
    return ensure_workspace(ix, sp, nq, (cudaStream_t)stream);
}

int vq_index_search_segmax(vq_index *ix, const void *xq, int64_t nq, int dtype, int64_t k, int64_t *video,
                           int64_t *row, float *D, void *stream) {
    VQ_TRY(check_search_args(ix, xq, nq, dtype, k));
    VQ_CHECK(video && row && D, VQ_ERR_INVALID, "NULL output");
    VQ_CHECK(!ix->seg_start.empty(), VQ_ERR_STATE, "vq_index_set_segments has not been called");
    const int64_t nseg = (int64_t)ix->seg_start.size() - 1;
    VQ_CHECK(k <= nseg, VQ_ERR_INVALID, "k=%lld exceeds the number of videos %lld", (long long)k, (long long)nseg);
    VQ_CHECK(k <= kMaxKPrime, VQ_ERR_CAPACITY, "k=%lld exceeds %d", (long long)k, kMaxKPrime);
    if (nq == 0) return VQ_OK;
    cudaStream_t st = (cudaStream_t)stream;
    In q;
    Out oV, oR, oD;
    VQ_TRY(q.bind(xq, (size_t)nq * ix->d * esize(dtype), st));
    VQ_TRY(oV.bind(video, (size_t)nq * k * 8, st));
    VQ_TRY(oR.bind(row, (size_t)nq * k * 8, st));
    VQ_TRY(oD.bind(D, (size_t)nq * k * 4, st));
    VQ_TRY(search_core(ix, q.dev, nq, dtype, (int)k, true, nullptr, 0, oD.as<float>(), oR.as<int64_t>(),
                       oV.as<int64_t>(), st));
    VQ_TRY(oV.finish(st));
    VQ_TRY(oR.finish(st));
    VQ_TRY(oD.finish(st));
    if (oV.host || oR.host || oD.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_index_eval_trimmed(vq_index *ix, const void *xq, int64_t nq, int dtype, const int32_t *q_label,
                          const int32_t *ignore_video, int n_ignore, double *ap, double *recall, int64_t *top_idx,
                          void *stream) {
    VQ_TRY(check_search_args(ix, xq, nq, dtype, 1));
    VQ_CHECK(q_label && ap && recall, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(ix->label, VQ_ERR_STATE, "vq_index_set_meta (labels) has not been called");
    VQ_CHECK(n_ignore >= 0 && n_ignore <= kMaxIgnore, VQ_ERR_INVALID, "n_ignore=%d (max %d)", n_ignore, kMaxIgnore);
    const int n_ign = ignore_video ? n_ignore : 0;
    VQ_CHECK(n_ign == 0 || ix->video, VQ_ERR_STATE, "ignore list given but no video ids set");
    if (nq == 0) return VQ_OK;
    cudaStream_t st = (cudaStream_t)stream;
    In q, ql, ig;
    Out oA, oR, oT;
    VQ_TRY(q.bind(xq, (size_t)nq * ix->d * esize(dtype), st));
    VQ_TRY(ql.bind(q_label, (size_t)nq * 4, st));
    VQ_TRY(ig.bind(ignore_video, (size_t)nq * n_ign * 4, st));
    VQ_TRY(oA.bind(ap, (size_t)nq * 8, st));
    VQ_TRY(oR.bind(recall, (size_t)nq * 3 * 8, st));
    VQ_TRY(oT.bind(top_idx, (size_t)nq * 100 * 8, st));
    // full-list AP / R@N: scores for a tile of queries, then the rank-count reduction
    const int64_t ld = (ix->n + 3) / 4 * 4;
    const int64_t tile_q = std::max<int64_t>(1, std::min<int64_t>(nq, score_tile_bytes() / std::max<int64_t>(1, ld * 4)));
    const int64_t ntile = (nq + tile_q - 1) / tile_q;
    Tmp d2, ovf, redo, redo_cnt;
    VQ_TRY(d2.alloc((size_t)tile_q * ld * 4, st));
    VQ_TRY(ovf.alloc(4, st));
    int launches = 0;
    // Tensor-core form (ap_rows_tc_kernel): coarse scores from the tcgen05 GEMM, canonical arithmetic only for the
    // positives and for the survivors inside an eps band of one of them; queries it hands back (list overflow) make
    // their whole tile go through the canonical path below.
    const bool use_tc = !g_hooks.no_tc_eval && !ix->tc_eval_off && resolve_prec(ix) != VQ_PREC_EXACT;
    std::vector<int32_t> tile_redo((size_t)ntile, use_tc ? 0 : 1);
    if (use_tc) {
        VQ_TRY(redo.alloc((size_t)nq * 4, st));
        VQ_TRY(redo_cnt.alloc((size_t)ntile * 4, st));
        VQ_CUDA(cudaMemsetAsync(redo_cnt.p, 0, (size_t)ntile * 4, st));
        for (int64_t q0 = 0, ti = 0; q0 < nq; q0 += tile_q, ++ti) {
            const int64_t tq = std::min(tile_q, nq - q0);
            const char *qt = (const char *)q.dev + (size_t)q0 * ix->d * esize(dtype);
            VQ_TRY(coarse_core(ix, qt, tq, dtype, (float *)d2.p, ld, st));
            ApTcParams t{};
            t.nq_tile = tq; t.n = ix->n; t.ld = ld; t.d = ix->d; t.dtype = dtype; t.coarse = (const float *)d2.p;
            t.q = qt; t.g = ix->raw; t.g_label = ix->label; t.g_video = ix->video;
            t.q_label = ql.as<int32_t>() + q0;
            t.q_ignore = n_ign ? ig.as<int32_t>() + q0 * n_ign : nullptr;
            t.n_ign = n_ign;
            error_model(resolve_prec(ix), ix->dtype, ix->dp, &t.eps_rel, &t.eps_abs);
            t.gmax_norm2_bits = ix->gmax_bits;
            t.ap = oA.as<double>() + q0; t.recall = oR.as<double>() + q0 * 3;
            t.redo = (int32_t *)redo.p + q0; t.redo_count = (int32_t *)redo_cnt.p + ti;
            VQ_TRY(launch_ap_rows_tc(t, st));
            launches += 3;
        }
        VQ_CUDA(cudaMemcpyAsync(tile_redo.data(), redo_cnt.p, (size_t)ntile * 4, cudaMemcpyDeviceToHost, st));
        VQ_CUDA(cudaStreamSynchronize(st));
    }
    bool any_redo = false;
    for (int32_t r : tile_redo) any_redo |= r != 0;
    ix->stats[4] = 0;
    for (int32_t r : tile_redo) ix->stats[4] += use_tc ? r : 0;  // queries handed back to the canonical path
    // positives spread through the bulk of the list (labels unrelated to the features): the eps bands cover too much of
    // the gallery for the tensor-core form to pay -- later calls on this index go straight to the canonical path
    if (use_tc && ix->stats[4] * 4 > nq) ix->tc_eval_off = true;
    for (int attempt = 0; attempt < 2 && any_redo; ++attempt) {
        const int pcap2 = attempt == 0 ? 1024 : kMaxPositives;
        VQ_CUDA(cudaMemsetAsync(ovf.p, 0, 4, st));
        for (int64_t q0 = 0, ti = 0; q0 < nq; q0 += tile_q, ++ti) {
            if (!tile_redo[(size_t)ti]) continue;
            const int64_t tq = std::min(tile_q, nq - q0);
            const char *qt = (const char *)q.dev + (size_t)q0 * ix->d * esize(dtype);
            VQ_TRY(launch_l2sq_all(qt, ix->raw, dtype, tq, ix->n, ix->d, (float *)d2.p, st));
            ApRowsParams a{};
            a.nq_tile = tq; a.n = ix->n; a.d2 = (const float *)d2.p; a.g_label = ix->label; a.g_video = ix->video;
            a.q_label = ql.as<int32_t>() + q0;
            a.q_ignore = n_ign ? ig.as<int32_t>() + q0 * n_ign : nullptr;
            a.n_ign = n_ign; a.ap = oA.as<double>() + q0; a.recall = oR.as<double>() + q0 * 3;
            a.top_idx = nullptr; a.overflow = (int32_t *)ovf.p;
            VQ_TRY(launch_ap_rows(a, pcap2, st));
            launches += 2;
        }
        int32_t h = 0;
        VQ_CUDA(cudaMemcpyAsync(&h, ovf.p, 4, cudaMemcpyDeviceToHost, st));
        VQ_CUDA(cudaStreamSynchronize(st));
        if (!h) break;
        VQ_CHECK(attempt == 0, VQ_ERR_CAPACITY, "a query has more than %d positives in the gallery", kMaxPositives - 1);
    }
    if (top_idx) {  // the first 100 survivors in rank order (cm_dict side data, :437-442)
        const int64_t k = std::min<int64_t>(100, ix->n);
        Tmp tD, tI;
        VQ_TRY(tD.alloc((size_t)nq * k * 4, st));
        VQ_TRY(tI.alloc((size_t)nq * k * 8, st));
        VQ_TRY(search_core(ix, q.dev, nq, dtype, (int)k, false, ig.as<int32_t>(), n_ign, (float *)tD.p, (int64_t *)tI.p,
                           nullptr, st));
        launches += (int)ix->stats[5];
        if (k == 100) {
            VQ_CUDA(cudaMemcpyAsync(oT.dev, tI.p, (size_t)nq * 100 * 8, cudaMemcpyDeviceToDevice, st));
        } else {
            VQ_CUDA(cudaMemsetAsync(oT.dev, 0xff, (size_t)nq * 100 * 8, st));
            VQ_CUDA(cudaMemcpy2DAsync(oT.dev, 100 * 8, tI.p, (size_t)k * 8, (size_t)k * 8, (size_t)nq,
                                      cudaMemcpyDeviceToDevice, st));
        }
    }
    ix->stats[5] = launches;
    VQ_TRY(oA.finish(st));
    VQ_TRY(oR.finish(st));
    VQ_TRY(oT.finish(st));
    VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_index_eval_moment(vq_index *ix, const void *xq, int64_t nq, int dtype, const int32_t *q_label,
                         const int32_t *ignore_video, int n_ignore, double nms_thresh, double *ap, double *recall,
                         int64_t *n_kept, uint8_t *kept_mask, void *stream) {
    VQ_TRY(check_search_args(ix, xq, nq, dtype, 1));
    VQ_CHECK(q_label && ap && recall, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(ix->m_start, VQ_ERR_STATE, "vq_index_set_moments has not been called");
    VQ_CHECK(!ix->seg_start.empty() && ix->seg_of_row, VQ_ERR_STATE, "vq_index_set_segments has not been called");
    VQ_CHECK(ix->video, VQ_ERR_STATE, "vq_index_set_meta (video ids) has not been called");
    VQ_CHECK(ix->seg_max_len <= kMomentBlockCap, VQ_ERR_CAPACITY, "a video has %lld moments (max %d)",
             (long long)ix->seg_max_len, kMomentBlockCap);
    VQ_CHECK(n_ignore >= 0 && n_ignore <= kMaxIgnore, VQ_ERR_INVALID, "n_ignore=%d (max %d)", n_ignore, kMaxIgnore);
    const int n_ign = ignore_video ? n_ignore : 0;
    if (nq == 0) return VQ_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t nseg = (int64_t)ix->seg_start.size() - 1;
    In q, ql, ig;
    Out oA, oR, oN, oK;
    VQ_TRY(q.bind(xq, (size_t)nq * ix->d * esize(dtype), st));
    VQ_TRY(ql.bind(q_label, (size_t)nq * 4, st));
    VQ_TRY(ig.bind(ignore_video, (size_t)nq * n_ign * 4, st));
    VQ_TRY(oA.bind(ap, (size_t)nq * 8, st));
    VQ_TRY(oR.bind(recall, (size_t)nq * 3 * 8, st));
    VQ_TRY(oN.bind(n_kept, (size_t)nq * 8, st));
    VQ_TRY(oK.bind(kept_mask, (size_t)nq * ix->n, st));
    int64_t tile_q = std::max<int64_t>(1, std::min<int64_t>(nq, score_tile_bytes() / std::max<int64_t>(1, ix->n * 4)));
    tile_q = std::min<int64_t>(tile_q, 32768);  // grid.y of the NMS kernel
    Tmp d2, kept, vbest, ovf;
    VQ_TRY(d2.alloc((size_t)tile_q * ix->n * 4, st));
    if (!kept_mask) VQ_TRY(kept.alloc((size_t)tile_q * ix->n, st));
    VQ_TRY(vbest.alloc((size_t)tile_q * nseg * 8, st));
    VQ_TRY(ovf.alloc(4, st));
    int launches = 0;
    if (ix->n_seg_small > 0 && !g_hooks.no_nms_masks && (!ix->m_sup || ix->m_sup_thresh != nms_thresh)) {
        // the suppression relation between the moments of a video does not depend on the query: once per threshold
        if (!ix->m_sup) VQ_CUDA(cudaMalloc(&ix->m_sup, (size_t)ix->n * 8));
        VQ_TRY(launch_nms_masks(ix->seg_start_dev, ix->seg_small, ix->n_seg_small, ix->m_start, ix->m_end, nms_thresh, ix->m_sup, st));
        ix->m_sup_thresh = nms_thresh;
        ++launches;
    }
    for (int attempt = 0; attempt < 2; ++attempt) {
        const int pcap2 = attempt == 0 ? 1024 : kMaxPositives;
        VQ_CUDA(cudaMemsetAsync(ovf.p, 0, 4, st));
        for (int64_t q0 = 0; q0 < nq; q0 += tile_q) {
            const int64_t tq = std::min(tile_q, nq - q0);
            const char *qt = (const char *)q.dev + (size_t)q0 * ix->d * esize(dtype);
            uint8_t *kp = kept_mask ? oK.as<uint8_t>() + (size_t)q0 * ix->n : (uint8_t *)kept.p;
            VQ_TRY(launch_l2sq_all(qt, ix->raw, dtype, tq, ix->n, ix->d, (float *)d2.p, st));
            MomentNmsParams m{};
            m.nq_tile = tq; m.n = ix->n; m.nseg = nseg; m.d2 = (const float *)d2.p; m.seg_start = ix->seg_start_dev;
            m.start = ix->m_start; m.end = ix->m_end; m.thresh = nms_thresh; m.kept = kp; m.vbest = (uint64_t *)vbest.p;
            m.sup = (!g_hooks.no_nms_masks && ix->m_sup_thresh == nms_thresh) ? ix->m_sup : nullptr;
            VQ_TRY(launch_moment_nms(m, ix->seg_small, ix->n_seg_small, ix->seg_large, ix->n_seg_large, st));
            ApRowsParams a{};
            a.nq_tile = tq; a.n = ix->n; a.d2 = (const float *)d2.p; a.g_label = ix->m_hit; a.g_video = ix->video;
            a.q_label = ql.as<int32_t>() + q0;
            a.q_ignore = n_ign ? ig.as<int32_t>() + q0 * n_ign : nullptr;
            a.n_ign = n_ign; a.ap = oA.as<double>() + q0; a.recall = oR.as<double>() + q0 * 3;
            a.top_idx = nullptr; a.overflow = (int32_t *)ovf.p;
            a.kept = kp; a.g_iou = ix->m_iou; a.vbest = (uint64_t *)vbest.p; a.seg_start = ix->seg_start_dev;
            a.seg_of_row = ix->seg_of_row; a.nseg = nseg; a.n_kept = n_kept ? oN.as<int64_t>() + q0 : nullptr;
            VQ_TRY(launch_ap_rows(a, pcap2, st));
            launches += 2 + (ix->n_seg_small > 0) + (ix->n_seg_large > 0);
        }
        int32_t h = 0;
        VQ_CUDA(cudaMemcpyAsync(&h, ovf.p, 4, cudaMemcpyDeviceToHost, st));
        VQ_CUDA(cudaStreamSynchronize(st));
        if (!h) break;
        VQ_CHECK(attempt == 0, VQ_ERR_CAPACITY, "a query has more than %d positives in the gallery", kMaxPositives - 1);
    }
    ix->stats[5] = launches;
    VQ_TRY(oA.finish(st));
    VQ_TRY(oR.finish(st));
    VQ_TRY(oN.finish(st));
    VQ_TRY(oK.finish(st));
    VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_average_precision(const int32_t *y_true, const double *y_score, const int64_t *offsets, int64_t nlists,
                         double *ap, void *stream) {
    VQ_CHECK(y_true && y_score && offsets && ap, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(!is_device_ptr(offsets), VQ_ERR_INVALID, "offsets must be host memory");
    VQ_TRY(check_device());
    if (nlists == 0) return VQ_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t total = offsets[nlists];
    In yt, ys, of;
    Out oA;
    Tmp ovf;
    VQ_TRY(yt.bind(y_true, (size_t)total * 4, st));
    VQ_TRY(ys.bind(y_score, (size_t)total * 8, st));
    VQ_TRY(of.bind(offsets, (size_t)(nlists + 1) * 8, st));
    VQ_TRY(oA.bind(ap, (size_t)nlists * 8, st));
    VQ_TRY(ovf.alloc(4, st));
    VQ_CUDA(cudaMemsetAsync(ovf.p, 0, 4, st));
    VQ_TRY(launch_average_precision(yt.as<int32_t>(), ys.as<double>(), of.as<int64_t>(), nlists, 0, oA.as<double>(),
                                    (int32_t *)ovf.p, st));
    int32_t h = 0;
    VQ_CUDA(cudaMemcpyAsync(&h, ovf.p, 4, cudaMemcpyDeviceToHost, st));
    VQ_TRY(oA.finish(st));
    VQ_CUDA(cudaStreamSynchronize(st));
    VQ_CHECK(!h, VQ_ERR_CAPACITY, "a list has more than %d positives", kMaxPositives);
    return VQ_OK;
}

int vq_nms1d_segmented(const double *dets, const int64_t *offsets, int64_t nseg, double thresh, uint8_t *keep,
                       void *stream) {
    VQ_CHECK(dets && offsets && keep, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(!is_device_ptr(offsets), VQ_ERR_INVALID, "offsets must be host memory");
    VQ_TRY(check_device());
    if (nseg == 0) return VQ_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t total = offsets[nseg];
    In dd, of;
    Out oK;
    Tmp dead;
    VQ_TRY(dd.bind(dets, (size_t)total * 3 * 8, st));
    VQ_TRY(of.bind(offsets, (size_t)(nseg + 1) * 8, st));
    VQ_TRY(oK.bind(keep, (size_t)total, st));
    VQ_TRY(dead.alloc((size_t)total, st));
    VQ_TRY(launch_nms1d(dd.as<double>(), of.as<int64_t>(), nseg, thresh, oK.as<uint8_t>(), (uint8_t *)dead.p, st));
    VQ_TRY(oK.finish(st));
    if (oK.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_window_means(const float *feat, int C, int64_t Ttot, const int64_t *vid_t0, const int64_t *win_start, int64_t V,
                    const int32_t *lo, const int32_t *hi, float *out, void *stream) {
    VQ_CHECK(feat && vid_t0 && win_start && lo && hi && out, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(!is_device_ptr(vid_t0) && !is_device_ptr(win_start), VQ_ERR_INVALID, "vid_t0 / win_start must be host memory");
    VQ_CHECK(C >= 1 && V >= 0 && Ttot >= 0, VQ_ERR_INVALID, "bad sizes");
    VQ_TRY(check_device());
    if (V == 0) return VQ_OK;
    VQ_CHECK(vid_t0[0] == 0 && vid_t0[V] == Ttot && win_start[0] == 0, VQ_ERR_INVALID, "offsets must start at 0 and cover Ttot");
    int64_t Tmax = 0;
    for (int64_t v = 0; v < V; ++v) {
        VQ_CHECK(vid_t0[v] <= vid_t0[v + 1] && win_start[v] <= win_start[v + 1], VQ_ERR_INVALID, "offsets not monotone at %lld",
                 (long long)v);
        Tmax = std::max(Tmax, vid_t0[v + 1] - vid_t0[v]);
    }
    const int64_t nw = win_start[V];
    cudaStream_t st = (cudaStream_t)stream;
    In f, t0, ws, l, h;
    Out o;
    VQ_TRY(f.bind(feat, (size_t)C * Ttot * 4, st));
    VQ_TRY(t0.bind(vid_t0, (size_t)(V + 1) * 8, st));
    VQ_TRY(ws.bind(win_start, (size_t)(V + 1) * 8, st));
    VQ_TRY(l.bind(lo, (size_t)nw * 4, st));
    VQ_TRY(h.bind(hi, (size_t)nw * 4, st));
    VQ_TRY(o.bind(out, (size_t)nw * C * 4, st));
    VQ_TRY(launch_window_means(f.as<float>(), C, Ttot, t0.as<int64_t>(), ws.as<int64_t>(), V, Tmax, l.as<int32_t>(),
                               h.as<int32_t>(), o.as<float>(), st));
    VQ_TRY(o.finish(st));
    VQ_CUDA(cudaStreamSynchronize(st));  // vid_t0 / win_start staging copies read host memory
    return VQ_OK;
}

int vq_margin_given(const float *emb, int64_t B, int d, const int32_t *trip, int64_t T, const float *beta_t, float margin,
                    float *loss, float *grad, float *grad_beta_t, float *pair_count, void *stream) {
    VQ_CHECK(emb && trip && beta_t && loss, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(B >= 1 && d >= 1 && T >= 0, VQ_ERR_INVALID, "bad sizes");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In e, tr, be;
    Out oL, oG, oB, oP;
    Tmp sc;
    VQ_TRY(e.bind(emb, (size_t)B * d * 4, st));
    VQ_TRY(tr.bind(trip, (size_t)T * 12, st));
    VQ_TRY(be.bind(beta_t, (size_t)T * 4, st));
    VQ_TRY(oL.bind(loss, 4, st));
    VQ_TRY(oG.bind(grad, (size_t)B * d * 4, st));
    VQ_TRY(oB.bind(grad_beta_t, (size_t)T * 4, st));
    VQ_TRY(oP.bind(pair_count, 4, st));
    VQ_TRY(sc.alloc((size_t)(4 * T + 4) * 4, st));
    VQ_TRY(launch_margin_given(e.as<float>(), B, d, tr.as<int32_t>(), T, be.as<float>(), 1, margin, oL.as<float>(),
                               grad ? oG.as<float>() : nullptr, grad_beta_t ? oB.as<float>() : nullptr,
                               pair_count ? oP.as<float>() : nullptr, (float *)sc.p, st));
    VQ_TRY(oL.finish(st));
    VQ_TRY(oG.finish(st));
    VQ_TRY(oB.finish(st));
    VQ_TRY(oP.finish(st));
    if (oL.host || oG.host || oB.host || oP.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_npair_given(const float *emb, int64_t B, int d, const int32_t *tuples, int64_t P, int N, float l2, float *loss,
                   float *grad, void *stream) {
    VQ_CHECK(emb && tuples && loss, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(B >= 1 && d >= 1 && P >= 0 && N >= 0 && P <= 65535, VQ_ERR_INVALID, "bad sizes");
    VQ_CHECK(((size_t)d + N) * 4 <= 200 * 1024, VQ_ERR_CAPACITY, "d + N too large for one block");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In e, tu;
    Out oL, oG;
    Tmp sc;
    VQ_TRY(e.bind(emb, (size_t)B * d * 4, st));
    VQ_TRY(tu.bind(tuples, (size_t)P * (2 + N) * 4, st));
    VQ_TRY(oL.bind(loss, 4, st));
    VQ_TRY(oG.bind(grad, (size_t)B * d * 4, st));
    VQ_TRY(sc.alloc(((size_t)P * N + 3 * P + 4) * 4, st));
    VQ_TRY(launch_npair_given(e.as<float>(), B, d, tu.as<int32_t>(), (int)P, N, l2, oL.as<float>(),
                              grad ? oG.as<float>() : nullptr, (float *)sc.p, st));
    VQ_TRY(oL.finish(st));
    VQ_TRY(oG.finish(st));
    if (oL.host || oG.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_proxynca(const float *emb, const float *proxies, const int32_t *labels, int64_t B, int64_t C, int d, float *loss,
                float *grad_emb, float *grad_proxies, void *stream) {
    VQ_CHECK(emb && proxies && labels && loss, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(B >= 1 && C >= 2 && d >= 1 && d <= 12288 && B <= (1 << 20) && C <= (1 << 20), VQ_ERR_INVALID, "bad sizes");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In e, pr, la;
    Out oL, oG, oP;
    Tmp sc;
    VQ_TRY(e.bind(emb, (size_t)B * d * 4, st));
    VQ_TRY(pr.bind(proxies, (size_t)C * d * 4, st));
    VQ_TRY(la.bind(labels, (size_t)B * 4, st));
    VQ_TRY(oL.bind(loss, 4, st));
    VQ_TRY(oG.bind(grad_emb, (size_t)B * d * 4, st));
    VQ_TRY(oP.bind(grad_proxies, (size_t)C * d * 4, st));
    VQ_TRY(sc.alloc(proxynca_scratch_floats((int)B, (int)C, d) * 4, st));
    VQ_TRY(launch_proxynca(e.as<float>(), pr.as<float>(), la.as<int32_t>(), (int)B, (int)C, d, oL.as<float>(),
                           grad_emb ? oG.as<float>() : nullptr, grad_proxies ? oP.as<float>() : nullptr, (float *)sc.p, st));
    VQ_TRY(oL.finish(st));
    VQ_TRY(oG.finish(st));
    VQ_TRY(oP.finish(st));
    if (oL.host || oG.host || oP.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_sphere_weights(const float *dist, const int32_t *labels, int64_t B, int dim, float *q, void *stream) {
    VQ_CHECK(dist && labels && q, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(B >= 1 && B <= 65535 && dim >= 1, VQ_ERR_INVALID, "bad sizes");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In di, la;
    Out o;
    VQ_TRY(di.bind(dist, (size_t)B * B * 4, st));
    VQ_TRY(la.bind(labels, (size_t)B * 4, st));
    VQ_TRY(o.bind(q, (size_t)B * B * 4, st));
    VQ_TRY(launch_sphere_weights(di.as<float>(), la.as<int32_t>(), (int)B, dim, o.as<float>(), st));
    VQ_TRY(o.finish(st));
    if (o.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_memory_logits(const float *x, const float *mem, int64_t B, int64_t K, int d, float temperature, float *logits,
                     const float *grad_logits, float *grad_x, float *grad_mem, void *stream) {
    VQ_CHECK(x && mem && (logits || grad_logits), VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(B >= 1 && K >= 1 && d >= 1 && d <= 12288 && B <= (1 << 20) && K <= (1 << 20), VQ_ERR_INVALID, "bad sizes");
    VQ_CHECK(temperature > 0.f, VQ_ERR_INVALID, "temperature must be positive");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In xi, mi, gi;
    Out oL, oX, oM;
    Tmp sc;
    VQ_TRY(xi.bind(x, (size_t)B * d * 4, st));
    VQ_TRY(mi.bind(mem, (size_t)K * d * 4, st));
    VQ_TRY(gi.bind(grad_logits, (size_t)B * K * 4, st));
    VQ_TRY(oL.bind(logits, (size_t)B * K * 4, st));
    VQ_TRY(oX.bind(grad_x, (size_t)B * d * 4, st));
    VQ_TRY(oM.bind(grad_mem, (size_t)K * d * 4, st));
    VQ_TRY(sc.alloc(memory_logits_scratch_floats((int)B, (int)K, d) * 4, st));
    VQ_TRY(launch_memory_logits(xi.as<float>(), mi.as<float>(), (int)B, (int)K, d, temperature,
                                logits ? oL.as<float>() : nullptr, gi.as<float>(), grad_x ? oX.as<float>() : nullptr,
                                grad_mem ? oM.as<float>() : nullptr, (float *)sc.p, st));
    VQ_TRY(oL.finish(st));
    VQ_TRY(oX.finish(st));
    VQ_TRY(oM.finish(st));
    if (oL.host || oX.host || oM.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_pdist(const float *A, int64_t B, int d, float eps, float *dist, void *stream) {
    VQ_CHECK(A && dist, VQ_ERR_INVALID, "NULL argument");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In a;
    Out o;
    VQ_TRY(a.bind(A, (size_t)B * d * 4, st));
    VQ_TRY(o.bind(dist, (size_t)B * B * 4, st));
    VQ_TRY(launch_pdist(a.as<float>(), B, d, eps, o.as<float>(), st));
    VQ_TRY(o.finish(st));
    if (o.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

static int32_t *g_triplet_handed_back = nullptr;

int vq_triplet_last_handed_back(void) {
    int32_t h = 0;
    if (!g_triplet_handed_back) return 0;
    if (cudaDeviceSynchronize() != cudaSuccess || cudaMemcpy(&h, g_triplet_handed_back, 4, cudaMemcpyDeviceToHost) != cudaSuccess)
        return VQ_ERR_CUDA;
    return h;
}

int vq_triplet_batchhard(const float *emb, const int32_t *labels, int64_t G, int64_t B, int d, float margin,
                         float *loss, float *grad, int32_t *trip, void *stream) {
    VQ_CHECK(emb && labels && loss, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(G >= 1 && B >= 1 && d >= 1, VQ_ERR_INVALID, "bad shape G=%lld B=%lld d=%d", (long long)G, (long long)B, d);
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In e, l;
    Out oL, oG, oT;
    Tmp lt, tr;
    VQ_TRY(e.bind(emb, (size_t)G * B * d * 4, st));
    VQ_TRY(l.bind(labels, (size_t)G * B * 4, st));
    VQ_TRY(oL.bind(loss, (size_t)G * 4, st));
    VQ_TRY(oG.bind(grad, (size_t)G * B * d * 4, st));
    VQ_TRY(oT.bind(trip, (size_t)G * B * 3 * 4, st));
    VQ_TRY(lt.alloc((size_t)G * B * 4, st));
    int32_t *tdev = oT.as<int32_t>();
    if (!tdev) {
        VQ_TRY(tr.alloc((size_t)G * B * 3 * 4, st));
        tdev = (int32_t *)tr.p;
    }
    Tmp wk;
    VQ_TRY(wk.alloc((size_t)(G + 1) * 4, st));
    VQ_CUDA(cudaMemsetAsync(wk.p, 0, 4, st));
    VQ_TRY(launch_triplet_batchhard(e.as<float>(), l.as<int32_t>(), G, B, d, margin, oL.as<float>(), oG.as<float>(),
                                    tdev, (float *)lt.p, (int32_t *)wk.p, st));
    if (!g_triplet_handed_back) VQ_CUDA(cudaMalloc(&g_triplet_handed_back, 4));
    VQ_CUDA(cudaMemcpyAsync(g_triplet_handed_back, wk.p, 4, cudaMemcpyDeviceToDevice, st));
    VQ_TRY(oL.finish(st));
    VQ_TRY(oG.finish(st));
    VQ_TRY(oT.finish(st));
    if (oL.host || oG.host || oT.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_triplet_given(const float *emb, int64_t B, int d, const int32_t *trip, int64_t T, float margin, float *loss,
                     float *grad, void *stream) {
    VQ_CHECK(emb && loss && (trip || T == 0), VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(B >= 1 && d >= 1 && T >= 0, VQ_ERR_INVALID, "bad shape");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    if (!is_device_ptr(trip))
        for (int64_t i = 0; i < T * 3; ++i)
            VQ_CHECK(trip[i] >= 0 && trip[i] < B, VQ_ERR_INVALID, "triplet index %d out of range", trip[i]);
    In e, t;
    Out oL, oG;
    Tmp lt;
    VQ_TRY(e.bind(emb, (size_t)B * d * 4, st));
    VQ_TRY(t.bind(trip, (size_t)T * 3 * 4, st));
    VQ_TRY(oL.bind(loss, 4, st));
    VQ_TRY(oG.bind(grad, (size_t)B * d * 4, st));
    VQ_TRY(lt.alloc((size_t)std::max<int64_t>(T, 1) * 4, st));
    VQ_TRY(launch_triplet_given(e.as<float>(), B, d, t.as<int32_t>(), T, margin, oL.as<float>(), oG.as<float>(),
                                (float *)lt.p, st));
    VQ_TRY(oL.finish(st));
    VQ_TRY(oG.finish(st));
    if (oL.host || oG.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_pool_normalize(const float *x, int64_t B, int C, int T, float *out, void *stream) {
    VQ_CHECK(x && out, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(B >= 0 && C >= 1 && T >= 1 && T <= 8192, VQ_ERR_INVALID, "bad shape");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In a;
    Out o;
    VQ_TRY(a.bind(x, (size_t)B * C * T * 4, st));
    VQ_TRY(o.bind(out, (size_t)B * C * 4, st));
    VQ_TRY(launch_pool_normalize(a.as<float>(), B, C, T, o.as<float>(), st));
    VQ_TRY(o.finish(st));
    if (o.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_topk_merge(const float *d2, const int64_t *idx, int G, int64_t nq, int64_t k, float *out_d2, int64_t *out_idx,
                  void *stream) {
    VQ_CHECK(d2 && idx && out_d2 && out_idx, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(G >= 1 && nq >= 0 && k >= 1, VQ_ERR_INVALID, "bad shape");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In a, b;
    Out oD, oI;
    VQ_TRY(a.bind(d2, (size_t)G * nq * k * 4, st));
    VQ_TRY(b.bind(idx, (size_t)G * nq * k * 8, st));
    VQ_TRY(oD.bind(out_d2, (size_t)nq * k * 4, st));
    VQ_TRY(oI.bind(out_idx, (size_t)nq * k * 8, st));
    VQ_TRY(launch_topk_merge(a.as<float>(), b.as<int64_t>(), G, nq, k, oD.as<float>(), oI.as<int64_t>(), st));
    VQ_TRY(oD.finish(st));
    VQ_TRY(oI.finish(st));
    if (oD.host || oI.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_topk_pack(const float *d2, const int64_t *idx, int64_t count, int64_t row_offset, uint64_t *keys, void *stream) {
    VQ_CHECK(d2 && idx && keys, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(count >= 0 && row_offset >= 0, VQ_ERR_INVALID, "bad shape");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In a, b;
    Out o;
    VQ_TRY(a.bind(d2, (size_t)count * 4, st));
    VQ_TRY(b.bind(idx, (size_t)count * 8, st));
    VQ_TRY(o.bind(keys, (size_t)count * 8, st));
    VQ_TRY(launch_topk_pack(a.as<float>(), b.as<int64_t>(), count, row_offset, o.as<uint64_t>(), st));
    VQ_TRY(o.finish(st));
    if (o.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_topk_merge_keys(const uint64_t *keys, int G, int64_t nq, int64_t k, float *out_d2, int64_t *out_idx, void *stream) {
    VQ_CHECK(keys && out_d2 && out_idx, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(G >= 1 && nq >= 0 && k >= 1, VQ_ERR_INVALID, "bad shape");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In a;
    Out oD, oI;
    VQ_TRY(a.bind(keys, (size_t)G * nq * k * 8, st));
    VQ_TRY(oD.bind(out_d2, (size_t)nq * k * 4, st));
    VQ_TRY(oI.bind(out_idx, (size_t)nq * k * 8, st));
    VQ_TRY(launch_topk_merge_keys(a.as<uint64_t>(), G, nq, k, oD.as<float>(), oI.as<int64_t>(), st));
    VQ_TRY(oD.finish(st));
    VQ_TRY(oI.finish(st));
    if (oD.host || oI.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_topk_merge_heads(const uint64_t *keys, int G, int64_t nq, int64_t m, int64_t k, float *out_d2, int64_t *out_idx,
                        int32_t *incomplete, void *stream) {
    VQ_CHECK(keys && out_d2 && out_idx && incomplete, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(G >= 1 && nq >= 0 && m >= 1 && k >= 1 && (int64_t)G * m >= k, VQ_ERR_INVALID, "bad shape (G*m must reach k)");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In a;
    Out oD, oI, oC;
    VQ_TRY(a.bind(keys, (size_t)G * nq * m * 8, st));
    VQ_TRY(oD.bind(out_d2, (size_t)nq * k * 4, st));
    VQ_TRY(oI.bind(out_idx, (size_t)nq * k * 8, st));
    VQ_TRY(oC.bind(incomplete, 4, st));
    VQ_CUDA(cudaMemsetAsync(oC.dev, 0, 4, st));
    VQ_TRY(launch_topk_merge_heads(a.as<uint64_t>(), G, nq, (int)m, (int)k, oD.as<float>(), oI.as<int64_t>(), oC.as<int32_t>(), st));
    VQ_TRY(oD.finish(st));
    VQ_TRY(oI.finish(st));
    VQ_TRY(oC.finish(st));
    if (oD.host || oI.host || oC.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

int vq_l2sq(const float *q, const float *g, int64_t nq, int64_t n, int d, float *out, void *stream) {
    VQ_CHECK(q && g && out, VQ_ERR_INVALID, "NULL argument");
    VQ_TRY(check_device());
    cudaStream_t st = (cudaStream_t)stream;
    In a, b;
    Out o;
    VQ_TRY(a.bind(q, (size_t)nq * d * 4, st));
    VQ_TRY(b.bind(g, (size_t)n * d * 4, st));
    VQ_TRY(o.bind(out, (size_t)nq * n * 4, st));
    VQ_TRY(launch_l2sq_all(a.dev, b.dev, VQ_F32, nq, n, d, o.as<float>(), st));
    VQ_TRY(o.finish(st));
    if (o.host) VQ_CUDA(cudaStreamSynchronize(st));
    return VQ_OK;
}

}  // extern "C"
