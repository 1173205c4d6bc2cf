// vq_gemm_topk.cu -- kernels (1)+(2): query x gallery similarity GEMM on tcgen05/TMEM tiles fed
// by TMA, with a fused epilogue that turns each accumulator tile into coarse scores
// (l2sq: c = 2<q,g> - |g|^2; the -|g|^2/2 term rides in a 16-column augmentation k-block, see
// vq_kernels.h, so the accumulator IS c/2) and streams them into a per-query candidate list
// (threshold filter + warp-cooperative select), optionally max-pooling over the moments of one
// untrimmed video first.  The Q x N score matrix never exists in HBM.
//
// Replaces the arithmetic of faiss.IndexFlatL2.search (dataloader_baseline.py:1592-1597,
// :935-940, :1268-1273) and the labelling "ignore" rule (:1548) for the candidates it emits.
// The candidates are only a PRE-FILTER: vq_select.cu re-scores them with the canonical fp32
// chain, proves the top-k complete against the discard threshold tau emitted here, and sends
// unproven queries to the exact fallback.
//
// Two kernels share one epilogue:
//   gemm_topk_pair_kernel (bf16 single pass, dim <= 512): a CTA PAIR per work item
//     (cluster of 2, tcgen05 cta_group::2, UMMA M=256 N=256 K=16).  Each CTA keeps its 128 query
//     rows resident in shared memory for the whole gallery chunk and streams only its HALF of
//     every gallery tile (128 rows x 64 k, 16 KB per stage); the MMA reads both halves.  Per SM
//     that is 16 KB of L2->SM traffic per 128x256x64 MMA block instead of 48 KB.
//   gemm_topk_kernel (split-bf16 fp32 mode, or dim > 512): one CTA per work item, A and B
//     k-blocks streamed together.
//   warp 0     TMA producer: k-blocks of 64 bf16 -> 128B-swizzled smem
//   warp 1     MMA issuer (pair: leader CTA only), fp32 accumulators in TMEM
//              (2 x 256 columns, double buffered against the epilogue)
//   warps 2-5  epilogue: tcgen05.ld 32 lanes x 32 columns, one query row per thread
#include "vq_common.cuh"
#include "vq_kernels.h"

namespace vq {

static constexpr int BM = kTileM, BN = kTileN, BK = 64;
static constexpr int kEpiThreads = 128;
static constexpr int kThreads = 64 + kEpiThreads;
static constexpr int kOwnerSlots = 8;    // scratch rows per epilogue warp (owners handled per pass)
static constexpr int kOwnerStride = 36;  // floats per scratch row: 128 B + 16 B pad, so the eight rows cover all banks
static constexpr int kScrBytes = 4 * kOwnerSlots * kOwnerStride * 4;  // four epilogue warps
static constexpr int A_BYTES = BM * BK * 2;  // 16 KB
static constexpr int B_BYTES = BN * BK * 2;  // 32 KB

template <int NPASS>
struct SmemLayout {
    static constexpr int NPLANE = NPASS == 1 ? 1 : 2;
    static constexpr int STAGE_BYTES = NPLANE * (A_BYTES + B_BYTES);
    static constexpr int STAGES = NPASS == 1 ? 4 : 2;
    static constexpr int OFF_SCR = STAGES * STAGE_BYTES;           // float[4 warps][kOwnerSlots][kOwnerStride] owner scratch rows
    static constexpr int OFF_IGN = OFF_SCR + kScrBytes;            // int[BM][kMaxIgnore]
    static constexpr int OFF_BAR = OFF_IGN + BM * kMaxIgnore * 4;  // mbarriers
    static constexpr int NBAR = 2 * STAGES + 4;
    static constexpr int OFF_TMEM = OFF_BAR + NBAR * 8;
    static constexpr int TOTAL = OFF_TMEM + 16 + 1024;  // + alignment slack
};

// pair kernel: resident A (up to 8 k-blocks + the augmentation block) + B half-tile stages
struct PairLayout {
    static constexpr int KB_MAX = 8;                         // dim <= 512
    static constexpr int BH_BYTES = (BN / 2) * BK * 2;       // 16 KB: this CTA's half of a gallery tile k-block
    static constexpr int STAGES = 4;
    static constexpr int OFF_B = (KB_MAX + 1) * A_BYTES;     // 144 KB
    static constexpr int OFF_SCR = OFF_B + STAGES * BH_BYTES;
    static constexpr int OFF_IGN = OFF_SCR + kScrBytes;
    static constexpr int OFF_BAR = OFF_IGN + BM * kMaxIgnore * 4;
    // full[S] empty[S] tfull[2] tempty[2] afull aempty
    static constexpr int NBAR = 2 * STAGES + 6;
    static constexpr int OFF_TMEM = OFF_BAR + NBAR * 8;
    static constexpr int TOTAL = OFF_TMEM + 16 + 1024;
};
static_assert(PairLayout::TOTAL <= 227 * 1024, "pair kernel shared memory");

// Warp-cooperative compaction of one query row's candidate buffer.  Finds a threshold word T by
// bisection on the ordered score bits such that kprime <= #{score >= T}; with keep_max > kprime
// the bisection stops as soon as the count also fits keep_max (the intermediate compactions only
// need SOME valid threshold; they run ~3x fewer steps), with keep_max == kprime it runs to the
// exact k'-th largest and ties at T are kept in buffer (= column) order up to the quota, so the
// result is deterministic.  Returns the new count; *tau_out = a score such that everything
// discarded is <= tau.
template <int E>
__device__ __forceinline__ int compact_row(uint64_t *buf, int cnt, int kprime, int keep_max, int lane,
                                           float *tau_out) {
    uint32_t hi[E], lo[E];
    __syncwarp();  // the lanes' appends to this row are visible to the whole warp
#pragma unroll
    for (int i = 0; i < E; ++i) {
        int idx = i * 32 + lane;
        uint64_t e = idx < cnt ? buf[idx] : 0ull;
        hi[i] = (uint32_t)(e >> 32);
        lo[i] = (uint32_t)e;
    }
    __syncwarp();
    // Bisection of the VALUE interval [smallest, largest] of the listed scores (ordered-uint images) instead of the 32 key
    // bits: the scores of one row sit in a narrow band, so bit-by-bit descent spent ~10 rounds on the bits they all
    // share; the interval version lands in the [k', keep_max] window after 3-6 rounds and still ends at the exact k'-th
    // largest when it has to (hi_ - lo == 1).
    uint32_t mx = 0u, mn = 0xffffffffu;
#pragma unroll
    for (int i = 0; i < E; ++i) {
        const bool valid = (i * 32 + lane) < cnt;
        mx = valid && hi[i] > mx ? hi[i] : mx;
        mn = valid && hi[i] < mn ? hi[i] : mn;
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    mn = __reduce_min_sync(0xffffffffu, mn);
    uint32_t T = mn;      // #{hi >= T} >= kprime holds throughout
    uint32_t hi_ = mx == 0xffffffffu ? mx : mx + 1u;  // #{hi >= hi_} < kprime (NaN-pattern keys excepted: they are kept)
    int cT = cnt;         // #{hi >= T}; the padding slots hold 0 and are never >= a midpoint
#pragma unroll 1
    while (hi_ - T > 1u && cT > keep_max) {
        const uint32_t mid = T + ((hi_ - T) >> 1);
        int c = 0;
#pragma unroll
        for (int i = 0; i < E; ++i) c += (hi[i] >= mid) ? 1 : 0;
        c = __reduce_add_sync(0xffffffffu, c);
        if (c >= kprime) {
            T = mid;
            cT = c;
        } else {
            hi_ = mid;
        }
    }
    const uint32_t lt_mask = (1u << lane) - 1u;
    int pos = 0;
    if (cT <= keep_max) {  // keep everything >= T
#pragma unroll
        for (int i = 0; i < E; ++i) {
            const bool keep = (i * 32 + lane) < cnt && hi[i] >= T;
            const uint32_t km = __ballot_sync(0xffffffffu, keep);
            if (keep) buf[pos + __popc(km & lt_mask)] = ((uint64_t)hi[i] << 32) | lo[i];
            pos += __popc(km);
        }
    } else {  // exact: T is the k'-th largest; ties at T by buffer order up to the quota
        int gt = 0;
#pragma unroll
        for (int i = 0; i < E; ++i) gt += (hi[i] > T) ? 1 : 0;
        gt = __reduce_add_sync(0xffffffffu, gt);
        const int quota_eq = kprime - gt;
        int eqseen = 0;
#pragma unroll
        for (int i = 0; i < E; ++i) {
            const bool valid = (i * 32 + lane) < cnt;
            const bool isgt = valid && hi[i] > T;
            const bool iseq = valid && hi[i] == T;
            const uint32_t eqm = __ballot_sync(0xffffffffu, iseq);
            const bool keep = isgt || (iseq && (eqseen + __popc(eqm & lt_mask)) < quota_eq);
            const uint32_t km = __ballot_sync(0xffffffffu, keep);
            if (keep) buf[pos + __popc(km & lt_mask)] = ((uint64_t)hi[i] << 32) | lo[i];
            pos += __popc(km);
            eqseen += __popc(eqm);
        }
    }
    __syncwarp();
    *tau_out = ordered_to_f32(T);
    return pos;
}

// Out-of-line dispatcher: compactions are rare once the threshold has warmed up, and keeping their
// three unrolled bodies out of the epilogue's hot loop keeps that loop inside the instruction cache.
__device__ __noinline__ int compact_row_any(uint64_t *buf, int cnt, int kprime, int keep_max, float *tau_out) {
    const int lane = threadIdx.x & 31;
    if (cnt <= 320) return compact_row<10>(buf, cnt, kprime, keep_max, lane, tau_out);
    if (cnt <= 576) return compact_row<18>(buf, cnt, kprime, keep_max, lane, tau_out);
    return compact_row<kCandCap / 32>(buf, cnt, kprime, keep_max, lane, tau_out);
}

// ---------------------------------------------------------------------------------------------
// Epilogue role (warps 2-5 of either kernel).  Segments [s0, s1) of this CTA (pair).
// Thread = accumulator row = query.  Per 32 columns: tcgen05.ld, max of the 32 values, ONE compare
// against the row's threshold.  Rows that hold a candidate are handled by the whole warp: the
// row's values are transposed through shared memory so that lane i tests column i, and the hits
// are appended to the row's list with one ballot.
// ---------------------------------------------------------------------------------------------
template <bool SEGMAX, bool PAIR>
__device__ __forceinline__ void epilogue_role(const GemmTopkParams &p, float *scr, int *ign, uint64_t *tfull_bar,
                                              uint64_t *tempty_bar, uint32_t tempty_leader, uint32_t tmem_base,
                                              int s0, int s1, int rank) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quarter = warp & 3;            // TMEM lane quarter this warp may read
    const int row = quarter * 32 + lane;     // accumulator row == query row in the block
    const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
    const uint32_t lt_mask = (1u << lane) - 1u;
    float *wscr = scr + quarter * (kOwnerSlots * kOwnerStride);  // this warp's owner scratch rows
    int as = 0;
    uint32_t aphase = 0;
    for (int sg_i = s0; sg_i < s1; ++sg_i) {
        // PAIR: segment of a pair of query blocks, this CTA takes block 2*group + rank
        const int grp = p.seg_group[sg_i];
        const int qb = PAIR ? 2 * grp + rank : grp;
        const int u = PAIR ? 2 * sg_i + rank : sg_i;
        const int64_t cs = p.seg_begin[sg_i], ce = p.seg_end[sg_i];
        const int ntiles = (int)((ce - cs + BN - 1) / BN);
        const int64_t q = (int64_t)qb * BM + row;
        const bool qvalid = q < p.nq;
        uint64_t *buf = p.cand + ((int64_t)u * BM + row) * kCandCap;
        uint64_t *wbuf = p.cand + ((int64_t)u * BM + quarter * 32) * kCandCap;
        // tau: everything discarded so far has coarse score <= tau; tauh = tau / 2 is what the
        // accumulator (= score / 2) is compared with
        float tau = qvalid ? -INFINITY : INFINITY;
        int cnt = 0;
        int cum = p.seg_pred[sg_i] == -1 ? 1 : 0;  // this unit's list covers every earlier gallery chunk of the row as well
        {   // CONTINUE the predecessor segment (same query rows, previous gallery chunk) if it has been published
            // (tau NaN = not yet): take over its final threshold AND its final list, so the row's stream goes on as
            // if the chunk boundary were not there -- the next compaction tightens tau after k' new candidates instead
            // of the list having to refill from empty -- and the select kernel reads ONE list per query.
            const int pr = p.seg_pred[sg_i];
            if (pr >= 0 && qvalid) {
                const int64_t pu = (int64_t)(PAIR ? 2 * pr + rank : pr) * BM + row;
                const float t0 = *(const volatile float *)(p.cand_tau + pu);
                if (t0 == t0) tau = t0;  // everything the predecessor discarded is <= its final threshold
                // its list is taken over only if it covers all earlier chunks itself: a list that continued a
                // non-cumulative one would be read by the select kernel TOGETHER with the lists it contains
                if (t0 == t0 && *(const volatile int32_t *)(p.cand_cum + pu) != 0) {
                    __threadfence();  // the list, its count and its flag were written before tau was published
                    cum = 1;
                    const int pc = min(*(const volatile int32_t *)(p.cand_cnt + pu), p.kprime);
                    const uint64_t *src = p.cand + pu * kCandCap;
                    for (int i0 = 0; i0 < pc; i0 += 8) {
                        uint64_t e[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) e[i] = i0 + i < pc ? __ldcg(src + i0 + i) : 0ull;
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            if (i0 + i < pc) buf[i0 + i] = e[i];
                    }
                    cnt = pc;
                }
            }
        }
        float tauh = 0.5f * tau;
        // group-max state (SEGMAX), in accumulator (half-score) units
        int cur_seg = -2;
        float cur_maxh = -INFINITY;
        float cur_sech = -INFINITY;  // runner-up of the current video among the elements looked at (p.seg_amb)
        uint32_t cur_arg = 0;
        bool grp_masked = false;
        const float band = (SEGMAX && p.seg_amb && qvalid) ? p.row_eps[q] : 0.f;  // accumulator units: acc1 - acc2 > eps proves the order
        __syncwarp();
        if (p.n_ign > 0) {
            for (int t = 0; t < p.n_ign; ++t) ign[row * kMaxIgnore + t] = qvalid ? p.q_ignore[q * p.n_ign + t] : -1;
        }
        __syncwarp();  // the warp-cooperative path reads other rows' ignore lists
        auto append = [&](float c, uint32_t j) {
            buf[cnt] = make_key_desc(c, j);
            ++cnt;
        };
        auto is_ignored = [&](int r, int v) {
            bool m = false;
            for (int z = 0; z < p.n_ign; ++z) m |= (v == ign[r * kMaxIgnore + z]);
            return m;
        };
        // E = buffer registers per lane a compaction scans: the trigger bounds the fill level
        auto compact_rows = [&](uint32_t need, int keep_max) {
            while (need) {
                const int src = __ffs(need) - 1;
                need &= need - 1;
                const int c_r = __shfl_sync(0xffffffffu, cnt, src);
                uint64_t *rb = wbuf + (int64_t)src * kCandCap;
                float t_r;
                const int nc = compact_row_any(rb, c_r, p.kprime, keep_max, &t_r);
                if (lane == src) {
                    cnt = nc;
                    tau = fmaxf(tau, t_r);
                    tauh = 0.5f * tau;
                }
            }
        };
        // Intermediate compactions are cheap (few registers, early-exit bisection) and fire
        // often so that tau stays close to the k'-th best seen so far.  A 32-column step can add
        // at most 33 entries, hence the cap.
        const int trigger = p.trigger > 0 ? min(p.trigger, kCandCap - 64) : min(2 * p.kprime, kCandCap - 64);
        const int keep_max = min(p.kprime + p.kprime / 2, trigger - 32);
        for (int t = 0; t < ntiles; ++t) {
            const int64_t j0 = cs + (int64_t)t * BN;
            const uint32_t jbase = (uint32_t)j0;
            const int ncols = (int)min((int64_t)BN, ce - j0);  // < BN only in the last tile of a chunk
            // SEGMAX: the video ids of the tile's columns (lane = column inside a 32-column step), fetched BEFORE the wait
            // for the accumulator -- inside the step the load's L2 latency sat on the critical path of every step (ncu:
            // 19 % of the kernel's stall samples on the shuffle that consumes it)
            int sg_pre[SEGMAX ? BN / 32 : 1];
            if (SEGMAX) {
#pragma unroll
                for (int ch = 0; ch < BN / 32; ++ch)
                    sg_pre[ch] = ch * 32 + lane < ncols ? __ldg(p.col_seg + j0 + ch * 32 + lane) : 0x7fffffff;
            }
            mbar_wait(&tfull_bar[as], aphase);
            tc_fence_after();
            const uint32_t tacc = tmem_base + lane_addr + (uint32_t)(as * BN);
#pragma unroll 1
            for (int ch = 0; ch < BN / 32; ++ch) {
                uint32_t r[32];
                tmem_ld32(tacc + ch * 32, r);
                tmem_wait_ld();
                if (ch == BN / 32 - 1) {  // accumulator fully read: hand TMEM back to the MMA thread
                    tc_fence_before();
                    if (PAIR) mbar_arrive_cluster(tempty_leader + 8u * as);  // it waits on the leader's barrier
                    else mbar_arrive(&tempty_bar[as]);
                }
                if (p.store) {
                    // STORE mode (full-list evaluators): the coarse scores themselves, 2 * acc = 2<q,g> - |g|^2, go to
                    // store[q, j]; no candidate handling.  A thread writes its row's 32 columns as eight 16-byte stores
                    // (one full 128-byte line per row and step).
                    if (qvalid) {
                        float *dst = p.store + q * p.store_ld + j0 + ch * 32;
                        if (ncols == BN && ((p.store_ld & 3) == 0)) {
#pragma unroll
                            for (int i = 0; i < 32; i += 4)
                                *reinterpret_cast<float4 *>(dst + i) =
                                    make_float4(2.0f * __uint_as_float(r[i]), 2.0f * __uint_as_float(r[i + 1]),
                                                2.0f * __uint_as_float(r[i + 2]), 2.0f * __uint_as_float(r[i + 3]));
                        } else {
#pragma unroll
                            for (int i = 0; i < 32; ++i)
                                if (ch * 32 + i < ncols) dst[i] = 2.0f * __uint_as_float(r[i]);
                        }
                    }
                    continue;
                }
                if (ncols < BN) {  // columns past the chunk end (zero-filled operands) can never qualify
                    if (ch * 32 >= ncols) continue;
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (ch * 32 + i >= ncols) r[i] = 0xff800000u;
                }
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
                // max of 32 (3-input max tree)
                float m[11];
#pragma unroll
                for (int i = 0; i < 10; ++i) m[i] = fmaxf(fmaxf(v[3 * i], v[3 * i + 1]), v[3 * i + 2]);
                m[10] = fmaxf(v[30], v[31]);
                const float m32 = fmaxf(fmaxf(fmaxf(fmaxf(m[0], m[1]), m[2]), fmaxf(fmaxf(m[3], m[4]), m[5])),
                                        fmaxf(fmaxf(fmaxf(m[6], m[7]), m[8]), fmaxf(m[9], m[10])));
                if (SEGMAX) {
                    // cur_maxh tracks the best element ABOVE tau of the current video; elements
                    // <= tau can never make a video a candidate and are skipped wholesale.
                    // Video boundaries inside these 32 columns are common to all rows: one coalesced load of the
                    // columns' video ids and a ballot give them as a bit mask (bit i: column i opens a new video).
                    const int colL = ch * 32 + lane;
                    int sgL = sg_pre[0];  // (the step loop is not unrolled: pick the register by a select chain)
#pragma unroll
                    for (int c2 = 1; c2 < BN / 32; ++c2) sgL = ch == c2 ? sg_pre[c2] : sgL;
                    int sgP = __shfl_up_sync(0xffffffffu, sgL, 1);
                    if (lane == 0) sgP = cur_seg;
                    const uint32_t bnd = __ballot_sync(0xffffffffu, colL < ncols && sgL != sgP);
                    const int last = min(31, ncols - 1 - ch * 32);
                    const int seg_first = __shfl_sync(0xffffffffu, sgL, 0), seg_last = __shfl_sync(0xffffffffu, sgL, last);
                    auto flush_to = [&](int sg, int col) {
                        if (cur_maxh > tauh && !grp_masked) {
                            append(2.0f * cur_maxh, cur_arg);
                            // The arg-max is the video's exact best moment if nothing else can reach it: the runner-up
                            // of the inspected elements and every skipped element (all <= tau at the time, tau only
                            // grows) lie more than the error band below.  Otherwise the video is marked.
                            if (p.seg_amb && qvalid) {
                                const float lim = cur_maxh - band;
                                if (!(cur_sech < lim && tauh < lim))
                                    atomicOr(p.seg_amb + q * p.amb_words + (cur_seg >> 5), 1u << (cur_seg & 31));
                            }
                        }
                        cur_seg = sg;
                        cur_maxh = -INFINITY;
                        cur_sech = -INFINITY;
                        grp_masked = p.n_ign > 0 ? is_ignored(row, __ldg(p.col_video + j0 + col)) : false;
                    };
                    if ((bnd & ~1u) == 0u) {  // warp-uniform: the 32 columns belong to one video
                        if (bnd & 1u) flush_to(seg_first, ch * 32);
                        if (m32 > tauh && m32 > cur_maxh) {  // the step's maximum, first column that attains it
                            // Through the max tree instead of a pass over the 32 values (the branch runs for the whole warp
                            // whenever one row qualifies -- about every second step; ncu: 38 % of the kernel's
                            // instructions were the two 32-element passes): first group of three that attains m32, first
                            // element inside it; runner-up = best of the other groups and the group's other elements (a
                            // duplicate of m32 counts).
                            int gsel = 10;
#pragma unroll
                            for (int gi = 9; gi >= 0; --gi) gsel = m[gi] == m32 ? gi : gsel;
                            float e0 = v[30], e1 = v[31], e2 = -INFINITY, others = -INFINITY;
#pragma unroll
                            for (int gi = 0; gi < 10; ++gi) {
                                const bool me = gi == gsel;
                                e0 = me ? v[3 * gi] : e0;
                                e1 = me ? v[3 * gi + 1] : e1;
                                e2 = me ? v[3 * gi + 2] : e2;
                                others = fmaxf(others, me ? -INFINITY : m[gi]);
                            }
                            others = fmaxf(others, gsel == 10 ? -INFINITY : m[10]);
                            const int sub = e0 == m32 ? 0 : (e1 == m32 ? 1 : 2);
                            if (p.seg_amb) {
                                const float r2 = sub == 0 ? fmaxf(e1, e2) : (sub == 1 ? fmaxf(e0, e2) : fmaxf(e0, e1));
                                cur_sech = fmaxf(fmaxf(cur_sech, cur_maxh), fmaxf(others, r2));
                            }
                            cur_maxh = m32;
                            cur_arg = jbase + (uint32_t)(ch * 32 + 3 * gsel + sub);
                        } else if (m32 > tauh) {
                            cur_sech = fmaxf(cur_sech, m32);
                        }
                    } else if (m32 > tauh) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) {
                            const int col = ch * 32 + i;
                            if (col < ncols) {  // warp-uniform
                                if ((bnd >> i) & 1u) flush_to(__ldg(p.col_seg + j0 + col), col);
                                if (v[i] > tauh && v[i] > cur_maxh) {
                                    cur_sech = fmaxf(cur_sech, cur_maxh);
                                    cur_maxh = v[i];
                                    cur_arg = jbase + col;
                                } else {
                                    cur_sech = fmaxf(cur_sech, v[i]);
                                }
                            }
                        }
                    } else {
                        flush_to(seg_last, ch * 32 + last);  // boundaries, but nothing above tau in this row
                    }
                    const uint32_t need = __ballot_sync(0xffffffffu, cnt > trigger);
                    if (need) compact_rows(need, keep_max);
                } else {
                    uint32_t bal = __ballot_sync(0xffffffffu, m32 > tauh);
                    if (p.dbg_no_insert) bal = 0u;  // timing probe only (results are wrong): the kernel without candidate handling
                    if (bal) {  // warp-uniform
                        // Rows that hold a candidate in these 32 columns ("owners") handle themselves, all at once: an
                        // owner parks its 32 values in a scratch row of its own, forms the hit mask from its registers
                        // (straight-line compares) and appends the hits to its own list, reading them back by index.
                        // No cross-lane traffic, no barrier, no per-row serialisation: a step with events costs the
                        // same ~120 instructions whether one row or eight rows hit (the first version passed one row at
                        // a time through the warp -- transpose, shuffles, ballot: ~500 cycles per row on a warp that is
                        // alone on its scheduler; ncu source view, profiles/r02_ncu_gemm_topk_n8shape_source.txt).
                        const bool owner = (bal >> lane) & 1u;
                        const int pos = __popc(bal & lt_mask);  // this owner's ordinal
                        const int nown = __popc(bal);
                        for (int base = 0; base < nown; base += kOwnerSlots) {  // warp-uniform trip count
                            if (owner && pos >= base && pos < base + kOwnerSlots) {
                                float *my = wscr + (pos - base) * kOwnerStride;
#pragma unroll
                                for (int i = 0; i < 8; ++i)
                                    *reinterpret_cast<float4 *>(my + 4 * i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                                uint32_t hm = 0u;
#pragma unroll
                                for (int i = 0; i < 32; ++i) hm |= v[i] > tauh ? (1u << i) : 0u;
                                const uint32_t jb = jbase + (uint32_t)(ch * 32);
                                while (hm) {
                                    const int i = __ffs(hm) - 1;
                                    hm &= hm - 1;
                                    const uint32_t j = jb + (uint32_t)i;
                                    if (p.n_ign > 0 && is_ignored(row, __ldg(p.col_video + j))) continue;
                                    buf[cnt] = make_key_desc(2.0f * my[i], j);
                                    ++cnt;
                                }
                            }
                            if (nown > kOwnerSlots) __syncwarp();  // the scratch rows change hands between passes
                        }
                        const uint32_t need = __ballot_sync(0xffffffffu, cnt > trigger);
                        if (need) compact_rows(need, keep_max);
                    }
                }
            }
            if (++as == 2) {
                as = 0;
                aphase ^= 1u;
            }
        }
        if (SEGMAX) {
            if (cur_maxh > tauh && !grp_masked) {
                append(2.0f * cur_maxh, cur_arg);
                if (p.seg_amb && qvalid) {
                    const float lim = cur_maxh - band;
                    if (!(cur_sech < lim && tauh < lim)) atomicOr(p.seg_amb + q * p.amb_words + (cur_seg >> 5), 1u << (cur_seg & 31));
                }
            }
        }
        __syncwarp();
        // final compaction of this unit to exactly <= k' entries per row
        compact_rows(__ballot_sync(0xffffffffu, cnt > p.kprime), p.kprime);
        p.cand_cnt[(int64_t)u * BM + row] = qvalid ? cnt : 0;
        p.cand_cum[(int64_t)u * BM + row] = cum;
        __syncwarp();     // the row's entries were written by the whole warp (appends, compactions) ...
        __threadfence();  // ... and are visible device-wide before the threshold that publishes them
        *(volatile float *)(p.cand_tau + (int64_t)u * BM + row) = tau;
    }
}

template <int NPASS, bool SEGMAX>
__global__ void __launch_bounds__(kThreads, 1)
gemm_topk_kernel(const __grid_constant__ CUtensorMap tm_a_hi, const __grid_constant__ CUtensorMap tm_a_lo,
                 const __grid_constant__ CUtensorMap tm_b_hi, const __grid_constant__ CUtensorMap tm_b_lo,
                 const GemmTopkParams p) {
    using L = SmemLayout<NPASS>;
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment (SWIZZLE_128B) by offset arithmetic: keeps the pointer in the shared window (LDS/STS)
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    float *scr = (float *)(smem + L::OFF_SCR);
    int *ign = (int *)(smem + L::OFF_IGN);
    uint64_t *full_bar = (uint64_t *)(smem + L::OFF_BAR);
    uint64_t *empty_bar = full_bar + L::STAGES;
    uint64_t *tfull_bar = empty_bar + L::STAGES;
    uint64_t *tempty_bar = tfull_bar + 2;
    uint32_t *tmem_slot = (uint32_t *)(smem + L::OFF_TMEM);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int seg0 = p.cta_seg_start[blockIdx.x], seg1 = p.cta_seg_start[blockIdx.x + 1];
    const int KB = p.kb_count;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_a_hi);
        tma_prefetch_desc(&tm_b_hi);
        if (NPASS > 1) {
            tma_prefetch_desc(&tm_a_lo);
            tma_prefetch_desc(&tm_b_lo);
        }
        for (int s = 0; s < L::STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tfull_bar[s], 1);
            mbar_init(&tempty_bar[s], kEpiThreads);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================ TMA producer =================================
        // The whole warp runs the loops (warp-uniform control flow keeps the TMA operands in
        // uniform registers); one elected lane issues.
        int stage = 0;
        uint32_t phase = 0;
        for (int sg_i = seg0; sg_i < seg1; ++sg_i) {
            const int qb = p.seg_group[sg_i];
            const int64_t cs = p.seg_begin[sg_i], ce = p.seg_end[sg_i];
            const int ntiles = (int)((ce - cs + BN - 1) / BN);
            const int q0 = qb * BM;
            for (int t = 0; t < ntiles; ++t) {
                const int row0 = (int)(cs + (int64_t)t * BN);
                for (int kb = 0; kb <= KB; ++kb) {  // kb == KB: the augmentation block (zero-filled past 16 columns)
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    uint8_t *sa = smem + stage * L::STAGE_BYTES;
                    if (elect_one()) {
                        mbar_expect_tx(&full_bar[stage], L::STAGE_BYTES);
                        tma_load_2d(&tm_a_hi, &full_bar[stage], sa, kb * BK, q0, kEvictLast);
                        if (NPASS > 1) {
                            tma_load_2d(&tm_a_lo, &full_bar[stage], sa + A_BYTES, kb * BK, q0, kEvictLast);
                            tma_load_2d(&tm_b_hi, &full_bar[stage], sa + 2 * A_BYTES, kb * BK, row0, kEvictNormal);
                            tma_load_2d(&tm_b_lo, &full_bar[stage], sa + 2 * A_BYTES + B_BYTES, kb * BK, row0, kEvictNormal);
                        } else {
                            tma_load_2d(&tm_b_hi, &full_bar[stage], sa + A_BYTES, kb * BK, row0, kEvictNormal);
                        }
                    }
                    __syncwarp();
                    if (++stage == L::STAGES) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ===================================
        constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
        int stage = 0;
        uint32_t phase = 0;
        int as = 0;
        uint32_t aphase = 0;
        for (int sg_i = seg0; sg_i < seg1; ++sg_i) {
            const int64_t cs = p.seg_begin[sg_i], ce = p.seg_end[sg_i];
            const int ntiles = (int)((ce - cs + BN - 1) / BN);
            for (int t = 0; t < ntiles; ++t) {
                mbar_wait(&tempty_bar[as], aphase ^ 1u);
                tc_fence_after();
                const uint32_t tacc = tmem_base + (uint32_t)(as * BN);
                for (int kb = 0; kb <= KB; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * L::STAGE_BYTES);
                    if (elect_one()) {  // warp-uniform operands + elect.sync: no per-lane waterfall around the MMAs
                        const int nk = kb == KB ? kAugCols / 16 : BK / 16;  // augmentation block: one K=16 step, hi x hi only
                        if (NPASS == 1) {
                            const uint64_t ad = umma_desc_sw128(sa), bd = umma_desc_sw128(sa + A_BYTES);
                            for (int k = 0; k < nk; ++k) umma_bf16(tacc, ad + 2 * k, bd + 2 * k, idesc, (kb | k) != 0);
                        } else {
                            const uint64_t ah = umma_desc_sw128(sa), al = umma_desc_sw128(sa + A_BYTES);
                            const uint64_t bh = umma_desc_sw128(sa + 2 * A_BYTES);
                            const uint64_t bl = umma_desc_sw128(sa + 2 * A_BYTES + B_BYTES);
                            if (kb == KB) {
                                umma_bf16(tacc, ah, bh, idesc, 1);
                            } else {
#pragma unroll
                                for (int k = 0; k < BK / 16; ++k) {
                                    umma_bf16(tacc, al + 2 * k, bh + 2 * k, idesc, (kb | k) != 0);
                                    umma_bf16(tacc, ah + 2 * k, bl + 2 * k, idesc, 1);
                                    umma_bf16(tacc, ah + 2 * k, bh + 2 * k, idesc, 1);
                                }
                            }
                        }
                        umma_commit(&empty_bar[stage]);  // frees the smem stage when the MMAs retire
                        if (kb == KB) umma_commit(&tfull_bar[as]);
                    }
                    __syncwarp();
                    if (++stage == L::STAGES) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
                if (++as == 2) {
                    as = 0;
                    aphase ^= 1u;
                }
            }
        }
    } else {
        epilogue_role<SEGMAX, false>(p, scr, ign, tfull_bar, tempty_bar, 0u, tmem_base, seg0, seg1, 0);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------
// CTA-pair kernel (see the header comment).  Barriers: full[s] / afull / tempty[a] are waited on by
// the leader's MMA thread and live in the LEADER's shared memory (the peer reaches them through
// the cluster window); empty[s] / aempty / tfull[a] are signalled by the leader's multicast
// tcgen05.commit into BOTH CTAs and each CTA waits on its own copy.
// ---------------------------------------------------------------------------------------------
template <bool SEGMAX>
__global__ void __launch_bounds__(kThreads, 1)
gemm_topk_pair_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                      const GemmTopkParams p) {
    using L = PairLayout;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    float *scr = (float *)(smem + L::OFF_SCR);
    int *ign = (int *)(smem + L::OFF_IGN);
    uint64_t *full_bar = (uint64_t *)(smem + L::OFF_BAR);
    uint64_t *empty_bar = full_bar + L::STAGES;
    uint64_t *tfull_bar = empty_bar + L::STAGES;
    uint64_t *tempty_bar = tfull_bar + 2;
    uint64_t *afull_bar = tempty_bar + 2;
    uint64_t *aempty_bar = afull_bar + 1;
    uint32_t *tmem_slot = (uint32_t *)(smem + L::OFF_TMEM);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int pair = (int)blockIdx.x >> 1;
    const int seg0 = p.cta_seg_start[pair], seg1 = p.cta_seg_start[pair + 1];
    const int KB = p.kb_count;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_a);
        tma_prefetch_desc(&tm_b);
        for (int s = 0; s < L::STAGES; ++s) {
            mbar_init(&full_bar[s], 1);   // leader's producer arrives with the byte count of BOTH halves
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tfull_bar[s], 1);
            mbar_init(&tempty_bar[s], 2 * kEpiThreads);  // both CTAs' epilogue threads
        }
        mbar_init(afull_bar, 1);
        mbar_init(aempty_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc_cg2(tmem_slot, 512);
        tmem_relinquish_cg2();
    }
    tc_fence_before();
    cluster_sync_all();  // barriers of both CTAs initialised before any remote arrive / multicast commit
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================ TMA producer (both CTAs) =====================
        const uint32_t full_leader = mapa_u32(full_bar, 0);
        const uint32_t afull_leader = mapa_u32(afull_bar, 0);
        int stage = 0;
        uint32_t phase = 0, aphase = 0;
        for (int sg_i = seg0; sg_i < seg1; ++sg_i) {
            const int qb = 2 * p.seg_group[sg_i] + rank;
            const int64_t cs = p.seg_begin[sg_i], ce = p.seg_end[sg_i];
            const int ntiles = (int)((ce - cs + BN - 1) / BN);
            // this CTA's 128 query rows, resident for the whole segment
            mbar_wait(aempty_bar, aphase ^ 1u);
            aphase ^= 1u;
            if (elect_one()) {
                if (rank == 0) mbar_expect_tx(afull_bar, 2u * (uint32_t)(KB + 1) * A_BYTES);
                for (int kb = 0; kb <= KB; ++kb)  // kb == KB: the augmentation block
                    tma_load_2d_cg2(&tm_a, afull_leader, smem + kb * A_BYTES, kb * BK, qb * BM, kEvictLast);
            }
            __syncwarp();
            // this CTA's half of every gallery tile
            for (int t = 0; t < ntiles; ++t) {
                const int row0 = (int)(cs + (int64_t)t * BN) + rank * (BN / 2);
                for (int kb = 0; kb <= KB; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    if (elect_one()) {
                        if (rank == 0) mbar_expect_tx(&full_bar[stage], 2u * L::BH_BYTES);
                        tma_load_2d_cg2(&tm_b, full_leader + 8u * stage, smem + L::OFF_B + stage * L::BH_BYTES, kb * BK,
                                        row0, kEvictNormal);
                    }
                    __syncwarp();
                    if (++stage == L::STAGES) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
        // the leader's last multicast commit has landed in THIS CTA before anyone may leave
        mbar_wait(aempty_bar, aphase ^ 1u);
    } else if (warp == 1) {
        // ================================ MMA issuer (leader CTA) ======================
        if (rank == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(2 * BM, BN);
            int stage = 0;
            uint32_t phase = 0, aphase = 0;
            int as = 0;
            uint32_t tphase = 0;
            for (int sg_i = seg0; sg_i < seg1; ++sg_i) {
                const int64_t cs = p.seg_begin[sg_i], ce = p.seg_end[sg_i];
                const int ntiles = (int)((ce - cs + BN - 1) / BN);
                mbar_wait(afull_bar, aphase);
                aphase ^= 1u;
                tc_fence_after();
                for (int t = 0; t < ntiles; ++t) {
                    mbar_wait(&tempty_bar[as], tphase ^ 1u);
                    tc_fence_after();
                    const uint32_t tacc = tmem_base + (uint32_t)(as * BN);
                    for (int kb = 0; kb <= KB; ++kb) {
                        mbar_wait(&full_bar[stage], phase);
                        tc_fence_after();
                        const uint64_t ad = umma_desc_sw128(smem_u32(smem + kb * A_BYTES));
                        const uint64_t bd = umma_desc_sw128(smem_u32(smem + L::OFF_B + stage * L::BH_BYTES));
                        if (elect_one()) {  // warp-uniform operands + elect.sync: no per-lane waterfall around the MMAs
                            if (kb == KB) {  // augmentation block: one K=16 step
                                umma_bf16_cg2(tacc, ad, bd, idesc, 1);
                            } else {
#pragma unroll
                                for (int k = 0; k < BK / 16; ++k)
                                    umma_bf16_cg2(tacc, ad + 2 * k, bd + 2 * k, idesc, (kb | k) != 0);
                            }
                            umma_commit_cg2(&empty_bar[stage]);
                            if (kb == KB) {
                                umma_commit_cg2(&tfull_bar[as]);
                                if (t == ntiles - 1) umma_commit_cg2(aempty_bar);  // resident A may be replaced
                            }
                        }
                        __syncwarp();
                        if (++stage == L::STAGES) {
                            stage = 0;
                            phase ^= 1u;
                        }
                    }
                    if (++as == 2) {
                        as = 0;
                        tphase ^= 1u;
                    }
                }
            }
        }
    } else {
        epilogue_role<SEGMAX, true>(p, scr, ign, tfull_bar, tempty_bar, mapa_u32(tempty_bar, 0), tmem_base, seg0,
                                    seg1, rank);
    }

    __syncwarp();
    tc_fence_before();
    cluster_sync_all();  // no CTA leaves while its peer can still touch its barriers / shared memory
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_cg2(tmem_base, 512);
    }
}

template <int NPASS, bool SEGMAX>
static int launch_t(const CUtensorMap *maps, const GemmTopkParams &p, int grid, cudaStream_t st) {
    using L = SmemLayout<NPASS>;
    auto kern = gemm_topk_kernel<NPASS, SEGMAX>;
    VQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    kern<<<grid, kThreads, L::TOTAL, st>>>(maps[0], maps[1], maps[2], maps[3], p);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

int launch_gemm_topk(const CUtensorMap *maps, const GemmTopkParams &p, int npass, bool segmax, int grid,
                     cudaStream_t st) {
    if (npass == 1) return segmax ? launch_t<1, true>(maps, p, grid, st) : launch_t<1, false>(maps, p, grid, st);
    return segmax ? launch_t<3, true>(maps, p, grid, st) : launch_t<3, false>(maps, p, grid, st);
}

static int g_pair_slots = 0;

template <bool SEGMAX>
static int launch_pair_t(const CUtensorMap *maps, const GemmTopkParams &p, int n_pairs, cudaStream_t st) {
    auto kern = gemm_topk_pair_kernel<SEGMAX>;
    VQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PairLayout::TOTAL));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * n_pairs);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = PairLayout::TOTAL;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    VQ_CUDA(cudaLaunchKernelEx(&cfg, kern, maps[0], maps[1], p));
    return VQ_OK;
}

// co-resident CTA pairs (GPC shapes can strand an SM): the persistent grid is sized to this
int gemm_topk_pair_slots() {
    if (g_pair_slots) return g_pair_slots;
    auto kern = gemm_topk_pair_kernel<false>;
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PairLayout::TOTAL);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(sms / 2 * 2);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = PairLayout::TOTAL;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int mc = 0;
    if (cudaOccupancyMaxActiveClusters(&mc, kern, &cfg) != cudaSuccess || mc <= 0) {
        cudaGetLastError();
        mc = sms / 2;
    }
    g_pair_slots = mc < sms / 2 ? mc : sms / 2;
    return g_pair_slots;
}

// maps2 = {A box [64 x 128], B box [64 x 128]}; p.n_qblocks must be even; p.kb_count <= 8
int launch_gemm_topk_pair(const CUtensorMap *maps2, const GemmTopkParams &p, bool segmax, int n_pairs,
                          cudaStream_t st) {
    VQ_CHECK(p.n_qblocks % 2 == 0 && p.kb_count <= PairLayout::KB_MAX, VQ_ERR_INVALID, "pair kernel preconditions");
    return segmax ? launch_pair_t<true>(maps2, p, n_pairs, st) : launch_pair_t<false>(maps2, p, n_pairs, st);
}

}  // namespace vq
