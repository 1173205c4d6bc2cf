// vq_metrics.cu -- kernel (4): segmented AP / R@N reduction in rank-count form, plus temporal
// NMS and the multi-GPU top-k merge.
//
// ap_rows   one block per query over its canonical d2 row: the reference's per-candidate
//           labelling loop (dataloader_baseline.py:1537-1586: tp = label match :1553, ignore =
//           shared video id :1548) becomes two integer compares per element; AP
//           (evaluation_metric.add2dict :431-461 -> sklearn average_precision_score) and
//           R@30/50/100 (:475-481) are computed WITHOUT sorting the row:
//             rank(p) = #{survivors ahead of p};  for each tie group of positive scores
//             AP += (tps/P - prev_recall) * (tps / #{survivors with d2 <= t})      (float64)
//           with the last-ranked survivor forced positive (:434, AP only).
// ap_lists  sklearn.metrics.average_precision_score on a batch of (y_true, y_score) lists.
// nms1d     misc_utils/utils_models.py:nms_cpu, one warp per segment.
// topk_merge  k-way merge of per-shard top-k lists.
#include "vq_blocksel.cuh"
#include "vq_kernels.h"

namespace vq {

static constexpr int kApThreads = 256;
static constexpr int kApTab = 1024;  // buckets of the positives' key range (ap_rows_kernel, pass 2)

// exclusive prefix sum of a[0..n) in shared memory, in place; returns the total.  part: int[kApThreads+1]
__device__ __forceinline__ int block_exclusive_scan(int *a, int n, int *part) {
    const int tid = threadIdx.x;
    const int per = (n + kApThreads - 1) / kApThreads;
    const int b = tid * per, e = min(n, b + per);
    int s = 0;
    for (int i = b; i < e; ++i) s += a[i];
    __syncthreads();
    part[tid] = s;
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int i = 0; i < kApThreads; ++i) {
            int v = part[i];
            part[i] = run;
            run += v;
        }
        part[kApThreads] = run;
    }
    __syncthreads();
    int run = part[tid];
    for (int i = b; i < e; ++i) {
        int v = a[i];
        a[i] = run;
        run += v;
    }
    __syncthreads();
    return part[kApThreads];
}

__device__ __forceinline__ bool is_ignored(int v, const int *ign, int n_ign) {
    bool m = false;
    for (int z = 0; z < n_ign; ++z) m |= (v == ign[z]);
    return m;
}

// MOMENT (dataloader_baseline.py:1283-1330): survivors are additionally the moments kept by
// the per-video NMS (p.kept), tp = closest-hit label match AND IoU >= 0.5 (:1249-1255, :1315-1319), and the list
// handed to multiprocess_calculate (:386-402) is GROUPED: videos in the order of their best moment, kept moments
// in ranked order inside a video.  AP depends on the (score, tp) multiset only, except for the forced positive
// y_true[-1] (:389) = last kept moment of the last video; R@N counts positions in the grouped list.
// dynamic smem: u64 pos[pcap2] | int cntA[pcap2+1] | int cntE[pcap2] | int gstart[pcap2] | int part[257] | misc
template <bool MOMENT>
__global__ void __launch_bounds__(kApThreads) ap_rows_kernel(const ApRowsParams p, int pcap2) {
    extern __shared__ __align__(16) uint8_t sm[];
    uint64_t *pos = (uint64_t *)sm;
    int *cntA = (int *)(pos + pcap2);
    int *cntE = cntA + pcap2 + 1;
    int *gstart = cntE + pcap2;
    int *part = gstart + pcap2;
    int *misc = part + kApThreads + 1;  // [0]=npos [1]=nvalid, [8..16)=ignore list
    unsigned long long *last_key = (unsigned long long *)(misc + 16);  // [0] last survivor  [1] MOMENT: last video
    double *res = (double *)(last_key + 2);  // [0]=ap
    int *topn = (int *)(res + 1);            // [3] true positives within rank 30/50/100
    uint64_t *tv = (uint64_t *)(topn + 4);   // MOMENT: [128] best-moment keys of the first videos
    int *tcnt = (int *)(tv + 128);           // MOMENT: [128] kept moments per such video, then their offsets
    int *tab = tcnt + 128;                   // [kApTab + 1] bucket table of pass 2
    const int tid = threadIdx.x;
    const int64_t q = blockIdx.x;
    const float *row = p.d2 + q * p.n;
    const uint8_t *keptrow = MOMENT ? p.kept + q * p.n : nullptr;
    const int ql = p.q_label[q];
    auto is_tp = [&](int64_t j) { return p.g_label[j] == ql && (!MOMENT || p.g_iou[j] >= 0.5f); };

    if (tid < kMaxIgnore) misc[8 + tid] = (p.q_ignore && tid < p.n_ign) ? p.q_ignore[q * p.n_ign + tid] : -1;
    if (tid == 0) {
        misc[0] = 0;
        misc[1] = 0;
        misc[2] = 0;
        misc[4] = 0;
        last_key[0] = 0ull;
        last_key[1] = 0ull;
        topn[0] = topn[1] = topn[2] = 0;
    }
    for (int i = tid; i < pcap2; i += kApThreads) {
        pos[i] = ~0ull;
        cntA[i] = 0;
        cntE[i] = 0;
    }
    if (tid == 0) cntA[pcap2] = 0;
    __syncthreads();
    const int *ign = misc + 8;
    const int n_ign = p.q_ignore ? p.n_ign : 0;
    // the ignore list in registers for the two passes over the row (unused slots hold INT_MIN, never a video id): v is
    // ignored iff min over z of (v ^ ig[z]) == 0 -- no shared-memory read per element (ncu: the loop over the list in
    // shared memory was a quarter of the kernel's instructions and a third of its shared-memory wavefronts)
    uint32_t ig[kMaxIgnore];
#pragma unroll
    for (int z = 0; z < kMaxIgnore; ++z) ig[z] = z < n_ign ? (uint32_t)ign[z] : 0x80000000u;
    auto ignored_r = [&](int v) {
        uint32_t m = 0xffffffffu;
#pragma unroll
        for (int z = 0; z < kMaxIgnore; ++z) m = min(m, (uint32_t)v ^ ig[z]);
        return m == 0u;
    };

    // ---- pass 1: positives, survivor count, last-ranked survivor ----
    int nvalid_t = 0, nkept_t = 0;
    unsigned long long lk = 0ull;
    // Both passes stream the row four columns per thread with every load of a step issued up front (16-byte loads of
    // the score, video, label and IoU columns, 4 bytes of the kept mask): the block is alone with its row, so what
    // bounds it is the number of loads in flight, not bandwidth.  Rows that are not 16-byte aligned walk column by column.
    const bool vec4 = (p.n & 3) == 0 && ((((uintptr_t)p.d2) | ((uintptr_t)p.g_label) | ((uintptr_t)p.g_video) |
                                          (MOMENT ? ((uintptr_t)p.g_iou) : 0)) & 15) == 0 &&
                      (!MOMENT || (((uintptr_t)p.kept) & 3) == 0);
    auto scan_row = [&](auto &&body) {  // body(j, kept, d2, video, label, iou)
        if (vec4) {
            for (int64_t j = (int64_t)tid * 4; j < p.n; j += kApThreads * 4) {
                const uchar4 k4 = MOMENT ? *reinterpret_cast<const uchar4 *>(keptrow + j) : make_uchar4(1, 1, 1, 1);
                const float4 r4 = *reinterpret_cast<const float4 *>(row + j);
                const int4 v4 = n_ign ? __ldg(reinterpret_cast<const int4 *>(p.g_video + j)) : make_int4(0, 0, 0, 0);
                const int4 l4 = __ldg(reinterpret_cast<const int4 *>(p.g_label + j));
                const float4 i4 = MOMENT ? __ldg(reinterpret_cast<const float4 *>(p.g_iou + j)) : make_float4(0.f, 0.f, 0.f, 0.f);
                body(j, k4.x, r4.x, v4.x, l4.x, i4.x);
                body(j + 1, k4.y, r4.y, v4.y, l4.y, i4.y);
                body(j + 2, k4.z, r4.z, v4.z, l4.z, i4.z);
                body(j + 3, k4.w, r4.w, v4.w, l4.w, i4.w);
            }
        } else {
            for (int64_t j = tid; j < p.n; j += kApThreads)
                body(j, MOMENT ? keptrow[j] : (uint8_t)1, row[j], n_ign ? p.g_video[j] : 0, p.g_label[j],
                     MOMENT ? p.g_iou[j] : 0.f);
        }
    };
    scan_row([&](int64_t j, uint8_t kp, float d2v, int vid, int lab, float iou) {
        if (MOMENT) {
            if (!kp) return;
            ++nkept_t;
        }
        if (n_ign && ignored_r(vid)) return;
        ++nvalid_t;
        const unsigned long long key = make_key_asc(d2v, (uint32_t)j);
        lk = key > lk ? key : lk;
        if (lab == ql && (!MOMENT || iou >= 0.5f)) {
            const int slot = atomicAdd(&misc[0], 1);
            if (slot < pcap2) pos[slot] = key;
        }
    });
    atomicAdd(&misc[1], nvalid_t);
    if (MOMENT) atomicAdd(&misc[2], nkept_t);
    if (!MOMENT) atomicMax(last_key, lk);
    __syncthreads();
    int nvideos = 0;  // MOMENT: videos that are neither empty nor ignored
    uint64_t *vb = MOMENT ? p.vbest + q * p.nseg : nullptr;
    if (MOMENT) {
        if (tid == 0 && p.n_kept) p.n_kept[q] = misc[2];
        // last video of the grouped list = largest best-moment key; ignored videos leave the list (:387)
        unsigned long long lv = 0ull;
        int nv_t = 0;
        for (int64_t v = tid; v < p.nseg; v += kApThreads) {
            unsigned long long key = vb[v];
            if (key == ~0ull) continue;
            if (n_ign && is_ignored(p.g_video[(uint32_t)key], ign, n_ign)) {
                vb[v] = ~0ull;
                continue;
            }
            ++nv_t;
            lv = key > lv ? key : lv;
        }
        atomicMax(last_key + 1, lv);
        nvideos = block_sum(nv_t, part);
        __syncthreads();
        if (nvideos > 0) {  // its last kept moment
            const int64_t v = p.seg_of_row[(uint32_t)last_key[1]];
            unsigned long long lm = 0ull;
            for (int64_t j = p.seg_start[v] + tid; j < p.seg_start[v + 1]; j += kApThreads)
                if (keptrow[j]) {
                    const unsigned long long key = make_key_asc(row[j], (uint32_t)j);
                    lm = key > lm ? key : lm;
                }
            atomicMax(last_key, lm);
        }
        __syncthreads();
    }
    const int ntrue = misc[0];
    const int nvalid = misc[1];
    if (ntrue + 1 > pcap2) {  // uniform
        if (tid == 0) {
            atomicExch(p.overflow, 1);
            p.ap[q] = 0.0;
            p.recall[q * 3] = p.recall[q * 3 + 1] = p.recall[q * 3 + 2] = 0.0;
        }
        return;
    }
    if (nvalid == 0) {  // the reference would raise on y_true[-1] of an empty list
        if (tid == 0) {
            p.ap[q] = 0.0;
            p.recall[q * 3] = p.recall[q * 3 + 1] = p.recall[q * 3 + 2] = 0.0;
        }
        return;
    }
    // forced positive: the last survivor in ranked order (dataloader_baseline.py:434), AP only
    const uint64_t lastk = *last_key;
    const bool last_is_true = is_tp((uint32_t)lastk);
    int npos = ntrue;
    if (!last_is_true) {
        if (tid == 0) pos[ntrue] = lastk;
        npos = ntrue + 1;
    }
    __syncthreads();
    const int n2 = next_pow2(npos);
    block_bitonic_sort(pos, n2);  // ascending (d2, row): best first

    // ---- pass 2: bucket every survivor against the sorted positive d2 values ----
    // A table over the positives' key range (kApTab buckets, tab[b] = lower bound of the bucket's first key) cuts the
    // bisection from log2(npos) dependent shared-memory reads to the few entries of one bucket.
    const uint32_t kmin = (uint32_t)(pos[0] >> 32), kmax = (uint32_t)(pos[npos - 1] >> 32);
    int tshift = 0;
    while (((kmax - kmin) >> tshift) >= (uint32_t)kApTab) ++tshift;
    for (int b = tid; b <= kApTab; b += kApThreads) {
        const uint64_t first = (uint64_t)kmin + ((uint64_t)b << tshift);  // (may pass 2^32 in the last buckets)
        int lo = 0, hi = npos;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if ((uint64_t)(uint32_t)(pos[mid] >> 32) < first) lo = mid + 1; else hi = mid;
        }
        tab[b] = lo;
    }
    __syncthreads();
    scan_row([&](int64_t, uint8_t kp, float d2v, int vid, int, float) {
        if (MOMENT && !kp) return;
        if (n_ign && ignored_r(vid)) return;
        const uint32_t s = f32_to_ordered(d2v);
        int lo = 0, hi = npos;  // lower bound: first i with pos_d2[i] >= s
        if (s > kmax) {
            lo = npos;
        } else if (s >= kmin) {
            const uint32_t b = (s - kmin) >> tshift;
            lo = tab[b];
            hi = tab[b + 1];
        } else {
            hi = 0;
        }
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if ((uint32_t)(pos[mid] >> 32) < s) lo = mid + 1; else hi = mid;
        }
        atomicAdd(&cntA[lo], 1);
        if (lo < npos && (uint32_t)(pos[lo] >> 32) == s) atomicAdd(&cntE[lo], 1);
    });
    __syncthreads();
    block_exclusive_scan(cntA, npos + 1, part);  // cntA[i] = #{survivors with lower bound < i}
    // inclusive count LE(g) = cntA[g+1] = #{survivors with d2 <= pos_d2[g]} for a group start g

    // ---- AP: sequential over tie groups (float64, sklearn's order of operations) ----
    if (tid == 0) {
        const double P = (double)npos;
        double ap = 0.0, tps = 0.0, prev = 0.0;
        int i = 0;
        while (i < npos) {
            const int g = i;
            const uint32_t t = (uint32_t)(pos[g] >> 32);
            while (i < npos && (uint32_t)(pos[i] >> 32) == t) {
                gstart[i] = g;
                tps += 1.0;
                ++i;
            }
            const double jcount = (double)cntA[g + 1];
            const double recall = tps / P;
            const double precision = tps / jcount;
            ap += (recall - prev) * precision;
            prev = recall;
        }
        res[0] = ap;
    }
    __syncthreads();

    if (MOMENT) {
        // ---- R@N over the grouped list: the first <= 100 videos by best moment, their kept moments in order ----
        const int K = nvideos < 100 ? nvideos : 100;
        if (K > 0) {
            const uint64_t kth = block_kth_smallest(vb, (int)p.nseg, K, part);
            if (tid < 128) tv[tid] = ~0ull;
            if (tid == 0) misc[3] = 0;
            __syncthreads();
            for (int64_t v = tid; v < p.nseg; v += kApThreads)
                if (vb[v] <= kth) tv[atomicAdd(&misc[3], 1)] = vb[v];  // keys are distinct: exactly K of them
            __syncthreads();
            block_bitonic_sort(tv, 128);
            const int warp = tid >> 5, lane = tid & 31;
            for (int t = warp; t < K; t += kApThreads / 32) {
                const int64_t v = p.seg_of_row[(uint32_t)tv[t]];
                int c = 0;
                for (int64_t j = p.seg_start[v] + lane; j < p.seg_start[v + 1]; j += 32) c += keptrow[j] ? 1 : 0;
                c = __reduce_add_sync(0xffffffffu, c);
                if (lane == 0) tcnt[t] = c;
            }
            __syncthreads();
            if (tid == 0) {
                int run = 0;
                for (int t = 0; t < K; ++t) {
                    const int c = tcnt[t];
                    tcnt[t] = run;
                    run += c;
                }
            }
            __syncthreads();
            for (int t = warp; t < K; t += kApThreads / 32) {
                const int off = tcnt[t];
                if (off >= 100) continue;
                const int64_t v = p.seg_of_row[(uint32_t)tv[t]];
                const int64_t s0 = p.seg_start[v], s1 = p.seg_start[v + 1];
                for (int64_t jb = s0; jb < s1; jb += 32) {
                    const int64_t j = jb + lane;
                    const bool mine = j < s1 && keptrow[j] && is_tp(j);
                    if (!__any_sync(0xffffffffu, mine)) continue;
                    const unsigned long long key = mine ? make_key_asc(row[j], (uint32_t)j) : 0ull;
                    int ahead = 0;  // kept moments of this video ranked before mine
                    for (int64_t jj = s0; jj < s1; ++jj)
                        if (keptrow[jj]) ahead += make_key_asc(row[jj], (uint32_t)jj) < key ? 1 : 0;
                    if (mine) {
                        const int rank = off + ahead + 1;
                        if (rank <= 30) atomicAdd(&topn[0], 1);
                        if (rank <= 50) atomicAdd(&topn[1], 1);
                        if (rank <= 100) atomicAdd(&topn[2], 1);
                    }
                }
            }
        }
    }
    // ---- R@N: exact rank of every TRUE positive under the (d2, row) order ----
    // A positive whose d2 is shared by other survivors (float32 collisions are common: ~n^2 / 2^23 pairs per row)
    // needs #{survivors with equal d2 and row <= mine}.  Those positives are queued and each one is resolved by the
    // WHOLE block in one sweep of the row (the first version let the owning thread walk the row alone: up to n
    // dependent loads in one thread, ~1 ms for a 20 480-row gallery, hit by every second query).
    int *tie_n = misc + 4;
    for (int i = tid; !MOMENT && i < npos; i += kApThreads) {
        const uint64_t key = pos[i];
        if (!last_is_true && key == lastk) continue;  // the forced one is not a true positive
        const int g = gstart[i];
        if (cntE[g] == 1) {
            const int rank = cntA[g + 1];  // survivors strictly ahead by d2, plus itself
            if (rank <= 30) atomicAdd(&topn[0], 1);
            if (rank <= 50) atomicAdd(&topn[1], 1);
            if (rank <= 100) atomicAdd(&topn[2], 1);
        } else if (cntA[g + 1] - cntE[g] < 100) {  // it can still rank inside the first 100
            const int slot = atomicAdd(tie_n, 1);
            if (slot < 128) {
                tcnt[slot] = i;  // (tcnt / tv are free in the trimmed mode)
            } else {  // > 128 tied positives near the top (a gallery of exact duplicates): walk the row alone
                const uint32_t s = (uint32_t)(key >> 32), myrow = (uint32_t)key;
                int c = 0;
                for (int64_t j = 0; j <= myrow; ++j)
                    if (f32_to_ordered(row[j]) == s && !(n_ign && is_ignored(p.g_video[j], ign, n_ign))) ++c;
                const int rank = cntA[g + 1] - cntE[g] + c;
                if (rank <= 30) atomicAdd(&topn[0], 1);
                if (rank <= 50) atomicAdd(&topn[1], 1);
                if (rank <= 100) atomicAdd(&topn[2], 1);
            }
        }
    }
    __syncthreads();
    if (!MOMENT) {
        const int nt = min(*tie_n, 128);
        // more than 128 tied positives inside the first 100 ranks cannot happen: each takes a distinct rank
        for (int t = 0; t < nt; ++t) {
            const int i = tcnt[t];
            const uint64_t key = pos[i];
            const int g = gstart[i];
            const uint32_t s = (uint32_t)(key >> 32), myrow = (uint32_t)key;
            int c = 0;
            for (int64_t j = tid; j <= myrow; j += kApThreads) {
                if (f32_to_ordered(row[j]) != s) continue;
                if (n_ign && is_ignored(p.g_video[j], ign, n_ign)) continue;
                ++c;
            }
            c = block_sum(c, part);
            if (tid == 0) {
                const int rank = cntA[g + 1] - cntE[g] + c;
                if (rank <= 30) atomicAdd(&topn[0], 1);
                if (rank <= 50) atomicAdd(&topn[1], 1);
                if (rank <= 100) atomicAdd(&topn[2], 1);
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        p.ap[q] = res[0];
        const double npos_true = (double)ntrue + 1e-10;  // :476
        for (int r = 0; r < 3; ++r) p.recall[q * 3 + r] = (double)topn[r] * 1.0 / npos_true;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// ap_rows_tc: the trimmed / clip evaluation on TENSOR-CORE scores.  Input: the coarse score row c[j] = 2<q,g> - |g|^2
// of a query (gemm_topk STORE mode; |c - (|q|^2 - d2)| <= eps, the proven bound of vq_api_search.inl:error_model).
// AP and R@N need, for every positive p, the number of survivors ahead of it in the CANONICAL (d2, row) order:
//   * the positives' canonical d2 come from a gather of their rows (P of them, ~n / classes);
//   * a non-positive survivor whose coarse distance e = |q|^2 - c is more than eps away from every positive's d2 falls
//     into a definite gap between two positives: one counter increment, no canonical arithmetic;
//   * the others -- within eps of some positive, or within 2 eps of the largest coarse distance (the forced last
//     positive of :434 is the canonically LAST survivor) -- are queued, re-scored canonically (gather) and placed exactly.
// With class-structured features the queue holds ~1 % of the gallery; a query whose queue or positive list overflows
// is reported in `redo` and evaluated by the canonical all-pairs path.  Results are the same numbers as ap_rows_kernel.
// ---------------------------------------------------------------------------------------------------------------
#ifndef VQ_TC_THREADS
#define VQ_TC_THREADS 256
#endif
static constexpr int kTcThreads = VQ_TC_THREADS, kTcPos = 512, kTcAmb = 1024;  // 256 threads, 80 registers, 56 KB: three blocks = 24 warps per SM (128 threads x four blocks = 16 warps measured 1.24 ms against 1.15 ms at configs[0])
static constexpr int kTcRescoreWarps = 4;  // warps that own a re-scoring tile (the row scans use every warp)

template <typename T>
__global__ void __launch_bounds__(kTcThreads, kTcThreads == 256 ? 3 : 4) ap_rows_tc_kernel(const ApTcParams p) {
    extern __shared__ __align__(16) uint8_t sm[];
    uint64_t *pos = (uint64_t *)sm;                       // [kTcPos] exact keys of the positives (sorted), then + forced last
    uint64_t *ambk = pos + kTcPos;                        // [kTcAmb] exact keys of the queued survivors
    unsigned long long *lastp = (unsigned long long *)(ambk + kTcAmb);  // [2] key of the canonically last survivor (+ pad)
    float *qs = (float *)(lastp + 2);                     // [d rounded up to 4] 16-byte aligned: read as float4
    uint32_t *ambr = (uint32_t *)(qs + ((p.d + 3) & ~3)); // [kTcAmb] rows of the queued survivors
    int *gap = (int *)(ambr + kTcAmb);                    // [kTcPos + 2] non-positive survivors between positives i-1 and i (by d2)
    int *eqc = gap + kTcPos + 2;                          // [kTcPos + 1] ... of which with d2 equal to positive i's
    uint32_t *posr = (uint32_t *)eqc;                     // [kTcPos] rows of the positives until their keys exist (eqc is zeroed after)
    int *part = eqc + kTcPos + 1;                         // [40]
    int *misc = part + 40;                                // [0] npos [1] nvalid [2] namb [3] bad [4] cmin bits [5..7] topn  [8..16) ignore
    uint8_t *tile = (uint8_t *)((((uintptr_t)(misc + 16)) + 127) & ~(uintptr_t)127);  // [warps][kRescoreTileBytes]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t q = blockIdx.x;
    const float *row = p.coarse + q * p.ld;
    const int ql = p.q_label[q];
    const int n_ign = p.q_ignore ? p.n_ign : 0;
    const T *qrow = (const T *)p.q + q * p.d;
    for (int i = tid; i < p.d; i += kTcThreads) qs[i] = to_f32<T>(qrow[i]);
    if (tid < kMaxIgnore) misc[8 + tid] = tid < n_ign ? p.q_ignore[q * p.n_ign + tid] : -1;
    if (tid < 8) misc[tid] = tid == 4 ? -1 : 0;  // [4]: ordered-uint image of the smallest coarse score, starts at the maximum
    for (int i = tid; i < kTcPos + 2; i += kTcThreads) gap[i] = 0;
    __syncthreads();
    const int *ign = misc + 8;
    float qn = 0.f;
    for (int i = tid; i < p.d; i += kTcThreads) qn = fmaf(qs[i], qs[i], qn);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) qn += __shfl_xor_sync(0xffffffffu, qn, o);
    if (lane == 0) part[32 + warp] = __float_as_int(qn);
    __syncthreads();
    qn = 0.f;
    for (int w = 0; w < kTcThreads / 32; ++w) qn += __int_as_float(part[32 + w]);
    const float gmax2 = __uint_as_float(*p.gmax_norm2_bits);
    const float eps = p.eps_rel * sqrtf(qn) * sqrtf(gmax2) + p.eps_abs * (qn + gmax2);

    // Both passes stream the row four columns per thread with every load of a step issued up front (16-byte loads of
    // the score, video and label columns): one block is alone with its row, so what bounds a pass is the number of
    // loads in flight (the first version walked column by column with dependent loads: 4.4 ms for 4 096 queries).
    const bool vec4 = ((((uintptr_t)p.coarse) | ((uintptr_t)p.g_label) | ((uintptr_t)p.g_video)) & 15) == 0 && (p.ld & 3) == 0;
    const int64_t n4 = vec4 ? (p.n & ~(int64_t)3) : 0;
    auto scan_row = [&](auto &&body) {  // body(j, coarse, video, label)
        for (int64_t j = (int64_t)tid * 4; j < n4; j += kTcThreads * 8) {
            const int64_t j2 = j + kTcThreads * 4;
            const bool two = j2 < n4;
            const float4 r4 = *reinterpret_cast<const float4 *>(row + j);
            const int4 v4 = n_ign ? __ldg(reinterpret_cast<const int4 *>(p.g_video + j)) : make_int4(0, 0, 0, 0);
            const int4 l4 = __ldg(reinterpret_cast<const int4 *>(p.g_label + j));
            const float4 s4 = two ? *reinterpret_cast<const float4 *>(row + j2) : make_float4(0.f, 0.f, 0.f, 0.f);
            const int4 w4 = two && n_ign ? __ldg(reinterpret_cast<const int4 *>(p.g_video + j2)) : make_int4(0, 0, 0, 0);
            const int4 m4 = two ? __ldg(reinterpret_cast<const int4 *>(p.g_label + j2)) : make_int4(0, 0, 0, 0);
            body(j, r4.x, v4.x, l4.x);
            body(j + 1, r4.y, v4.y, l4.y);
            body(j + 2, r4.z, v4.z, l4.z);
            body(j + 3, r4.w, v4.w, l4.w);
            if (two) {
                body(j2, s4.x, w4.x, m4.x);
                body(j2 + 1, s4.y, w4.y, m4.y);
                body(j2 + 2, s4.z, w4.z, m4.z);
                body(j2 + 3, s4.w, w4.w, m4.w);
            }
        }
        for (int64_t j = n4 + tid; j < p.n; j += kTcThreads) body(j, row[j], n_ign ? p.g_video[j] : 0, p.g_label[j]);
    };
    // the ignore list in registers (as in ap_rows_kernel: min of xors, no shared-memory read per element -- the loop over
    // the shared list was a third of this kernel's stall samples and a quarter of its instructions)
    uint32_t ig[kMaxIgnore];
#pragma unroll
    for (int z = 0; z < kMaxIgnore; ++z) ig[z] = z < n_ign ? (uint32_t)ign[z] : 0x80000000u;
    auto ignored_r = [&](int v) {
        uint32_t m = 0xffffffffu;
#pragma unroll
        for (int z = 0; z < kMaxIgnore; ++z) m = min(m, (uint32_t)v ^ ig[z]);
        return m == 0u;
    };
    // ---- pass 1: survivors, positives, smallest coarse score ----
    int nvalid_t = 0;
    float cmin_t = INFINITY;
    scan_row([&](int64_t j, float c, int vid, int lab) {
        if (n_ign && ignored_r(vid)) return;
        ++nvalid_t;
        cmin_t = fminf(cmin_t, c);
        if (lab == ql) {
            const int slot = atomicAdd(&misc[0], 1);
            if (slot < kTcPos - 1) posr[slot] = (uint32_t)j;
        }
    });
    atomicAdd(&misc[1], nvalid_t);
    // float min through the ordered-uint image (coarse scores may be negative)
    atomicMin((unsigned int *)&misc[4], f32_to_ordered(cmin_t));
    __syncthreads();
    const int ntrue = misc[0], nvalid = misc[1];
    if (nvalid == 0) {  // the reference would raise on y_true[-1] of an empty list
        if (tid == 0) {
            p.ap[q] = 0.0;
            p.recall[q * 3] = p.recall[q * 3 + 1] = p.recall[q * 3 + 2] = 0.0;
        }
        return;
    }
    if (ntrue > kTcPos - 1) {  // uniform: the canonical path takes this query
        if (tid == 0) p.redo[atomicAdd(p.redo_count, 1)] = (int32_t)q;
        return;
    }
    const float cmin = ordered_to_f32((unsigned int)misc[4]);
    const float emax = qn - cmin;  // largest coarse distance of a survivor
    const bool tiled = ((size_t)p.d * sizeof(T)) % 128 == 0 && (((uintptr_t)p.g) & 15) == 0;
    uint8_t *wtile = tile + (warp % kTcRescoreWarps) * kRescoreTileBytes;
    auto rescore = [&](const uint32_t *rows, uint64_t *keys, int cnt) {  // canonical (d2, row) keys of gathered rows
        if (tiled) {
            for (int i0 = warp * 32; warp < kTcRescoreWarps && i0 < cnt; i0 += 32 * kTcRescoreWarps) {
                const int i = i0 + lane;
                const uint32_t r = rows[i < cnt ? i : i0];
                const float d2 = warp_canon_d2_rows<T>(qs, (const T *)p.g + (int64_t)r * p.d, p.d, wtile, lane);
                if (i < cnt) keys[i] = make_key_asc(d2, r);
            }
        } else {
            for (int i = tid; i < cnt; i += kTcThreads)
                keys[i] = make_key_asc(canon_d2_row<T>(qs, (const T *)p.g + (int64_t)rows[i] * p.d, p.d), rows[i]);
        }
    };
    // ---- the positives' canonical keys, sorted ----
    const int n2p = next_pow2(ntrue + 1);
    for (int i = ntrue + tid; i < n2p; i += kTcThreads) pos[i] = ~0ull;
    rescore(posr, pos, ntrue);
    __syncthreads();
    for (int i = tid; i < kTcPos + 1; i += kTcThreads) eqc[i] = 0;  // (posr's storage)
    block_bitonic_sort(pos, n2p);
    auto posd = [&](int i) { return (uint32_t)(pos[i] >> 32); };  // ordered image of positive i's d2

    // ---- pass 2: every non-positive survivor either falls into a definite gap or is queued ----
    int tail_t = 0;
    const uint32_t last_pos_o = ntrue > 0 ? posd(ntrue - 1) : 0u;
    scan_row([&](int64_t j, float c, int vid, int lab) {
        if (lab == ql) return;
        if (n_ign && ignored_r(vid)) return;
        const float e = qn - c;
        bool queue = e >= emax - 2.0f * eps;  // could be the canonically last survivor
        int lb = 0;
        if (!queue) {
            const uint32_t lo_o = f32_to_ordered(e - eps), hi_o = f32_to_ordered(e + eps);
            if (lo_o > last_pos_o) {  // behind every positive (the bulk of a class-structured gallery): no search, no atomic
                ++tail_t;
                return;
            }
            int lo = 0, hi = ntrue;  // first positive with d2 >= e - eps
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (posd(mid) < lo_o) lo = mid + 1; else hi = mid;
            }
            lb = lo;
            queue = lb < ntrue && posd(lb) <= hi_o;  // a positive inside [e - eps, e + eps]
        }
        if (!queue) {
            atomicAdd(&gap[lb], 1);
        } else {
            const int slot = atomicAdd(&misc[2], 1);
            if (slot < kTcAmb) ambr[slot] = (uint32_t)j;
        }
    });
    if (tail_t) atomicAdd(&gap[ntrue], tail_t);
    __syncthreads();
    const int namb = misc[2];
    if (namb > kTcAmb) {  // uniform
        if (tid == 0) p.redo[atomicAdd(p.redo_count, 1)] = (int32_t)q;
        return;
    }
    rescore(ambr, ambk, namb);
    __syncthreads();
    // ---- the last survivor in canonical order (forced positive, AP only) ----
    unsigned long long lk = 0ull;
    for (int i = tid; i < namb; i += kTcThreads) lk = ambk[i] > lk ? ambk[i] : lk;
    if (tid == 0) *lastp = ntrue > 0 ? pos[ntrue - 1] : 0ull;
    __syncthreads();
    atomicMax(lastp, lk);
    __syncthreads();
    const uint64_t lastk = *lastp;
    const bool last_is_true = ntrue > 0 && lastk == pos[ntrue - 1];
    int npos = ntrue;
    if (!last_is_true) {
        if (tid == 0) pos[ntrue] = lastk;  // the largest key: the list stays sorted
        npos = ntrue + 1;
    }
    __syncthreads();
    // ---- queued survivors: exact place among the positives (by d2), ties noted ----
    for (int i = tid; i < namb; i += kTcThreads) {
        const uint64_t k = ambk[i];
        if (!last_is_true && k == lastk) continue;  // promoted to a positive
        const uint32_t s = (uint32_t)(k >> 32);
        int lo = 0, hi = npos;  // first positive with d2 >= mine
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (posd(mid) < s) lo = mid + 1; else hi = mid;
        }
        atomicAdd(&gap[lo], 1);
        if (lo < npos && posd(lo) == s) atomicAdd(&eqc[lo], 1);
    }
    __syncthreads();
    // inclusive scan: gap[i] = non-positive survivors with d2 < positive i's d2, or equal to it when i starts its tie group
    if (warp == 0) {
        int run = 0;
        for (int c0 = 0; c0 <= npos; c0 += 32) {
            const int v = c0 + lane <= npos ? gap[c0 + lane] : 0;
            int inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += u;
            }
            if (c0 + lane <= npos) gap[c0 + lane] = run + inc;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    __syncthreads();
    // ---- AP over tie groups (float64, sklearn's order of operations): the groups' terms in parallel -- the divisions --,
    // then ONE thread adds them up in list order (the sum is what has an order; a non-final entry of a group adds +0.0) ----
    double *dr = (double *)ambr;   // recall steps: the queued rows are no longer needed ([kTcPos] doubles = kTcAmb * 4 bytes)
    double *pr = (double *)tile;   // precisions: neither are the re-scoring tiles
    for (int i = tid; i < npos; i += kTcThreads) {
        const uint32_t t = posd(i);
        double d_ = 0.0, p_ = 0.0;
        if (i == npos - 1 || posd(i + 1) != t) {  // last positive of its tie group
            int f = i;
            while (f > 0 && posd(f - 1) == t) --f;
            const double P = (double)npos, tps = (double)(i + 1);
            const double jcount = (double)gap[f] + tps;  // survivors with d2 <= t: non-positives incl. ties + positives so far
            d_ = tps / P - (f > 0 ? (double)f / P : 0.0);
            p_ = tps / jcount;
        }
        dr[i] = d_;
        pr[i] = p_;
    }
    __syncthreads();
    double apv = 0.0;
    if (tid == 0)  // ap += (recall - prev) * precision, contracted to one fma per group as in ap_rows_kernel: same bits
        for (int i = 0; i < npos; ++i) apv = fma(dr[i], pr[i], apv);
    // ---- R@N: exact rank of every TRUE positive under the (d2, row) order ----
    for (int i = tid; i < ntrue; i += kTcThreads) {
        const uint32_t t = posd(i);
        int f = i;
        while (f > 0 && posd(f - 1) == t) --f;
        int ahead = gap[f] - eqc[f];  // non-positives with d2 < t
        if (ahead + i >= 100) continue;
        if (eqc[f] > 0) {  // non-positives with the same d2 and a lower row (rare): look them up among the queued keys
            const uint64_t key = pos[i];
            for (int a = 0; a < namb; ++a) {
                const uint64_t k = ambk[a];
                if ((uint32_t)(k >> 32) == t && k < key && !(!last_is_true && k == lastk)) ++ahead;
            }
        }
        const int rank = ahead + i + 1;
        if (rank <= 30) atomicAdd(&misc[5], 1);
        if (rank <= 50) atomicAdd(&misc[6], 1);
        if (rank <= 100) atomicAdd(&misc[7], 1);
    }
    __syncthreads();
    if (tid == 0) {
        p.ap[q] = apv;
        const double npos_true = (double)ntrue + 1e-10;  // :476
        for (int r = 0; r < 3; ++r) p.recall[q * 3 + r] = (double)misc[5 + r] * 1.0 / npos_true;
    }
}

size_t ap_rows_tc_smem(int d) {
    return (size_t)kTcPos * 8 + (size_t)kTcAmb * 8 + 16 + (size_t)kTcAmb * 4 + (size_t)(kTcPos + 2) * 4 +
           (size_t)(kTcPos + 1) * 4 + (size_t)((d + 3) & ~3) * 4 + 40 * 4 + 16 * 4 + 128 + (size_t)kTcRescoreWarps * kRescoreTileBytes + 64;
}

int launch_ap_rows_tc(const ApTcParams &p, cudaStream_t st) {
    if (p.nq_tile == 0) return VQ_OK;
    const size_t smem = ap_rows_tc_smem(p.d);
    VQ_CHECK(smem <= 200 * 1024, VQ_ERR_CAPACITY, "ap_rows_tc: dimension %d too large", p.d);
    if (p.dtype == VQ_F32) {
        VQ_CUDA(cudaFuncSetAttribute(ap_rows_tc_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ap_rows_tc_kernel<float><<<(unsigned)p.nq_tile, kTcThreads, smem, st>>>(p);
    } else {
        VQ_CUDA(cudaFuncSetAttribute(ap_rows_tc_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ap_rows_tc_kernel<__nv_bfloat16><<<(unsigned)p.nq_tile, kTcThreads, smem, st>>>(p);
    }
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

static size_t ap_rows_smem(int pcap2) {
    return (size_t)pcap2 * 8 + (size_t)(pcap2 + 1) * 4 + (size_t)pcap2 * 4 * 2 + (kApThreads + 1) * 4 + 16 * 4 + 16 +
           8 + 16 + 128 * 8 + 128 * 4 + (kApTab + 1) * 4 + 64;
}

int launch_ap_rows(const ApRowsParams &p, int pcap2, cudaStream_t st) {
    if (p.nq_tile == 0) return VQ_OK;
    const size_t smem = ap_rows_smem(pcap2);
    if (p.kept) {
        VQ_CUDA(cudaFuncSetAttribute(ap_rows_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ap_rows_kernel<true><<<(unsigned)p.nq_tile, kApThreads, smem, st>>>(p, pcap2);
    } else {
        VQ_CUDA(cudaFuncSetAttribute(ap_rows_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ap_rows_kernel<false><<<(unsigned)p.nq_tile, kApThreads, smem, st>>>(p, pcap2);
    }
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

// ---------------------------------------------------------------------------------------
// sklearn average_precision_score on a batch of lists (any order in, ties grouped).
// One block per list; positives' scores sorted in shared memory, every element bucketed.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t f64_to_ordered_desc(double s) {
    // larger score -> SMALLER key (so ascending key order = descending score)
    uint64_t u = (uint64_t)__double_as_longlong(s);
    u = (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);  // ascending in s
    return ~u;
}

__global__ void __launch_bounds__(kApThreads) ap_lists_kernel(const int32_t *__restrict__ y_true,
                                                             const double *__restrict__ y_score,
                                                             const int64_t *__restrict__ offsets, double *ap,
                                                             int pcap2, int32_t *overflow) {
    extern __shared__ __align__(16) uint8_t sm[];
    uint64_t *pos = (uint64_t *)sm;
    int *cntA = (int *)(pos + pcap2);
    int *part = cntA + pcap2 + 1;
    int *misc = part + kApThreads + 1;
    const int tid = threadIdx.x;
    const int64_t b = offsets[blockIdx.x], e = offsets[blockIdx.x + 1];
    if (tid == 0) misc[0] = 0;
    for (int i = tid; i < pcap2; i += kApThreads) {
        pos[i] = ~0ull;
        cntA[i] = 0;
    }
    if (tid == 0) cntA[pcap2] = 0;
    __syncthreads();
    for (int64_t j = b + tid; j < e; j += kApThreads)
        if (y_true[j] != 0) {
            const int slot = atomicAdd(&misc[0], 1);
            if (slot < pcap2) pos[slot] = f64_to_ordered_desc(y_score[j]);
        }
    __syncthreads();
    const int npos = misc[0];
    if (npos > pcap2) {
        if (tid == 0) {
            atomicExch(overflow, 1);
            ap[blockIdx.x] = 0.0;
        }
        return;
    }
    if (npos == 0) {
        if (tid == 0) ap[blockIdx.x] = 0.0;
        return;
    }
    block_bitonic_sort(pos, next_pow2(npos));
    for (int64_t j = b + tid; j < e; j += kApThreads) {
        const uint64_t s = f64_to_ordered_desc(y_score[j]);
        int lo = 0, hi = npos;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (pos[mid] < s) lo = mid + 1; else hi = mid;
        }
        atomicAdd(&cntA[lo], 1);
    }
    __syncthreads();
    block_exclusive_scan(cntA, npos + 1, part);
    if (tid == 0) {
        const double P = (double)npos;
        double a = 0.0, tps = 0.0, prev = 0.0;
        int i = 0;
        while (i < npos) {
            const int g = i;
            while (i < npos && pos[i] == pos[g]) {
                tps += 1.0;
                ++i;
            }
            const double recall = tps / P, precision = tps / (double)cntA[g + 1];
            a += (recall - prev) * precision;
            prev = recall;
        }
        ap[blockIdx.x] = a;
    }
}

int launch_average_precision(const int32_t *y_true, const double *y_score, const int64_t *offsets,
                             int64_t nlists, int64_t max_len, double *ap, int32_t *overflow, cudaStream_t st) {
    (void)max_len;
    if (nlists == 0) return VQ_OK;
    const int pcap2 = kMaxPositives;
    const size_t smem = (size_t)pcap2 * 8 + (size_t)(pcap2 + 1) * 4 + (kApThreads + 1) * 4 + 64;
    VQ_CUDA(cudaFuncSetAttribute(ap_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ap_lists_kernel<<<(unsigned)nlists, kApThreads, smem, st>>>(y_true, y_score, offsets, ap, pcap2, overflow);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

// ---------------------------------------------------------------------------------------
// nms_cpu, one warp per segment.  The warp walks the segment in (score desc, row asc) order;
// lanes test the remaining candidates against each newly kept one.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) nms1d_kernel(const double *__restrict__ dets,
                                                    const int64_t *__restrict__ offsets, int64_t nseg,
                                                    double thresh, uint8_t *__restrict__ keep,
                                                    uint8_t *__restrict__ dead) {
    const int lane = threadIdx.x & 31;
    const int64_t seg = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (seg >= nseg) return;
    const int64_t b = offsets[seg], e = offsets[seg + 1];
    const int64_t n = e - b;
    for (int64_t i = lane; i < n; i += 32) {
        keep[b + i] = 0;
        dead[b + i] = 0;
    }
    __syncwarp();
    // repeatedly pick the best not-yet-dead row: (score desc, row asc)
    for (;;) {
        double bs = -INFINITY;
        int64_t bi = -1;
        for (int64_t i = lane; i < n; i += 32) {
            if (dead[b + i]) continue;
            const double s = dets[(b + i) * 3 + 2];
            if (bi < 0 || s > bs) {  // rows visited ascending per lane: first max wins
                bs = s;
                bi = i;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double os = __shfl_xor_sync(0xffffffffu, bs, o);
            const int64_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (oi >= 0 && (bi < 0 || os > bs || (os == bs && oi < bi))) {
                bs = os;
                bi = oi;
            }
        }
        if (bi < 0) break;
        const double x1 = dets[(b + bi) * 3], x2 = dets[(b + bi) * 3 + 1], li = x2 - x1 + 1.0;
        if (lane == 0) {
            keep[b + bi] = 1;
            dead[b + bi] = 1;
        }
        for (int64_t i = lane; i < n; i += 32) {
            if (dead[b + i] || i == bi) continue;
            const double y1 = dets[(b + i) * 3], y2 = dets[(b + i) * 3 + 1], lj = y2 - y1 + 1.0;
            const double xx1 = fmax(x1, y1), xx2 = fmin(x2, y2);
            const double inter = fmax(0.0, xx2 - xx1 + 1.0);
            const double ovr = inter / (li + lj - inter);
            if (!(ovr < thresh)) dead[b + i] = 1;
        }
        __syncwarp();
    }
}

int launch_nms1d(const double *dets, const int64_t *offsets, int64_t nseg, double thresh, uint8_t *keep,
                         uint8_t *dead, cudaStream_t st) {
    if (nseg == 0) return VQ_OK;
    nms1d_kernel<<<(unsigned)((nseg + 7) / 8), 256, 0, st>>>(dets, offsets, nseg, thresh, keep, dead);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

// ---------------------------------------------------------------------------------------
// G-way merge of per-shard lists, one block per query.  Every shard's list arrives sorted by
// (d2, row) -- that is how vq_index_search emits it -- so the merge is a RANK computation, not a sort: the final
// position of element i of list g is  i + sum over the other lists of #{keys < mine}  (keys are distinct: a global row
// lives on one shard), one 7-step binary search per other list, no barrier between elements.  At G = 8, k = 100 that is
// ~6 k warp instructions per query against ~21 k + 55 barriers for the bitonic sort of the 1024-padded buffer.  Lists
// that turn out NOT to be sorted (a caller's own) take the bitonic path.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void merge_sorted_lists(uint64_t *keys, int G, int k, int n2, int64_t q, float *__restrict__ od,
                                                   int64_t *__restrict__ oi) {
    const int total = G * k;
    int bad = 0;
    for (int i = threadIdx.x; i < total; i += blockDim.x)
        if (i % k != k - 1 && keys[i] > keys[i + 1]) bad = 1;
    if (__syncthreads_or(bad)) {
        block_bitonic_sort(keys, n2);
        for (int i = threadIdx.x; i < k; i += blockDim.x) {
            const uint64_t kk = keys[i];
            const bool ok = kk != ~0ull;
            od[q * k + i] = ok ? ordered_to_f32((uint32_t)(kk >> 32)) : INFINITY;
            oi[q * k + i] = ok ? (int64_t)(uint32_t)kk : -1;
        }
        return;
    }
    int valid_t = 0;
    for (int e = threadIdx.x; e < total; e += blockDim.x) {
        const uint64_t mine = keys[e];
        if (mine == ~0ull) continue;
        ++valid_t;
        const int g = e / k;
        int rank = e - g * k;
        for (int h = 0; h < G && rank < k; ++h) {
            if (h == g) continue;
            const uint64_t *lst = keys + h * k;
            int lo = 0, hi = k;  // first index with lst[i] >= mine (empties are all-ones: never below)
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (lst[mid] < mine) lo = mid + 1; else hi = mid;
            }
            rank += lo;
        }
        if (rank < k) {
            od[q * k + rank] = ordered_to_f32((uint32_t)(mine >> 32));
            oi[q * k + rank] = (int64_t)(uint32_t)mine;
        }
    }
    // fewer than k valid entries in total: pad the tail
    __shared__ int s_valid;
    if (threadIdx.x == 0) s_valid = 0;
    __syncthreads();
    if (valid_t) atomicAdd(&s_valid, valid_t);
    __syncthreads();
    for (int i = s_valid + threadIdx.x; i < k; i += blockDim.x) {
        od[q * k + i] = INFINITY;
        oi[q * k + i] = -1;
    }
}

__global__ void __launch_bounds__(128) topk_merge_kernel(const float *__restrict__ d2, const int64_t *__restrict__ idx,
                                                         int G, int64_t nq, int64_t k, float *__restrict__ od,
                                                         int64_t *__restrict__ oi, int n2) {
    extern __shared__ __align__(16) uint8_t sm[];
    uint64_t *keys = (uint64_t *)sm;
    const int64_t q = blockIdx.x;
    const int total = (int)(G * k);
    for (int i = threadIdx.x; i < n2 + 1; i += blockDim.x) {
        uint64_t kk = ~0ull;
        if (i < total) {
            const int g = i / (int)k, j = i % (int)k;
            const int64_t o = ((int64_t)g * nq + q) * k + j;
            if (idx[o] >= 0) kk = make_key_asc(d2[o], (uint32_t)idx[o]);
        }
        keys[i] = kk;
    }
    __syncthreads();
    merge_sorted_lists(keys, G, (int)k, n2, q, od, oi);
}

// Packed exchange: a rank's local list as ONE 8-byte key per entry, (ordered d2 << 32 | global row),
// so the all-gather moves 8 instead of 12 bytes per entry in one collective and the merge ranks what arrives.
__global__ void __launch_bounds__(256) topk_pack_kernel(const float *__restrict__ d2, const int64_t *__restrict__ idx,
                                                        int64_t count, int64_t row_offset, uint64_t *__restrict__ keys) {
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= count) return;
    const int64_t r = idx[i];
    keys[i] = r < 0 ? ~0ull : make_key_asc(d2[i], (uint32_t)(r + row_offset));
}

__global__ void __launch_bounds__(128) topk_merge_keys_kernel(const uint64_t *__restrict__ in, int G, int64_t nq, int64_t k,
                                                              float *__restrict__ od, int64_t *__restrict__ oi, int n2) {
    extern __shared__ __align__(16) uint8_t sm[];
    uint64_t *keys = (uint64_t *)sm;
    const int64_t q = blockIdx.x;
    const int total = (int)(G * k);
    for (int i = threadIdx.x; i < n2 + 1; i += blockDim.x) {
        const int g = i / (int)k, j = i % (int)k;
        keys[i] = i < total ? in[((int64_t)g * nq + q) * k + j] : ~0ull;
    }
    __syncthreads();
    merge_sorted_lists(keys, G, (int)k, n2, q, od, oi);
}

// Head exchange (two rounds): with G shards a query's global top-k takes ~k/G entries from each shard, so
// the ranks first exchange only the first m < k entries of their lists.  The merge of the G heads is FINAL for a query
// when every head either ended inside its list (all-ones padding) or ends at / behind the merged k-th key: everything a
// rank did not send is worse than its own last sent key.  Queries for which that fails are counted in `incomplete`; the
// caller then falls back to the full exchange (ops.exchange_merge).  Heads that are not sorted count as incomplete.
__global__ void __launch_bounds__(128) topk_merge_heads_kernel(const uint64_t *__restrict__ in, int G, int64_t nq, int m, int k,
                                                               float *__restrict__ od, int64_t *__restrict__ oi,
                                                               int32_t *__restrict__ incomplete) {
    extern __shared__ __align__(16) uint8_t sm[];
    uint64_t *keys = (uint64_t *)sm;
    __shared__ unsigned long long s_kth;
    __shared__ int s_valid, s_bad;
    const int64_t q = blockIdx.x;
    const int total = G * m;
    if (threadIdx.x == 0) {
        s_kth = ~0ull;
        s_valid = 0;
        s_bad = 0;
    }
    for (int i = threadIdx.x; i < total; i += blockDim.x) keys[i] = in[((int64_t)(i / m) * nq + q) * m + (i % m)];
    __syncthreads();
    int bad = 0, valid_t = 0;
    for (int e = threadIdx.x; e < total; e += blockDim.x) {
        const uint64_t mine = keys[e];
        const int g = e / m, i = e - g * m;
        if (i + 1 < m && mine > keys[e + 1]) bad = 1;
        if (mine == ~0ull) continue;
        ++valid_t;
        int rank = i;
        for (int h = 0; h < G && rank < k; ++h) {
            if (h == g) continue;
            const uint64_t *lst = keys + h * m;
            int lo = 0, hi = m;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (lst[mid] < mine) lo = mid + 1; else hi = mid;
            }
            rank += lo;
        }
        if (rank < k) {
            od[q * k + rank] = ordered_to_f32((uint32_t)(mine >> 32));
            oi[q * k + rank] = (int64_t)(uint32_t)mine;
            if (rank == k - 1) s_kth = mine;
        }
    }
    if (valid_t) atomicAdd(&s_valid, valid_t);
    if (bad) s_bad = 1;
    __syncthreads();
    for (int i = s_valid + threadIdx.x; i < k; i += blockDim.x) {
        od[q * k + i] = INFINITY;
        oi[q * k + i] = -1;
    }
    if (threadIdx.x < G) {
        const uint64_t last = keys[threadIdx.x * m + m - 1];
        if (last != ~0ull && last < s_kth) s_bad = 1;  // this shard may hold better entries it has not sent
    }
    __syncthreads();
    if (threadIdx.x == 0 && s_bad) atomicAdd(incomplete, 1);
}

int launch_topk_merge_heads(const uint64_t *keys, int G, int64_t nq, int m, int k, float *od, int64_t *oi, int32_t *incomplete,
                            cudaStream_t st) {
    if (nq == 0) return VQ_OK;
    const size_t smem = (size_t)(G * m + 1) * 8;
    VQ_CHECK(smem <= 200 * 1024 && G <= 128, VQ_ERR_CAPACITY, "topk_merge_heads: G*m=%lld too large", (long long)G * m);
    VQ_CUDA(cudaFuncSetAttribute(topk_merge_heads_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    topk_merge_heads_kernel<<<(unsigned)nq, 128, smem, st>>>(keys, G, nq, m, k, od, oi, incomplete);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

int launch_topk_pack(const float *d2, const int64_t *idx, int64_t count, int64_t row_offset, uint64_t *keys, cudaStream_t st) {
    if (count == 0) return VQ_OK;
    topk_pack_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(d2, idx, count, row_offset, keys);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

int launch_topk_merge_keys(const uint64_t *keys, int G, int64_t nq, int64_t k, float *od, int64_t *oi, cudaStream_t st) {
    if (nq == 0) return VQ_OK;
    const int n2 = next_pow2((int)(G * k));
    const size_t smem = (size_t)(n2 + 1) * 8;  // + one sentinel slot behind the last list
    VQ_CHECK(smem <= 200 * 1024, VQ_ERR_CAPACITY, "topk_merge: G*k=%lld too large", (long long)(G * k));
    VQ_CUDA(cudaFuncSetAttribute(topk_merge_keys_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    topk_merge_keys_kernel<<<(unsigned)nq, 128, smem, st>>>(keys, G, nq, k, od, oi, n2);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

int launch_topk_merge(const float *d2, const int64_t *idx, int G, int64_t nq, int64_t k, float *od, int64_t *oi,
                      cudaStream_t st) {
    if (nq == 0) return VQ_OK;
    const int n2 = next_pow2((int)(G * k));
    const size_t smem = (size_t)(n2 + 1) * 8;
    VQ_CHECK(smem <= 200 * 1024, VQ_ERR_CAPACITY, "topk_merge: G*k=%lld too large", (long long)(G * k));
    VQ_CUDA(cudaFuncSetAttribute(topk_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    topk_merge_kernel<<<(unsigned)nq, 128, smem, st>>>(d2, idx, G, nq, k, od, oi, n2);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

}  // namespace vq
