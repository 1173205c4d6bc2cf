// vq_moment.cu -- moment retrieval evaluation: the per-query body of
// ARV_Retrieval_Moment.ranking (dataloader_baseline.py:1225-1330) for a tile of queries.
//
//   moment_nms   one thread group per (query, video): the video's moments sorted by canonical
//                (d2, row) -- the order the ranked list visits them (:1268-1281) -- then the greedy
//                temporal NMS of do_nms / nms_cpu (:1283-1312, misc_utils/utils_models.py:153-174:
//                float64, inclusive length x2-x1+1, a later moment survives iff IoU < thresh).
//                Writes the survivor mask and the key of the video's best moment, which is the
//                video's place in the first-seen cluster order.
//   (AP and the grouped R@N over the survivors: ap_rows_kernel<true> in vq_metrics.cu)
//
// The gallery rows of one video must be contiguous (vq_index_set_segments), as the reference builds
// them (:1091-1151).  Groups: a warp for videos of <= 256 moments, a block for <= 4096.
#include "vq_blocksel.cuh"
#include "vq_kernels.h"

namespace vq {

template <int NT>
__device__ __forceinline__ void grp_sync() {
    if (NT == 32) __syncwarp(); else __syncthreads();
}

// nms_cpu's test for one pair (utils_models.py:165-172): ovr = inter / (la + lb - inter); the later moment is
// suppressed unless ovr < thresh.  No division on the common paths, same answer bit for bit:
//   r = fma(thresh, union, -inter) has the exact sign of thresh*union - inter (one rounding, zero only if equal);
//   r <= 0: the exact quotient is >= thresh, and rounding is monotonic with thresh representable -> suppressed
//           (covers the ties of regular window grids, IoU exactly = thresh);
//   r > 2^-50 * thresh * union: the exact quotient is more than an ulp below thresh -> rounds below it -> kept;
//   the sliver in between (and union <= 0 / NaN inputs) forms the quotient as the reference does.
__device__ __noinline__ bool nms_quotient(double inter, double uni, double thresh) {  // rare: keep it out of line
    const double ovr = inter / uni;
    return !(ovr < thresh);
}
__device__ __forceinline__ bool nms_suppressed(double xa1, double xa2, double la, double xb1, double xb2, double lb,
                                               double thresh) {
    const double xx1 = xa1 > xb1 ? xa1 : xb1, xx2 = xa2 < xb2 ? xa2 : xb2;
    double inter = xx2 - xx1 + 1.0;
    inter = inter < 0.0 ? 0.0 : inter;
    const double uni = la + lb - inter;
    if (uni > 0.0 && thresh > 0.0) {
        const double r = fma(thresh, uni, -inter);
        if (r <= 0.0) return true;
        if (r > 0x1p-50 * (thresh * uni)) return false;
    }
    return nms_quotient(inter, uni, thresh);
}

// Suppression relation of the moments of one video, as bit masks: sup[row] bit j = "moment j of the same video is
// suppressed by this one" (nms_suppressed: IoU >= thresh, the reference's float64 test).  It depends on the video and the
// threshold only -- not on the query -- so it is built once per (index, threshold) for the videos of <= 64 moments and
// the per-(query, video) greedy pass becomes bit arithmetic: dead |= sup[a] for every survivor a in ranked order.
__global__ void __launch_bounds__(256) nms_masks_kernel(const int64_t *__restrict__ seg_start, const int32_t *__restrict__ seg_list,
                                                        int nlist, const float *__restrict__ start, const float *__restrict__ end,
                                                        double thresh, unsigned long long *__restrict__ sup) {
    const int li = blockIdx.x * 8 + (threadIdx.x >> 5), t = threadIdx.x & 31;
    if (li >= nlist) return;
    const int v = seg_list[li];
    const int64_t s0 = seg_start[v];
    const int S = (int)(seg_start[v + 1] - s0);
    if (S > 64) return;
    for (int i = t; i < S; i += 32) {
        const double xa1 = (double)start[s0 + i], xa2 = (double)end[s0 + i], la = xa2 - xa1 + 1.0;
        unsigned long long m = 0ull;
        for (int j = 0; j < S; ++j) {
            const double xb1 = (double)start[s0 + j], xb2 = (double)end[s0 + j];
            if (j != i && nms_suppressed(xa1, xa2, la, xb1, xb2, xb2 - xb1 + 1.0, thresh)) m |= 1ull << j;
        }
        sup[s0 + i] = m;
    }
}

int launch_nms_masks(const int64_t *seg_start, const int32_t *small_list, int n_small, const float *start, const float *end,
                     double thresh, unsigned long long *sup, cudaStream_t st) {
    if (n_small == 0) return VQ_OK;
    nms_masks_kernel<<<(unsigned)((n_small + 7) / 8), 256, 0, st>>>(seg_start, small_list, n_small, start, end, thresh, sup);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

static constexpr int kNmsRankPitch = 68;  // bytes per video of a batch: 17 words, so that the lanes' word loads of phase B hit 32 banks
static constexpr int kNmsBatchBytes = 32 * kNmsRankPitch + 32 * 8 + 32 * 8 + 32;  // per warp: ranks | keep masks | first rows | lengths

template <int NT, int CAP>
__global__ void __launch_bounds__(256, NT == 32 ? 4 : 1) moment_nms_kernel(const MomentNmsParams p, const int32_t *__restrict__ seg_list,
                                                         int nlist, int per_group) {
    extern __shared__ __align__(16) uint8_t sm_raw[];
    constexpr int GPB = 256 / NT;
    constexpr int GBYTES = CAP * 8 + CAP * 4 + CAP * 4 + CAP;
    const int gid = threadIdx.x / NT, t = threadIdx.x % NT;
    uint8_t *base = sm_raw + (size_t)gid * GBYTES;
    uint64_t *keys = (uint64_t *)base;
    float *x1 = (float *)(base + CAP * 8);
    float *x2 = x1 + CAP;
    uint8_t *dead = (uint8_t *)(x2 + CAP);
    const int64_t q = blockIdx.y;
    const float *row = p.d2 + q * p.n;
    uint8_t *kept = p.kept + q * p.n;
    const int li0 = (blockIdx.x * GPB + gid) * per_group;
    // Videos of <= 64 moments with precomputed suppression masks, a batch of up to 32 per warp:
    //   A  the WARP sorts one video after the other (two keys per lane, bitonic network on registers) and leaves the
    //      ranked moment numbers in shared memory, one byte each;
    //   B  LANE j walks video j of the batch: greedy NMS as bit arithmetic, dead |= sup[a] for every survivor a in ranked
    //      order -- 64 dependent steps, but 32 videos per instruction instead of one (the walk was half of the
    //      warp-per-video form's instructions);
    //   C  the warp writes the survivor bytes of the batch, coalesced.
    // A video of the list that does not qualify (65 .. 256 moments) takes the per-video path below.
    static_assert(NT != 32 || kNmsBatchBytes <= GBYTES, "the batch arrays reuse the warp's sort / NMS arrays");
    uint8_t *fbase = base;  // (the per-video path below runs after phase C of a batch and before phase A of the next)
    uint8_t *b_rank = fbase;                                                          // [32][68] ranked moment numbers
    unsigned long long *b_keep = (unsigned long long *)(fbase + 32 * kNmsRankPitch);   // [32] survivor masks
    long long *b_s0 = (long long *)(b_keep + 32);                                     // [32] first row of the video
    uint8_t *b_S = (uint8_t *)(b_s0 + 32);                                            // [32] moments (0: not in the batch)
    const bool batched = NT == 32 && p.sup != nullptr;
    for (int lb = li0; lb < li0 + per_group && lb < nlist; lb += (batched ? 32 : 1)) {
      int li = lb;
      if (batched) {
        const int cnt = min(min(32, li0 + per_group - lb), nlist - lb);
        __syncwarp();
        for (int j = 0; j < cnt; ++j) {  // ---- A
            const int v = seg_list[lb + j];
            const int64_t s0 = p.seg_start[v];
            const int S = (int)(p.seg_start[v + 1] - s0);
            if (S == 0 || S > 64) {
                if (t == 0) b_S[j] = 0;
                if (S == 0 && t == 0) p.vbest[q * p.nseg + v] = ~0ull;
                continue;
            }
            // First on 32-bit keys: the score's ordered image with its low 6 bits replaced by the moment number -- one
            // shuffle and one min / max per compare-exchange instead of two shuffles and a 64-bit compare.  The order is
            // the exact (d2, moment) order unless two moments agree in the 26 score bits that were kept; such a pair ends
            // up adjacent, is detected, and the video is sorted again on the full 64-bit keys.
            const float d0 = t < S ? row[s0 + t] : 0.f, d1 = t + 32 < S ? row[s0 + t + 32] : 0.f;
            uint32_t c0 = t < S ? ((f32_to_ordered(d0) & ~63u) | (uint32_t)t) : 0xffffffffu;
            uint32_t c1 = t + 32 < S ? ((f32_to_ordered(d1) & ~63u) | (uint32_t)(t + 32)) : 0xffffffffu;
            auto cx32 = [&](uint32_t &x, int jj, bool asc) {
                const uint32_t o = __shfl_xor_sync(0xffffffffu, x, jj);
                const bool takemin = ((t & jj) == 0) == asc;
                x = takemin ? min(x, o) : max(x, o);
            };
#pragma unroll
            for (int k = 2; k <= 32; k <<= 1) {
                const bool a0 = k < 32 ? (t & k) == 0 : true, a1 = k < 32 ? a0 : false;
#pragma unroll
                for (int jj = k >> 1; jj > 0; jj >>= 1) {
                    cx32(c0, jj, a0);
                    cx32(c1, jj, a1);
                }
            }
            if (c1 < c0) {
                const uint32_t tmp = c0;
                c0 = c1;
                c1 = tmp;
            }
#pragma unroll
            for (int jj = 16; jj > 0; jj >>= 1) {
                cx32(c0, jj, true);
                cx32(c1, jj, true);
            }
            // rank r and r + 1 with equal kept score bits?  (ranks t / t + 1 and t + 32 / t + 33; pads are all ones)
            uint32_t n0 = __shfl_down_sync(0xffffffffu, c0, 1), n1 = __shfl_down_sync(0xffffffffu, c1, 1);
            const uint32_t c1_first = __shfl_sync(0xffffffffu, c1, 0);
            if (t == 31) {
                n0 = c1_first;
                n1 = 0xffffffffu;
            }
            const bool clash = (t + 1 < S && ((c0 ^ n0) & ~63u) == 0u) || (t + 33 < S && ((c1 ^ n1) & ~63u) == 0u);
            uint64_t k0, k1;
            if (__any_sync(0xffffffffu, clash)) {  // uniform, rare: the exact sort on (d2, moment)
                k0 = t < S ? make_key_asc(d0, (uint32_t)t) : ~0ull;
                k1 = t + 32 < S ? make_key_asc(d1, (uint32_t)(t + 32)) : ~0ull;
                auto cx = [&](uint64_t &x, int jj, bool asc) {
                    const uint64_t o = __shfl_xor_sync(0xffffffffu, x, jj);
                    const bool takemin = ((t & jj) == 0) == asc;
                    x = ((o < x) == takemin) ? o : x;
                };
#pragma unroll 1
                for (int k = 2; k <= 32; k <<= 1) {
                    const bool a0 = k < 32 ? (t & k) == 0 : true, a1 = k < 32 ? a0 : false;
#pragma unroll 1
                    for (int jj = k >> 1; jj > 0; jj >>= 1) {
                        cx(k0, jj, a0);
                        cx(k1, jj, a1);
                    }
                }
                if (k1 < k0) {
                    const uint64_t tmp = k0;
                    k0 = k1;
                    k1 = tmp;
                }
#pragma unroll 1
                for (int jj = 16; jj > 0; jj >>= 1) {
                    cx(k0, jj, true);
                    cx(k1, jj, true);
                }
            } else {  // moment numbers in the low bytes; the exact key of the best moment for vbest
                k0 = (uint64_t)(c0 & 63u);
                k1 = (uint64_t)(c1 & 63u);
                if (t == 0) k0 = make_key_asc(row[s0 + (c0 & 63u)], c0 & 63u);
            }
            b_rank[j * kNmsRankPitch + t] = (uint8_t)k0;        // moment (index inside the video) at ranks t, t + 32;
            b_rank[j * kNmsRankPitch + t + 32] = (uint8_t)k1;   // ranks >= S hold the pad key's low byte, never read
            if (t == 0) {
                b_S[j] = (uint8_t)S;
                b_s0[j] = s0;
                p.vbest[q * p.nseg + v] = (k0 & 0xffffffff00000000ull) | (uint64_t)(s0 + (uint32_t)k0);
            }
        }
        __syncwarp();
        {  // ---- B
            const int S = t < cnt ? (int)b_S[t] : 0;
            const int Smax = __reduce_max_sync(0xffffffffu, S);
            const unsigned long long *sup = p.sup + (t < cnt && S ? b_s0[t] : 0);
            const uint32_t *rk = (const uint32_t *)(b_rank + t * kNmsRankPitch);
            unsigned long long dead = 0ull, keep = 0ull;
            for (int r4 = 0; r4 < Smax; r4 += 4) {
                const uint32_t four = rk[r4 >> 2];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const uint32_t a = (four >> (8 * e)) & 63u;
                    if (r4 + e < S && !((dead >> a) & 1ull)) {
                        keep |= 1ull << a;
                        dead |= sup[a];
                    }
                }
            }
            b_keep[t] = keep;
        }
        __syncwarp();
        for (int j = 0; j < cnt; ++j) {  // ---- C
            const int S = b_S[j];
            if (S == 0) continue;
            const unsigned long long keep = b_keep[j];
            uint8_t *kd = kept + b_s0[j];
            if (t < S) kd[t] = (uint8_t)((keep >> t) & 1ull);
            if (t + 32 < S) kd[t + 32] = (uint8_t)((keep >> (t + 32)) & 1ull);
        }
      }
      for (; li < (batched ? lb + min(min(32, li0 + per_group - lb), nlist - lb) : lb + 1); ++li) {
        const int v = seg_list[li];
        const int64_t s0 = p.seg_start[v];
        const int S = (int)(p.seg_start[v + 1] - s0);
        if (batched && S <= 64) continue;  // done above
        if (S == 0) {
            if (t == 0) p.vbest[q * p.nseg + v] = ~0ull;
            continue;
        }
        int n2 = 1;
        while (n2 < S) n2 <<= 1;
        grp_sync<NT>();  // the previous segment's arrays are free
        if (NT == 32 && S <= 64) {
            // two keys per lane (ranks t and t + 32), bitonic network on registers: partners below 32 apart are a
            // shuffle away, the 32-apart partner is the lane's other register
            uint64_t k0 = t < S ? make_key_asc(row[s0 + t], (uint32_t)t) : ~0ull;
            uint64_t k1 = t + 32 < S ? make_key_asc(row[s0 + t + 32], (uint32_t)(t + 32)) : ~0ull;
            auto cx = [&](uint64_t &x, int j, bool asc) {
                const uint64_t o = __shfl_xor_sync(0xffffffffu, x, j);
                const bool takemin = ((t & j) == 0) == asc;
                x = ((o < x) == takemin) ? o : x;
            };
#pragma unroll
            for (int k = 2; k <= 32; k <<= 1) {
                const bool a0 = k < 32 ? (t & k) == 0 : true, a1 = k < 32 ? a0 : false;
#pragma unroll
                for (int j = k >> 1; j > 0; j >>= 1) {
                    cx(k0, j, a0);
                    cx(k1, j, a1);
                }
            }
            if (k1 < k0) {
                const uint64_t tmp = k0;
                k0 = k1;
                k1 = tmp;
            }
#pragma unroll
            for (int j = 16; j > 0; j >>= 1) {
                cx(k0, j, true);
                cx(k1, j, true);
            }
            keys[t] = k0;
            keys[t + 32] = k1;
        } else {
            for (int i = t; i < n2; i += NT) keys[i] = i < S ? make_key_asc(row[s0 + i], (uint32_t)i) : ~0ull;
            for (int k = 2; k <= n2; k <<= 1)
                for (int j = k >> 1; j > 0; j >>= 1) {
                    grp_sync<NT>();
                    for (int h = t; h < (n2 >> 1); h += NT) {  // pair h: i = h with a zero inserted at bit j
                        const int i = ((h & ~(j - 1)) << 1) | (h & (j - 1)), ixj = i | j;
                        const uint64_t a = keys[i], b = keys[ixj];
                        if ((a > b) == ((i & k) == 0)) {
                            keys[i] = b;
                            keys[ixj] = a;
                        }
                    }
                }
        }
        grp_sync<NT>();
        for (int i = t; i < S; i += NT) {
            const int64_t r = s0 + (uint32_t)keys[i];
            x1[i] = p.start[r];
            x2[i] = p.end[r];
            dead[i] = 0;
        }
        grp_sync<NT>();
        // greedy NMS in ranked order (nms_cpu :159-173; equal scores: earlier rank first)
        if (NT == 32) {
            // warp form: lane t owns ranks t, t+32, ... and keeps their dead bits in a register; the survivors of
            // the current 32-rank chunk are a ballot, so a round costs no shared-memory flag traffic and no barrier
            uint32_t mydead = 0;
            const int nchunk = (S + 31) >> 5;
            constexpr int RC = 2;  // the lane's first RC ranks stay in registers as doubles (videos of <= 64 moments: all)
            double bx1[RC], bx2[RC], blb[RC];
#pragma unroll
            for (int jj = 0; jj < RC; ++jj) {
                const int b = 32 * jj + t;
                bx1[jj] = b < S ? (double)x1[b] : 0.0;
                bx2[jj] = b < S ? (double)x2[b] : 0.0;
                blb[jj] = bx2[jj] - bx1[jj] + 1.0;
            }
            for (int j = 0; j < nchunk; ++j) {
                const uint32_t valid = (S - 32 * j) >= 32 ? 0xffffffffu : ((1u << (S - 32 * j)) - 1u);
                uint32_t todo = ~__ballot_sync(0xffffffffu, (mydead >> j) & 1u) & valid;
                while (todo) {
                    const int al = __ffs(todo) - 1, a = 32 * j + al;
                    const double xa1 = (double)x1[a], xa2 = (double)x2[a], la = xa2 - xa1 + 1.0;
#pragma unroll
                    for (int jj = 0; jj < RC; ++jj) {
                        const int b = 32 * jj + t;
                        if (jj >= j && b > a && b < S && !((mydead >> jj) & 1u) &&
                            nms_suppressed(xa1, xa2, la, bx1[jj], bx2[jj], blb[jj], p.thresh))
                            mydead |= 1u << jj;
                    }
                    for (int jj = j > RC ? j : RC; jj < nchunk; ++jj) {
                        const int b = 32 * jj + t;
                        if (b > a && b < S && !((mydead >> jj) & 1u)) {
                            const double xb1 = (double)x1[b], xb2 = (double)x2[b];
                            if (nms_suppressed(xa1, xa2, la, xb1, xb2, xb2 - xb1 + 1.0, p.thresh)) mydead |= 1u << jj;
                        }
                    }
                    todo = ~__ballot_sync(0xffffffffu, (mydead >> j) & 1u) & valid & ~((2u << al) - 1u);
                }
            }
            for (int jj = 0; jj < nchunk; ++jj)
                if (32 * jj + t < S) dead[32 * jj + t] = (mydead >> jj) & 1u;
            grp_sync<NT>();
        } else {
            for (int a = 0; a < S; ++a) {
                if (dead[a]) continue;
                const double xa1 = (double)x1[a], xa2 = (double)x2[a], la = xa2 - xa1 + 1.0;
                for (int b = a + 1 + t; b < S; b += NT) {
                    if (dead[b]) continue;
                    const double xb1 = (double)x1[b], xb2 = (double)x2[b];
                    if (nms_suppressed(xa1, xa2, la, xb1, xb2, xb2 - xb1 + 1.0, p.thresh)) dead[b] = 1;
                }
                grp_sync<NT>();
            }
        }
        for (int i = t; i < S; i += NT) kept[s0 + (uint32_t)keys[i]] = dead[i] ? 0 : 1;
        if (t == 0) p.vbest[q * p.nseg + v] = (keys[0] & 0xffffffff00000000ull) | (uint64_t)(s0 + (uint32_t)keys[0]);
      }
    }
}

int launch_moment_nms(const MomentNmsParams &p, const int32_t *small_list, int n_small, const int32_t *large_list,
                      int n_large, cudaStream_t st) {
    if (p.nq_tile == 0) return VQ_OK;
    if (n_small > 0) {
        constexpr int CAP = kMomentWarpCap, GB = CAP * 17;
        const int PER = p.sup ? 32 : 4;  // with suppression masks a warp takes its videos in batches of 32
        const int blocks = (n_small + 8 * PER - 1) / (8 * PER);
        moment_nms_kernel<32, CAP><<<dim3((unsigned)blocks, (unsigned)p.nq_tile), 256, 8 * GB, st>>>(p, small_list, n_small, PER);
        VQ_CUDA(cudaGetLastError());
    }
    if (n_large > 0) {
        constexpr int CAP = kMomentBlockCap, GB = CAP * 17;
        VQ_CUDA(cudaFuncSetAttribute(moment_nms_kernel<256, CAP>, cudaFuncAttributeMaxDynamicSharedMemorySize, GB));
        moment_nms_kernel<256, CAP><<<dim3((unsigned)n_large, (unsigned)p.nq_tile), 256, GB, st>>>(p, large_list, n_large, 1);
        VQ_CUDA(cudaGetLastError());
    }
    return VQ_OK;
}

}  // namespace vq
