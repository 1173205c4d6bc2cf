// vq_triplet.cu -- kernel (3): in-batch triplet loss forward + backward with fused mining.
//
//   pdist          Sampler.pdist, misc_utils/metric_learning_loss.py:276-289
//   bh_mine        batch-hard mining on that distance matrix (extension; masks as
//                  semihardsampling :173-176: pos = same label minus the anchor, neg = other
//                  labels; p = argmax dist, n = argmin dist, ties -> lowest index) fused with
//                  TripletLoss.triplet_distance :342-355 (squared L2 recomputed from the rows)
//   triplet_losses the same per-triplet loss for caller-sampled triplets
//   grad           loss mean (:379) and the autograd backward of :368-379 written as a
//                  deterministic GATHER: every gradient row scans the triplet list in order and
//                  accumulates acc = fmaf(2/T, diff, acc) -- no floating-point atomics.
// All arithmetic is the canonical sequential-fmaf form of oracle/vq_oracle.c, so loss, grad and
// the mined indices are bit-identical to the oracle.
#include "vq_blocksel.cuh"
#include "vq_kernels.h"

namespace vq {

// Canonical blocked chain of the loss path (oracle/vq_oracle.c: orc_dot / orc_d2_blocked): blocks of
// 32 k; chunk c = 4 consecutive k as a sequential fmaf chain from 0; fixed tree over the 8 chunks
// (c ^ 4, c ^ 2, c ^ 1); block sums added to the accumulator in order.
__device__ __forceinline__ float tree8(const float (&t)[8]) {
    const float u0 = __fadd_rn(t[0], t[4]), u1 = __fadd_rn(t[1], t[5]), u2 = __fadd_rn(t[2], t[6]),
                u3 = __fadd_rn(t[3], t[7]);
    return __fadd_rn(__fadd_rn(u0, u2), __fadd_rn(u1, u3));
}
template <bool SQDIFF>
__device__ __forceinline__ float blocked_chain(const float *a, const float *b, int d) {
    float acc = 0.f;
    for (int k0 = 0; k0 < d; k0 += 32) {
        float t[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            t[c] = 0.f;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int k = k0 + c * 4 + i;
                if (k < d) {
                    if (SQDIFF) {
                        const float df = __fsub_rn(a[k], b[k]);
                        t[c] = __fmaf_rn(df, df, t[c]);
                    } else {
                        t[c] = __fmaf_rn(a[k], b[k], t[c]);
                    }
                }
            }
        }
        acc = __fadd_rn(acc, tree8(t));
    }
    return acc;
}
__device__ __forceinline__ float dot_chain(const float *a, const float *b, int d) { return blocked_chain<false>(a, b, d); }
__device__ __forceinline__ float d2_chain(const float *a, const float *b, int d) { return blocked_chain<true>(a, b, d); }
__device__ __forceinline__ float pdist_formula(float ni, float nj, float prod, float eps) {
    float res = __fsub_rn(__fadd_rn(ni, nj), __fmul_rn(2.0f, prod));
    res = res < 0.f ? 0.f : res;
    res = res < eps ? eps : res;
    return __fsqrt_rn(res);
}

static constexpr int MT = 256;   // threads
static constexpr int MA = 32;    // anchors per block
static constexpr int MKC = 32;   // k chunk
static constexpr int MBMAX = 1024;  // max group size B

// grid (B/MA rounded up, G).  Dynamic smem: float nrm[B] | float sA[MKC][MA+1] | float sR[MKC][B+1]
// dist_out (optional): [G? no -- only when G == 1] B x B matrix.
// glist/gcount (optional): blockIdx.y indexes a device list of group ids (the groups the fused tensor-core
// kernel handed back, vq_triplet_tc.cu); blocks past *gcount exit.  NCMAX = columns per thread (B <= 8*NCMAX).
template <int NCMAX>
__global__ void __launch_bounds__(MT) bh_mine_kernel(const float *__restrict__ emb, const int32_t *__restrict__ labels,
                                                    int B, int d, float margin, float eps, int32_t *__restrict__ trip,
                                                    float *__restrict__ lt, float *__restrict__ dist_out,
                                                    const int32_t *__restrict__ glist, const int32_t *__restrict__ gcount) {
    extern __shared__ __align__(16) float smf[];
    float *nrm = smf;
    float *sA = nrm + B;
    float *sR = sA + MKC * (MA + 1);
    __shared__ int s_p[MA], s_n[MA];
    const int tid = threadIdx.x;
    const int nlist = glist ? *gcount : (int)gridDim.y;
    for (int li = blockIdx.y; li < nlist; li += gridDim.y) {
    const int g = glist ? glist[li] : li;
    __syncthreads();  // shared tiles of the previous group are no longer read
    const float *E = emb + (int64_t)g * B * d;
    const int32_t *L = labels ? labels + (int64_t)g * B : nullptr;
    const int a0 = blockIdx.x * MA;
    // row norms = diagonal of the Gram matrix, canonical chain (prod.diag(), :286).  Eight lanes per row, lane c =
    // chunk c of every 32-k block (dot_chain's t[c]); the xor-4 / 2 / 1 exchange is tree8 (fadd commutes, so every
    // lane of the eight ends up with the same bits), and a warp reads 4 x 128 contiguous bytes per step instead of
    // 32 rows at a 4*d-byte stride.
    {
        const int lane = tid & 31, c = lane & 7;
        for (int r0 = (tid >> 5) * 4; r0 < B; r0 += (MT / 32) * 4) {  // warp-uniform
            const int r = r0 + (lane >> 3);
            const float *row = E + (int64_t)min(r, B - 1) * d;
            float acc = 0.f;
            for (int k0 = 0; k0 < d; k0 += 32) {
                float t = 0.f;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int k = k0 + c * 4 + i;
                    const float v = k < d ? row[k] : 0.f;  // fmaf(0, 0, t) == t: same as dot_chain's skip
                    t = __fmaf_rn(v, v, t);
                }
                t = __fadd_rn(t, __shfl_xor_sync(0xffffffffu, t, 4));
                t = __fadd_rn(t, __shfl_xor_sync(0xffffffffu, t, 2));
                t = __fadd_rn(t, __shfl_xor_sync(0xffffffffu, t, 1));
                acc = __fadd_rn(acc, t);
            }
            if (c == 0 && r < B) nrm[r] = acc;
        }
    }
    const int al = tid >> 3, cg = tid & 7;  // anchor (local), column group
    const int ncol = (B + 7) / 8;           // columns per thread: cg, cg+8, ...
    float acc[NCMAX];
#pragma unroll
    for (int c = 0; c < NCMAX; ++c) acc[c] = 0.f;
    for (int k0 = 0; k0 < d; k0 += MKC) {
        __syncthreads();
        for (int i = tid; i < MA * MKC; i += MT) {
            const int r = i / MKC, k = i % MKC;
            sA[k * (MA + 1) + r] = (a0 + r < B && k0 + k < d) ? E[(int64_t)(a0 + r) * d + k0 + k] : 0.f;
        }
        for (int i = tid; i < B * MKC; i += MT) {
            const int r = i / MKC, k = i % MKC;
            sR[k * (B + 1) + r] = (k0 + k < d) ? E[(int64_t)r * d + k0 + k] : 0.f;
        }
        __syncthreads();
        // one canonical 32-block per k chunk (rows past d are zero in shared memory: no effect)
        float av[MKC];
#pragma unroll
        for (int kk = 0; kk < MKC; ++kk) av[kk] = sA[kk * (MA + 1) + al];
#pragma unroll
        for (int c = 0; c < NCMAX; ++c) {
            const int j = cg + 8 * c;
            if (c < ncol && j < B) {
                float t[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    t[q] = 0.f;
#pragma unroll
                    for (int i = 0; i < 4; ++i) t[q] = __fmaf_rn(av[4 * q + i], sR[(4 * q + i) * (B + 1) + j], t[q]);
                }
                acc[c] = __fadd_rn(acc[c], tree8(t));
            }
        }
    }
    __syncthreads();
    const int a = a0 + al;
    // distances + mining
    float bp = -1.f, bn = 0.f;
    int ip = -1, in_ = -1;
    if (a < B) {
        const int la = L ? L[a] : 0;
#pragma unroll
        for (int c = 0; c < NCMAX; ++c)
            if (c < ncol) {
                const int j = cg + 8 * c;
                if (j < B) {
                    const float dist = pdist_formula(nrm[a], nrm[j], acc[c], eps);
                    if (dist_out) dist_out[((int64_t)g * B + a) * B + j] = dist;
                    if (L && j != a) {
                        if (L[j] == la) {
                            if (ip < 0 || dist > bp) { bp = dist; ip = j; }  // j ascending: first max wins
                        } else {
                            if (in_ < 0 || dist < bn) { bn = dist; in_ = j; }
                        }
                    }
                }
            }
    }
    // reduce over the 8 column groups of this anchor (lanes al*8 .. al*8+7 are contiguous)
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        const float obp = __shfl_xor_sync(0xffffffffu, bp, o), obn = __shfl_xor_sync(0xffffffffu, bn, o);
        const int oip = __shfl_xor_sync(0xffffffffu, ip, o), oin = __shfl_xor_sync(0xffffffffu, in_, o);
        if (oip >= 0 && (ip < 0 || obp > bp || (obp == bp && oip < ip))) { bp = obp; ip = oip; }
        if (oin >= 0 && (in_ < 0 || obn < bn || (obn == bn && oin < in_))) { bn = obn; in_ = oin; }
    }
    if (!L) continue;
    if (cg == 0 && a < B) {
        s_p[al] = ip;
        s_n[al] = in_;
    }
    __syncthreads();
    // per-anchor loss: two canonical d2 chains, one thread each (cg 0 -> a-p, cg 1 -> a-n)
    float dd = 0.f;
    const bool valid = a < B && s_p[al] >= 0 && s_n[al] >= 0;
    if (valid && cg < 2) dd = d2_chain(E + (int64_t)a * d, E + (int64_t)(cg == 0 ? s_p[al] : s_n[al]) * d, d);
    const float dan = __shfl_down_sync(0xffffffffu, dd, 1);
    if (cg == 0 && a < B) {
        const int64_t o = (int64_t)g * B + a;
        if (valid) {
            const float v = __fadd_rn(__fsub_rn(dd, dan), margin);
            trip[o * 3] = a;
            trip[o * 3 + 1] = s_p[al];
            trip[o * 3 + 2] = s_n[al];
            lt[o] = v > 0.f ? v : 0.f;
        } else {
            trip[o * 3] = trip[o * 3 + 1] = trip[o * 3 + 2] = -1;
            lt[o] = 0.f;
        }
    }
    }  // group list
}

// per-triplet losses for a caller-provided list (one thread per (triplet, role))
__global__ void __launch_bounds__(256) triplet_losses_kernel(const float *__restrict__ emb, int d,
                                                             const int32_t *__restrict__ trip, int64_t T, float margin,
                                                             float *__restrict__ lt) {
    const int64_t i = (int64_t)blockIdx.x * 128 + (threadIdx.x >> 1);
    const int role = threadIdx.x & 1;
    float dd = 0.f;
    if (i < T) {
        const float *a = emb + (int64_t)trip[i * 3] * d;
        const float *o = emb + (int64_t)trip[i * 3 + 1 + role] * d;
        dd = d2_chain(a, o, d);
    }
    const float dan = __shfl_down_sync(0xffffffffu, dd, 1);
    if (i < T && role == 0) {
        const float v = __fadd_rn(__fsub_rn(dd, dan), margin);
        lt[i] = v > 0.f ? v : 0.f;
    }
}

// grid (ceil(B/8), G): one warp per gradient row.  T_in < 0: batch-hard layout (T slots = B per
// group, rows with trip[.] < 0 are skipped anchors); otherwise one group with T_in triplets.
__global__ void __launch_bounds__(256) triplet_grad_kernel(const float *__restrict__ emb, int B, int d,
                                                           const int32_t *__restrict__ trip, const float *__restrict__ lt,
                                                           int64_t T_in, float *__restrict__ loss,
                                                           float *__restrict__ grad, const int32_t *__restrict__ glist,
                                                           const int32_t *__restrict__ gcount) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nlist = glist ? *gcount : (int)gridDim.y;
    for (int li = blockIdx.y; li < nlist; li += gridDim.y) {
    const int g = glist ? glist[li] : li;
    const int64_t slots = T_in < 0 ? B : T_in;
    const int32_t *TR = trip + (int64_t)g * slots * 3;
    const float *LT = lt + (int64_t)g * slots;
    const float *E = emb + (int64_t)g * B * d;
    // number of real triplets and the loss mean (sequential, triplet order)
    int T = 0;
    for (int64_t t = lane; t < slots; t += 32) T += TR[t * 3] >= 0 ? 1 : 0;
    T = __reduce_add_sync(0xffffffffu, T);
    if (blockIdx.x == 0 && w == 0 && lane == 0 && loss) {
        float s = 0.f;
        for (int64_t t = 0; t < slots; ++t)
            if (TR[t * 3] >= 0) s = __fadd_rn(s, LT[t]);
        loss[g] = T > 0 ? __fdiv_rn(s, (float)T) : 0.f;
    }
    if (!grad) continue;
    const int r = blockIdx.x * 8 + w;
    if (r >= B) continue;
    const float scale = T > 0 ? __fdiv_rn(2.0f, (float)T) : 0.f;
    float *gr = grad + ((int64_t)g * B + r) * d;
    for (int c0 = 0; c0 < d; c0 += 32 * 8) {
        float acc[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) acc[u] = 0.f;
        for (int64_t t = 0; t < slots; ++t) {
            const int ia = TR[t * 3];
            if (ia < 0 || !(LT[t] > 0.f)) continue;
            const int ip = TR[t * 3 + 1], in_ = TR[t * 3 + 2];
            if (ia != r && ip != r && in_ != r) continue;
            const float *a = E + (int64_t)ia * d, *pp = E + (int64_t)ip * d, *nn = E + (int64_t)in_ * d;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int c = c0 + u * 32 + lane;
                if (c < d) {
                    if (ia == r) acc[u] = __fmaf_rn(scale, __fsub_rn(nn[c], pp[c]), acc[u]);
                    if (ip == r) acc[u] = __fmaf_rn(scale, __fsub_rn(pp[c], a[c]), acc[u]);
                    if (in_ == r) acc[u] = __fmaf_rn(scale, __fsub_rn(a[c], nn[c]), acc[u]);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int c = c0 + u * 32 + lane;
            if (c < d) gr[c] = acc[u];
        }
    }
    }  // group list
}

static size_t mine_smem(int B) { return ((size_t)B + MKC * (MA + 1) + (size_t)MKC * (B + 1)) * 4; }

template <int NCMAX>
static int launch_mine(const float *emb, const int32_t *labels, int64_t G, int B, int d, float margin, float eps,
                       int32_t *trip, float *lt, float *dist, const int32_t *glist, const int32_t *gcount,
                       cudaStream_t st) {
    const size_t smem = mine_smem(B);
    auto kern = bh_mine_kernel<NCMAX>;
    VQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)((B + MA - 1) / MA), (unsigned)G);
    kern<<<grid, MT, smem, st>>>(emb, labels, B, d, margin, eps, trip, lt, dist, glist, gcount);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

int launch_pdist(const float *A, int64_t B, int d, float eps, float *dist, cudaStream_t st) {
    if (B == 0) return VQ_OK;
    VQ_CHECK(B <= MBMAX, VQ_ERR_CAPACITY, "pdist: B=%lld exceeds %d", (long long)B, MBMAX);
    if (B <= 256) return launch_mine<32>(A, nullptr, 1, (int)B, d, 0.f, eps, nullptr, nullptr, dist, nullptr, nullptr, st);
    return launch_mine<MBMAX / 8>(A, nullptr, 1, (int)B, d, 0.f, eps, nullptr, nullptr, dist, nullptr, nullptr, st);
}

// CUDA-core path: any B <= 1024, any d.  glist/gcount: restrict to a device list of groups (G = list capacity).
int launch_triplet_batchhard_simt(const float *emb, const int32_t *labels, int64_t G, int64_t B, int d, float margin,
                                  float *loss, float *grad, int32_t *trip, float *scratch, const int32_t *glist,
                                  const int32_t *gcount, cudaStream_t st) {
    if (G == 0 || B == 0) return VQ_OK;
    VQ_CHECK(B <= MBMAX, VQ_ERR_CAPACITY, "triplet: B=%lld exceeds %d", (long long)B, MBMAX);
    VQ_CHECK(G <= 65535, VQ_ERR_CAPACITY, "triplet: G=%lld exceeds 65535 groups per call", (long long)G);
    if (B <= 256)
        VQ_TRY(launch_mine<32>(emb, labels, G, (int)B, d, margin, 1e-4f, trip, scratch, nullptr, glist, gcount, st));
    else
        VQ_TRY(launch_mine<MBMAX / 8>(emb, labels, G, (int)B, d, margin, 1e-4f, trip, scratch, nullptr, glist, gcount, st));
    dim3 grid2((unsigned)((B + 7) / 8), (unsigned)G);
    triplet_grad_kernel<<<grid2, 256, 0, st>>>(emb, (int)B, d, trip, scratch, -1, loss, grad, glist, gcount);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

int launch_triplet_given(const float *emb, int64_t B, int d, const int32_t *trip, int64_t T, float margin, float *loss,
                         float *grad, float *scratch, cudaStream_t st) {
    if (T > 0) {
        triplet_losses_kernel<<<(unsigned)((T + 127) / 128), 256, 0, st>>>(emb, d, trip, T, margin, scratch);
        VQ_CUDA(cudaGetLastError());
    }
    dim3 grid2((unsigned)((B + 7) / 8), 1);
    triplet_grad_kernel<<<grid2, 256, 0, st>>>(emb, (int)B, d, trip, scratch, T, loss, grad, nullptr, nullptr);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

}  // namespace vq
