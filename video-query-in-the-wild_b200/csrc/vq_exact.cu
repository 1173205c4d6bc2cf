// vq_exact.cu -- canonical (oracle-identical) CUDA-core scoring.
//   l2sq_all    : all-pairs d2 tile kernel, out[nq,n]; feeds the full-list AP reduction
//                 (vq_metrics.cu) and the k > kMaxKPrime search path.
//   exact_topk  : per-query exact streaming top-k over the whole gallery -- the fallback for
//                 queries the tensor-core pre-filter could not prove, and VQ_PREC_EXACT mode.
// Arithmetic: d2 = sequential fmaf chain over i (vq_common.cuh:canon_step), the restatement of
// faiss.IndexFlatL2.search's squared L2 (dataloader_baseline.py:1592-1597).
#include "vq_blocksel.cuh"
#include "vq_kernels.h"

namespace vq {

// ---------------------------------------------------------------------------------------
// all-pairs: 64 x 64 output tile per 256-thread block, 4 x 4 per thread, k staged 32 at a time
// ---------------------------------------------------------------------------------------
// k-block depth: 64 (half the barriers, but 124 registers and 51 KB of shared memory -> two blocks per SM instead of three)
// measured 4.98 ms against 3.48 ms at 4096 x 20480 x 512 -- occupancy matters more than the barrier count.
#ifndef VQ_L2SQ_AK
#define VQ_L2SQ_AK 32
#endif
static constexpr int AT = 64, AQ = 128, AK = VQ_L2SQ_AK, APAD = 4;
static constexpr size_t kL2Smem = sizeof(float) * AK * (AQ + APAD + AT + APAD);  // gallery rows / query rows / k per block tile

template <typename T>
__global__ void __launch_bounds__(256, AK == 64 ? 2 : 3) l2sq_all_kernel(const T *__restrict__ q, const T *__restrict__ g,
                                                          int64_t nq, int64_t n, int d, float *__restrict__ out) {
    // 128 queries x 64 gallery rows per block, 8 x 4 per thread: per k one thread reads 8 query and 4 gallery values
    // (three LDS.128) for 32 packed steps.  The {q, q} operand of a packed step is the scalar-broadcast form of FADD2
    // (R.F32): nothing is duplicated in shared memory -- with a 4 x 4 tile and duplicated queries the shared-memory
    // pipe was the limiter (ncu: l1tex 95 %, fma pipe 58 %).
    extern __shared__ __align__(16) uint8_t l2_sm[];  // dynamic: sQ[AK][AQ + APAD] | sG[AK][AT + APAD]
    float(*sQ)[AQ + APAD] = reinterpret_cast<float(*)[AQ + APAD]>(l2_sm);
    float(*sG)[AT + APAD] = reinterpret_cast<float(*)[AT + APAD]>(l2_sm + sizeof(float) * AK * (AQ + APAD));
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int64_t q0 = (int64_t)blockIdx.y * AQ, g0 = (int64_t)blockIdx.x * AT;
    unsigned long long acc[8][2];  // [query a][gallery pair]: {acc(a, 2p), acc(a, 2p+1)}
#pragma unroll
    for (int a = 0; a < 8; ++a) acc[a][0] = acc[a][1] = 0ull;
    // loaders: query row qr = tid / 2, 16 consecutive k; gallery row gr = tid / 4, 8 consecutive k.  The next k-block's
    // values are fetched into registers while this one is being multiplied; tile columns are XOR-swizzled by the k-group
    // so that the loader lanes of one row do not land in one bank.
    constexpr int QE = AK / 2, GE = AK / 4;            // k values a loader thread carries per k-block
    constexpr int QSH = AK == 64 ? 5 : 4, GSH = AK == 64 ? 4 : 3;
    static_assert(AK == 32 || AK == 64, "loader layout");
    const int qr = tid >> 1, qk = (tid & 1) * QE, q_sw = (tid & 1) << 4;
    const int gr = tid >> 2, gk = (tid & 3) * GE, g_sw = (tid & 3) << 3;
    float rq[QE], rg[GE];
    auto fetch = [&](int k0) {
#pragma unroll
        for (int e = 0; e < QE; ++e) {
            const int k = k0 + qk + e;
            rq[e] = (k < d && q0 + qr < nq) ? to_f32<T>(q[(q0 + qr) * d + k]) : 0.f;
        }
#pragma unroll
        for (int e = 0; e < GE; ++e) {
            const int k = k0 + gk + e;
            rg[e] = (k < d && g0 + gr < n) ? to_f32<T>(g[(g0 + gr) * d + k]) : 0.f;
        }
    };
    fetch(0);
    for (int k0 = 0; k0 < d; k0 += AK) {
        __syncthreads();
#pragma unroll
        for (int e = 0; e < QE; ++e) sQ[qk + e][qr ^ q_sw] = rq[e];
#pragma unroll
        for (int e = 0; e < GE; ++e) sG[gk + e][gr ^ g_sw] = rg[e];
        __syncthreads();
        if (k0 + AK < d) fetch(k0 + AK);
#pragma unroll
        for (int kk = 0; kk < AK; ++kk) {
            const int swq = (kk >> QSH) << 4, swg = (kk >> GSH) << 3;
            const float4 a0 = *reinterpret_cast<const float4 *>(&sQ[kk][(ty * 8) ^ swq]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&sQ[kk][((ty * 8) ^ swq) + 4]);
            const ulonglong2 b2 = *reinterpret_cast<const ulonglong2 *>(&sG[kk][(tx * 4) ^ swg]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
            for (int a = 0; a < 8; ++a) {
                unsigned long long qv;
                asm("mov.b64 %0, {%1, %1};" : "=l"(qv) : "f"(av[a]));
                acc[a][0] = canon_step2(acc[a][0], qv, b2.x);
                acc[a][1] = canon_step2(acc[a][1], qv, b2.y);
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        const int64_t qi = q0 + ty * 8 + a;
        if (qi >= nq) continue;
        const float r[4] = {__uint_as_float((uint32_t)acc[a][0]), __uint_as_float((uint32_t)(acc[a][0] >> 32)),
                            __uint_as_float((uint32_t)acc[a][1]), __uint_as_float((uint32_t)(acc[a][1] >> 32))};
        const int64_t gi0 = g0 + tx * 4;
        if (gi0 + 3 < n && ((qi * n + gi0) & 3) == 0) {
            *reinterpret_cast<float4 *>(out + qi * n + gi0) = make_float4(r[0], r[1], r[2], r[3]);
        } else {
#pragma unroll
            for (int b = 0; b < 4; ++b)
                if (gi0 + b < n) out[qi * n + gi0 + b] = r[b];
        }
    }
}

// Wide variant: 128 queries x 128 gallery rows per block, 8 x 8 per thread (64 packed chains): per k a thread reads 8 query
// and 8 gallery values (four LDS.128) for 64 packed steps -- 16 FP instructions per shared-memory load instead of 10.7;
// the 8 x 4 kernel above keeps the shared-memory pipe 80 % busy at 68 % of the FMA pipe.  A thread's gallery columns are
// tx*4 .. +3 and 64 + tx*4 .. +3, so that both reads of a half warp are 256 contiguous bytes.
static constexpr int WT = 128;
template <typename T>
__global__ void __launch_bounds__(256, 2) l2sq_all_wide_kernel(const T *__restrict__ q, const T *__restrict__ g, int64_t nq,
                                                               int64_t n, int d, float *__restrict__ out) {
    __shared__ __align__(16) float sQ[AK][AQ + APAD];
    __shared__ __align__(16) float sG[AK][WT + APAD];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int64_t q0 = (int64_t)blockIdx.y * AQ, g0 = (int64_t)blockIdx.x * WT;
    unsigned long long acc[8][4];  // [query a][gallery pair p]: pairs 0, 1 = columns tx*4 .. +3, pairs 2, 3 = 64 + tx*4 .. +3
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0ull;
    static_assert(AK == 32, "loader layout of the wide kernel");
    const int lr = tid >> 1, lk = (tid & 1) * 16, l_sw = (tid & 1) << 4;  // both tiles: row tid / 2, 16 consecutive k
    float rq[16], rg[16];
    auto fetch = [&](int k0) {
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            const int k = k0 + lk + e;
            rq[e] = (k < d && q0 + lr < nq) ? to_f32<T>(q[(q0 + lr) * d + k]) : 0.f;
            rg[e] = (k < d && g0 + lr < n) ? to_f32<T>(g[(g0 + lr) * d + k]) : 0.f;
        }
    };
    fetch(0);
    for (int k0 = 0; k0 < d; k0 += AK) {
        __syncthreads();
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            sQ[lk + e][lr ^ l_sw] = rq[e];
            sG[lk + e][lr ^ l_sw] = rg[e];
        }
        __syncthreads();
        if (k0 + AK < d) fetch(k0 + AK);
#pragma unroll
        for (int kk = 0; kk < AK; ++kk) {
            const int sw = (kk >> 4) << 4;
            const float4 a0 = *reinterpret_cast<const float4 *>(&sQ[kk][(ty * 8) ^ sw]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&sQ[kk][((ty * 8) ^ sw) + 4]);
            const ulonglong2 b0 = *reinterpret_cast<const ulonglong2 *>(&sG[kk][(tx * 4) ^ sw]);
            const ulonglong2 b1 = *reinterpret_cast<const ulonglong2 *>(&sG[kk][(64 + tx * 4) ^ sw]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
            for (int a = 0; a < 8; ++a) {
                unsigned long long qv;
                asm("mov.b64 %0, {%1, %1};" : "=l"(qv) : "f"(av[a]));
                acc[a][0] = canon_step2(acc[a][0], qv, b0.x);
                acc[a][1] = canon_step2(acc[a][1], qv, b0.y);
                acc[a][2] = canon_step2(acc[a][2], qv, b1.x);
                acc[a][3] = canon_step2(acc[a][3], qv, b1.y);
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        const int64_t qi = q0 + ty * 8 + a;
        if (qi >= nq) continue;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float r[4] = {__uint_as_float((uint32_t)acc[a][2 * h]), __uint_as_float((uint32_t)(acc[a][2 * h] >> 32)),
                                __uint_as_float((uint32_t)acc[a][2 * h + 1]), __uint_as_float((uint32_t)(acc[a][2 * h + 1] >> 32))};
            const int64_t gi0 = g0 + 64 * h + tx * 4;
            if (gi0 + 3 < n && ((qi * n + gi0) & 3) == 0) {
                *reinterpret_cast<float4 *>(out + qi * n + gi0) = make_float4(r[0], r[1], r[2], r[3]);
            } else {
#pragma unroll
                for (int b = 0; b < 4; ++b)
                    if (gi0 + b < n) out[qi * n + gi0 + b] = r[b];
            }
        }
    }
}

int launch_l2sq_all(const void *q, const void *g, int dtype, int64_t nq, int64_t n, int d, float *out,
                    cudaStream_t st) {
    if (nq == 0 || n == 0) return VQ_OK;
    dim3 grid((unsigned)((n + AT - 1) / AT), (unsigned)((nq + AQ - 1) / AQ));
    VQ_CHECK(grid.y <= 65535, VQ_ERR_INVALID, "l2sq_all: too many query tiles in one launch");
    if (!g_hooks.no_l2sq_wide) {  // default; NO_L2SQ_WIDE keeps the 8 x 4 kernel for A/B (3.48 vs 3.16 ms at 4096 x 20480 x 512)
        const dim3 gw((unsigned)((n + WT - 1) / WT), grid.y);
        if (dtype == VQ_F32)
            l2sq_all_wide_kernel<float><<<gw, 256, 0, st>>>((const float *)q, (const float *)g, nq, n, d, out);
        else
            l2sq_all_wide_kernel<__nv_bfloat16><<<gw, 256, 0, st>>>((const __nv_bfloat16 *)q, (const __nv_bfloat16 *)g, nq, n, d, out);
        VQ_CUDA(cudaGetLastError());
        return VQ_OK;
    }
    if (dtype == VQ_F32) {
        VQ_CUDA(cudaFuncSetAttribute(l2sq_all_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kL2Smem));
        l2sq_all_kernel<float><<<grid, 256, kL2Smem, st>>>((const float *)q, (const float *)g, nq, n, d, out);
    } else {
        VQ_CUDA(cudaFuncSetAttribute(l2sq_all_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kL2Smem));
        l2sq_all_kernel<__nv_bfloat16><<<grid, 256, kL2Smem, st>>>((const __nv_bfloat16 *)q, (const __nv_bfloat16 *)g,
                                                              nq, n, d, out);
    }
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

// ---------------------------------------------------------------------------------------
// exact streaming top-k, one block per query (grid-stride over the query list)
// ---------------------------------------------------------------------------------------
static constexpr int kExThreads = 256;
static constexpr int kExCap = 4096;  // key buffer; must be >= 2*k and leave kExThreads of headroom

template <typename T, bool SEGMAX>
__global__ void __launch_bounds__(kExThreads) exact_topk_kernel(const FallbackParams p) {
    extern __shared__ __align__(16) uint8_t sm[];
    uint64_t *keys = (uint64_t *)sm;          // [kExCap]
    float *qs = (float *)(keys + kExCap);     // [d]
    int *scratch = (int *)(qs + p.d);         // [40]; [36] = count, [37] = ignore list base
    int *ign = scratch + 40;                  // [kMaxIgnore]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // rows are scored 32 at a time through this warp's shared-memory tile (coalesced copies, vq_blocksel.cuh) when
    // they are whole 128-byte chunks
    const bool tiled = ((size_t)p.d * sizeof(T)) % 128 == 0 && (((uintptr_t)p.g) & 15) == 0;
    uint8_t *tile = (uint8_t *)((((uintptr_t)(ign + kMaxIgnore)) + 127) & ~(uintptr_t)127) + warp * kRescoreTileBytes;
    const int nlist = p.unproven_count ? *p.unproven_count : (int)p.nq;
    const int ksort = next_pow2(p.k);

    for (int li = blockIdx.x; li < nlist; li += gridDim.x) {
        const int64_t q = p.unproven ? p.unproven[li] : li;
        __syncthreads();
        const T *qrow = (const T *)p.q + q * p.d;
        for (int i = tid; i < p.d; i += kExThreads) qs[i] = to_f32<T>(qrow[i]);
        if (tid < kMaxIgnore) ign[tid] = (tid < p.n_ign) ? p.q_ignore[q * p.n_ign + tid] : -1;
        if (tid == 0) scratch[36] = 0;
        __syncthreads();
        uint64_t thr = ~0ull;  // keep keys < thr
        const int64_t nitems = SEGMAX ? p.nseg : p.n;
        const int per_round = SEGMAX ? kExThreads / 32 : kExThreads;
        for (int64_t base = 0; base < nitems; base += per_round) {
            uint64_t key = ~0ull;
            if (!SEGMAX) {
                const int64_t j = base + tid;
                bool want = j < p.n;
                if (want && p.n_ign > 0) {
                    const int v = p.col_video[j];
                    for (int z = 0; z < p.n_ign; ++z) want &= (v != ign[z]);
                }
                if (tiled) {
                    if (base + warp * 32 < p.n) {  // warp-uniform: at least one lane has a row; the others copy a valid one
                        const float d2v = warp_canon_d2_rows<T>(qs, (const T *)p.g + (j < p.n ? j : p.n - 1) * p.d, p.d, tile, lane);
                        if (want) key = make_key_asc(d2v, (uint32_t)j);
                    }
                } else if (want) {
                    key = make_key_asc(canon_d2_row<T>(qs, (const T *)p.g + j * p.d, p.d), (uint32_t)j);
                }
            } else {
                const int64_t v = base + warp;
                if (v < p.nseg) {
                    const int64_t s0 = p.seg_start[v], s1 = p.seg_start[v + 1];
                    bool masked = false;
                    if (p.n_ign > 0 && s0 < s1) {
                        const int vid = p.col_video[s0];
                        for (int z = 0; z < p.n_ign; ++z) masked |= (vid == ign[z]);
                    }
                    uint64_t best = ~0ull;
                    if (!masked && tiled) {
                        for (int64_t r0 = s0; r0 < s1; r0 += 32) {
                            const int64_t rr = r0 + lane;
                            const float d2v = warp_canon_d2_rows<T>(qs, (const T *)p.g + (rr < s1 ? rr : r0) * p.d, p.d, tile, lane);
                            const uint64_t kk = rr < s1 ? make_key_asc(d2v, (uint32_t)rr) : ~0ull;
                            best = kk < best ? kk : best;
                        }
                    } else if (!masked)
                        for (int64_t rr = s0 + lane; rr < s1; rr += 32) {
                            const uint64_t kk =
                                make_key_asc(canon_d2_row<T>(qs, (const T *)p.g + rr * p.d, p.d), (uint32_t)rr);
                            best = kk < best ? kk : best;
                        }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        const uint64_t other = __shfl_xor_sync(0xffffffffu, best, o);
                        best = other < best ? other : best;
                    }
                    if (lane == 0) key = best;
                }
            }
            if (key < thr) keys[atomicAdd(&scratch[36], 1)] = key;
            __syncthreads();
            const int cnt = scratch[36];
            if (cnt > kExCap - kExThreads) {  // uniform
                const uint64_t X = block_kth_smallest(keys, cnt, p.k, scratch);
                // keep keys <= X (exactly k: keys are unique)
                uint64_t mine[kExCap / kExThreads];
                int nm = 0;
                for (int i = tid; i < cnt; i += kExThreads) {
                    const uint64_t kk = keys[i];
                    if (kk <= X) mine[nm++] = kk;
                }
                __syncthreads();
                if (tid == 0) scratch[36] = 0;
                __syncthreads();
                const int b = atomicAdd(&scratch[36], nm);
                for (int i = 0; i < nm; ++i) keys[b + i] = mine[i];
                __syncthreads();
                thr = X + 1;  // strictly better than the current k-th, or it cannot enter the top-k
            }
        }
        __syncthreads();
        int cnt = scratch[36];
        if (cnt > p.k) {
            const uint64_t X = block_kth_smallest(keys, cnt, p.k, scratch);
            uint64_t mine[kExCap / kExThreads];
            int nm = 0;
            for (int i = tid; i < cnt; i += kExThreads) {
                const uint64_t kk = keys[i];
                if (kk <= X) mine[nm++] = kk;
            }
            __syncthreads();
            if (tid == 0) scratch[36] = 0;
            __syncthreads();
            const int b = atomicAdd(&scratch[36], nm);
            for (int i = 0; i < nm; ++i) keys[b + i] = mine[i];
            __syncthreads();
            cnt = scratch[36];
        }
        for (int i = cnt + tid; i < ksort; i += kExThreads) keys[i] = ~0ull;
        __syncthreads();
        block_bitonic_sort(keys, ksort);
        for (int i = tid; i < p.k; i += kExThreads) {
            const uint64_t kk = keys[i];
            const bool ok = i < cnt;
            const uint32_t row = (uint32_t)kk;
            p.out_d2[q * p.k + i] = ok ? ordered_to_f32((uint32_t)(kk >> 32)) : INFINITY;
            p.out_idx[q * p.k + i] = ok ? (int64_t)row : -1;
            if (SEGMAX && p.out_seg) p.out_seg[q * p.k + i] = ok ? (int64_t)p.seg_of_row[row] : -1;
        }
    }
}

template <typename T, bool SEGMAX>
static int launch_ex(const FallbackParams &p, int grid, cudaStream_t st) {
    const size_t smem = (size_t)kExCap * 8 + (size_t)p.d * 4 + 40 * 4 + kMaxIgnore * 4 + 16 + 128 +
                        (size_t)(kExThreads / 32) * kRescoreTileBytes;
    auto kern = exact_topk_kernel<T, SEGMAX>;
    VQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, kExThreads, smem, st>>>(p);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

int launch_exact_topk(const FallbackParams &p, int grid, cudaStream_t st) {
    if (p.nq == 0) return VQ_OK;
    VQ_CHECK(p.k >= 1 && 2 * p.k <= kExCap - kExThreads, VQ_ERR_CAPACITY, "exact_topk: k=%d too large", p.k);
    const bool seg = p.seg_start != nullptr;
    if (p.dtype == VQ_F32) return seg ? launch_ex<float, true>(p, grid, st) : launch_ex<float, false>(p, grid, st);
    return seg ? launch_ex<__nv_bfloat16, true>(p, grid, st) : launch_ex<__nv_bfloat16, false>(p, grid, st);
}

}  // namespace vq
