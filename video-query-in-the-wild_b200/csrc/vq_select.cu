// vq_select.cu -- second half of kernel (2): turns the pre-filter candidates of
// vq_gemm_topk.cu into the exact answer.
//   per query (one 128-thread block):
//     1. gather the <= k' coarse keys of every gallery segment of the query's group, keep the k' best (bisection select)
//     2. re-score the survivors with the canonical fp32 chain (oracle-identical bits); in
//        segment-max mode re-score every moment of each surviving video and take its exact best
//     3. bitonic-sort by (d2 asc, row asc)  -> the reference's ranked order
//        (faiss.IndexFlatL2.search, dataloader_baseline.py:1592-1597, ties -> lowest index)
//     4. prove completeness: everything the pre-filter dropped has coarse score <= tau, hence
//        d2 >= |q|^2 - tau - eps; if that bound does not clear the k-th exact d2 the query is
//        queued for the exact fallback (vq_exact.cu)
#include "vq_blocksel.cuh"
#include "vq_kernels.h"

namespace vq {

static constexpr int kSelThreads = 128;

// gather buffer: every candidate key of the query's segments; once the cut has kept k' of them the rest of the
// buffer holds the four warps' re-score tiles (128-byte aligned), so it is at least that large
__host__ __device__ __forceinline__ size_t select_keys_bytes(int segs_per_group, int kprime) {
    const size_t gather = (size_t)segs_per_group * kprime * 8;
    const size_t rescore = (size_t)kprime * 8 + 128 + (size_t)(kSelThreads / 32) * kRescoreTileBytes;
    return ((gather > rescore ? gather : rescore) + 15) & ~(size_t)15;
}

template <typename T, bool SEGMAX>
__global__ void __launch_bounds__(kSelThreads) select_kernel(const SelectParams p) {
    extern __shared__ __align__(16) uint8_t sm[];
    const int ksort = next_pow2(p.kprime);
    uint64_t *keys = (uint64_t *)sm;                     // [segs_per_group * k'] coarse keys; after the cut: k' keys + re-score tiles
    uint64_t *ex = (uint64_t *)(sm + select_keys_bytes(p.segs_per_group, p.kprime));  // [ksort] exact keys
    uint32_t *aux = (uint32_t *)(ex + ksort);            // [ksort] segmax: video per exact key slot
    float *qs = (float *)(aux + ksort);                  // [d]
    int *scratch = (int *)(qs + p.d);                    // [40]
    int *cnts = scratch + 40;                            // [nsg] candidates taken from each segment
    int *offs = cnts + p.segs_per_group;                 // [nsg + 1]
    int64_t *urs = (int64_t *)(offs + p.segs_per_group + 1 + ((p.d + 1) & 1));  // [nsg] candidate-list rows (8-byte aligned)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t q = blockIdx.x;
    const int qb = (int)(q / kTileM), r = (int)(q % kTileM);
    const int grp = qb / p.group_size, member = qb % p.group_size;
    const int32_t *segs = p.group_segs + (int64_t)grp * p.segs_per_group;  // this query's segments
    const int nsg = p.segs_per_group;
    const int E = nsg * p.kprime;
    const T *qrow = (const T *)p.q + q * p.d;

    // ---- 1. gather: per-segment counts first, then one flattened copy with independent loads (the lists of the
    // later segments are nearly empty once tau has tightened, so the compacted total is a small multiple of k') ----
    for (int i = tid; i < p.d; i += kSelThreads) qs[i] = to_f32<T>(qrow[i]);
    // A segment whose list CONTINUED its predecessor's (cand_cum, vq_gemm_topk.cu) already holds the best k' of every
    // earlier gallery chunk: only the segments from the last such one on are read -- normally just the final one.
    __shared__ int s_first;
    if (tid == 0) s_first = 0;
    __syncthreads();
    for (int c = tid; c < nsg; c += kSelThreads) {
        const int64_t ur = ((int64_t)segs[c] * p.group_size + member) * kTileM + r;
        urs[c] = ur;
        if (c > 0 && p.cand_cum[ur]) atomicMax(&s_first, c);
    }
    __syncthreads();
    const int first = s_first;
    float tau = -INFINITY;
    for (int c = tid; c < nsg; c += kSelThreads) {
        const int64_t ur = urs[c];
        cnts[c] = c >= first ? min(p.cand_cnt[ur], p.kprime) : 0;
        if (c >= first) tau = fmaxf(tau, p.cand_tau[ur]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tau = fmaxf(tau, __shfl_xor_sync(0xffffffffu, tau, o));
    if (lane == 0) scratch[32 + warp] = __float_as_int(tau);
    __syncthreads();
    tau = fmaxf(fmaxf(__int_as_float(scratch[32]), __int_as_float(scratch[33])),
                fmaxf(__int_as_float(scratch[34]), __int_as_float(scratch[35])));
    if (warp == 0) {  // exclusive scan of the counts
        int run = 0;
        for (int c0 = 0; c0 < nsg; c0 += 32) {
            const int v = c0 + lane < nsg ? cnts[c0 + lane] : 0;
            int inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += u;
            }
            if (c0 + lane < nsg) offs[c0 + lane] = run + inc - v;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) offs[nsg] = run;
    }
    __syncthreads();
    for (int idx = tid; idx < E; idx += kSelThreads) {
        const int c = idx / p.kprime, i = idx - c * p.kprime;
        if (i < cnts[c]) keys[offs[c] + i] = p.cand[urs[c] * kCandCap + i];
    }
    __syncthreads();
    int m = offs[nsg];
    if (m > p.kprime) {
        // keep the k' best coarse keys; ties at the cut resolved by the full key (score, then row)
        const int M0 = m;
        const uint32_t Th = block_kth_largest_hi(keys, M0, p.kprime, scratch);
        int gt = 0;
        for (int i = tid; i < M0; i += kSelThreads) gt += (uint32_t)(keys[i] >> 32) > Th ? 1 : 0;
        gt = block_sum(gt, scratch);
        // among hi == Th keep the (k' - gt) largest full keys: find that key by a second bisection
        const int need_eq = p.kprime - gt;
        uint32_t Tl = 0;
        for (int bit = 31; bit >= 0; --bit) {
            const uint32_t cnd = Tl | (1u << bit);
            int cn = 0;
            for (int i = tid; i < M0; i += kSelThreads)
                cn += ((uint32_t)(keys[i] >> 32) == Th && (uint32_t)keys[i] >= cnd) ? 1 : 0;
            cn = block_sum(cn, scratch);
            if (cn >= need_eq) Tl = cnd;
        }
        const uint64_t cut = ((uint64_t)Th << 32) | Tl;  // keep keys >= cut : exactly k' (keys are unique)
        __syncthreads();
        if (tid == 0) scratch[36] = 0;
        __syncthreads();
        // compact through registers (E <= kSelThreads * 64 guaranteed by the host)
        uint64_t mine[64];
        int nmine = 0;
        for (int i = tid; i < M0; i += kSelThreads) {
            uint64_t kk = keys[i];
            if (kk >= cut && nmine < 64) mine[nmine++] = kk;
        }
        __syncthreads();
        int base = atomicAdd(&scratch[36], nmine);
        for (int i = 0; i < nmine; ++i) keys[base + i] = mine[i];
        __syncthreads();
        m = scratch[36];
        tau = fmaxf(tau, ordered_to_f32(Th));
    }

    if (SEGMAX && p.pair_vid) {
        // video-major re-scoring (below): hand the surviving (query, video) pairs over instead of re-scoring them here
        for (int i = tid; i < p.kprime; i += kSelThreads) {
            int v = -1, one = -1;
            if (i < m) {
                const uint32_t row = key_desc_idx(keys[i]);
                v = p.seg_of_row[row];
                // the GEMM proved the coarse arg-max alone at the top of its video: one canonical score settles the pair
                if (p.seg_amb && !((p.seg_amb[q * p.amb_words + (v >> 5)] >> (v & 31)) & 1u)) {
                    one = (int)row;
                    v = -1;
                } else {
                    atomicAdd(p.vcount + v, 1);
                }
            }
            p.pair_vid[q * p.kprime + i] = v;
            if (p.pair_one) p.pair_one[q * p.kprime + i] = one;
        }
        if (tid == 0) {
            p.pair_m[q] = m;
            p.pair_tau[q] = tau;
        }
        return;
    }

    // ---- 2. canonical re-scoring ----
    for (int i = tid; i < ksort; i += kSelThreads) {
        ex[i] = ~0ull;
        aux[i] = 0xffffffffu;
    }
    __syncthreads();
    // rows are re-scored 32 at a time through this warp's shared-memory tile (coalesced copies, vq_blocksel.cuh)
    // when they are whole 128-byte chunks; the space is the part of the gather buffer the cut has freed
    const bool tiled = ((size_t)p.d * sizeof(T)) % 128 == 0 && (((uintptr_t)p.g) & 15) == 0;
    uint8_t *tile = (uint8_t *)((((uintptr_t)(keys + p.kprime)) + 127) & ~(uintptr_t)127) + warp * kRescoreTileBytes;
    if (!SEGMAX && tiled) {
        for (int i0 = warp * 32; i0 < m; i0 += kSelThreads) {
            const int i = i0 + lane;
            const uint32_t row = key_desc_idx(keys[i < m ? i : i0]);
            const float d2 = warp_canon_d2_rows<T>(qs, (const T *)p.g + (int64_t)row * p.d, p.d, tile, lane);
            if (i < m) ex[i] = make_key_asc(d2, row);
        }
    } else if (!SEGMAX) {
        for (int i = tid; i < m; i += kSelThreads) {
            const uint32_t row = key_desc_idx(keys[i]);
            const float d2 = canon_d2_row<T>(qs, (const T *)p.g + (int64_t)row * p.d, p.d);
            ex[i] = make_key_asc(d2, row);
        }
    } else if (tiled) {
        for (int i = warp; i < m; i += kSelThreads / 32) {
            const uint32_t arow = key_desc_idx(keys[i]);
            const int v = p.seg_of_row[arow];
            const int64_t s0 = p.seg_start[v], s1 = p.seg_start[v + 1];
            uint64_t best = ~0ull;
            for (int64_t r0 = s0; r0 < s1; r0 += 32) {
                const int64_t rr = r0 + lane;
                const float d2 = warp_canon_d2_rows<T>(qs, (const T *)p.g + (rr < s1 ? rr : r0) * p.d, p.d, tile, lane);
                const uint64_t kk = rr < s1 ? make_key_asc(d2, (uint32_t)rr) : ~0ull;
                best = kk < best ? kk : best;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const uint64_t other = __shfl_xor_sync(0xffffffffu, best, o);
                best = other < best ? other : best;
            }
            if (lane == 0) ex[i] = best;
        }
    } else {
        for (int i = warp; i < m; i += kSelThreads / 32) {
            const uint32_t arow = key_desc_idx(keys[i]);
            const int v = p.seg_of_row[arow];
            const int64_t s0 = p.seg_start[v], s1 = p.seg_start[v + 1];
            uint64_t best = ~0ull;
            for (int64_t rr = s0 + lane; rr < s1; rr += 32) {
                const float d2 = canon_d2_row<T>(qs, (const T *)p.g + rr * p.d, p.d);
                const uint64_t kk = make_key_asc(d2, (uint32_t)rr);
                best = kk < best ? kk : best;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const uint64_t other = __shfl_xor_sync(0xffffffffu, best, o);
                best = other < best ? other : best;
            }
            if (lane == 0) ex[i] = best;  // rows of lower videos are lower: row order == video order
        }
    }
    __syncthreads();

    // ---- 3. sort ----
    block_bitonic_sort(ex, ksort);

    // ---- 4. prove + emit ----
    float qn = 0.f;
    for (int i = tid; i < p.d; i += kSelThreads) qn = fmaf(qs[i], qs[i], qn);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) qn += __shfl_xor_sync(0xffffffffu, qn, o);
    __shared__ float qn_w[kSelThreads / 32];
    if (lane == 0) qn_w[warp] = qn;
    __syncthreads();
    qn = 0.f;
    for (int w = 0; w < kSelThreads / 32; ++w) qn += qn_w[w];

    const int kout = p.k;
    bool proven = true;
    if (tau != -INFINITY) {
        if (m < kout) {
            proven = false;
        } else {
            const float gmax2 = __uint_as_float(*p.gmax_norm2_bits);
            const float eps = p.eps_rel * sqrtf(qn) * sqrtf(gmax2) + p.eps_abs * (qn + gmax2);
            const float d2k = ordered_to_f32((uint32_t)(ex[kout - 1] >> 32));
            const float lower = qn - tau - eps;  // every dropped row has d2 >= lower
            proven = lower > d2k;
        }
    }
    for (int i = tid; i < kout; i += kSelThreads) {
        const uint64_t kk = ex[i];
        const bool ok = (i < m) && kk != ~0ull;
        const uint32_t row = (uint32_t)kk;
        p.out_d2[q * kout + i] = ok ? ordered_to_f32((uint32_t)(kk >> 32)) : INFINITY;
        p.out_idx[q * kout + i] = ok ? (int64_t)row : -1;
        if (SEGMAX && p.out_seg) p.out_seg[q * kout + i] = ok ? (int64_t)p.seg_of_row[row] : -1;
    }
    if (!proven && tid == 0) {
        const int slot = atomicAdd(p.unproven_count, 1);
        p.unproven[slot] = (int32_t)q;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// SEGMAX, video-major re-scoring.  Block-per-query re-scoring reads 128 candidate videos x all their moments per query
// -- 69 GB of rows for 4 096 queries x (20 480 videos x 64 moments x 2 KB), far outside L2: HBM-bound.  Here the
// surviving (query, video) pairs are bucketed by VIDEO (count in the select kernel, scan, fill), each video's rows are
// read once and scored against all the queries that kept it with the all-pairs register tile (packed canonical
// chains, same bits), and a last per-query kernel sorts the exact keys and proves completeness.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) scan_i32_kernel(const int32_t *__restrict__ in, int32_t *__restrict__ out, int64_t n) {
    __shared__ int part[1024];
    const int tid = threadIdx.x;
    const int64_t per = (n + 1023) / 1024, b = tid * per, e = b + per < n ? b + per : n;
    int s = 0;
    for (int64_t i = b; i < e; ++i) s += in[i];
    part[tid] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = tid >= o ? part[tid - o] : 0;
        __syncthreads();
        part[tid] += v;
        __syncthreads();
    }
    int run = part[tid] - s;
    for (int64_t i = b; i < e; ++i) {
        const int c = in[i];
        out[i] = run;
        run += c;
    }
    if (tid == 1023) out[n] = part[1023];
}

__global__ void __launch_bounds__(256) segmax_fill_kernel(const int32_t *__restrict__ pair_vid, int64_t npairs,
                                                          const int32_t *__restrict__ voff, int32_t *__restrict__ vfill,
                                                          int32_t *__restrict__ plist) {
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= npairs) return;
    const int v = pair_vid[i];
    if (v >= 0) plist[voff[v] + atomicAdd(vfill + v, 1)] = (int32_t)i;
}

// one 128-thread block per video: 32 queries x 64 moments per tile, 4 x 4 per thread, k staged 32 at a time
static constexpr int VQT = 32, VMT = 64, VK = 32, VPAD = 4, VTHREADS = 128;

template <typename T>
__global__ void __launch_bounds__(VTHREADS) segmax_rescore_kernel(const SelectParams p, const int32_t *__restrict__ voff,
                                                                  const int32_t *__restrict__ plist, uint64_t *__restrict__ exact) {
    __shared__ __align__(16) float sQ[VK][VQT + VPAD];
    __shared__ __align__(16) float sG[VK][VMT + VPAD];
    __shared__ int s_pair[VQT];
    const int v = blockIdx.x, tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int p0 = voff[v], p1 = voff[v + 1];
    if (p0 == p1) return;
    const int64_t s0 = p.seg_start[v], s1 = p.seg_start[v + 1];
    const T *gq = (const T *)p.q, *gg = (const T *)p.g;
    const int d = p.d;
    // loaders: query slot qr = tid / 4 (32 slots), 8 consecutive k; moment row gr = tid / 2 (64 rows), 16 consecutive k
    const int qr = tid >> 2, qk = (tid & 3) * 8, q_sw = (tid & 3) << 3;
    const int gr = tid >> 1, gk = (tid & 1) * 16, g_sw = (tid & 1) << 4;
    for (int pc = p0; pc < p1; pc += VQT) {  // chunk of the video's queries
        __syncthreads();
        if (tid < VQT) s_pair[tid] = pc + tid < p1 ? plist[pc + tid] : -1;
        __syncthreads();
        const int pr = s_pair[qr];
        const int64_t qrow = pr >= 0 ? (int64_t)(pr / p.kprime) : -1;
        uint64_t best[4] = {~0ull, ~0ull, ~0ull, ~0ull};  // running (d2, row) minima of queries ty*4 .. ty*4+3 over lane tx's moments
        for (int64_t mc = s0; mc < s1; mc += VMT) {  // chunk of the video's moments
            unsigned long long acc[4][2];
#pragma unroll
            for (int a = 0; a < 4; ++a) acc[a][0] = acc[a][1] = 0ull;
            float rq[8], rg[16];
            auto fetch = [&](int k0) {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const int k = k0 + qk + e;
                    rq[e] = (k < d && qrow >= 0) ? to_f32<T>(gq[qrow * d + k]) : 0.f;
                }
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    const int k = k0 + gk + e;
                    rg[e] = (k < d && mc + gr < s1) ? to_f32<T>(gg[(mc + gr) * d + k]) : 0.f;
                }
            };
            fetch(0);
            for (int k0 = 0; k0 < d; k0 += VK) {
                __syncthreads();
#pragma unroll
                for (int e = 0; e < 8; ++e) sQ[qk + e][qr ^ q_sw] = rq[e];
#pragma unroll
                for (int e = 0; e < 16; ++e) sG[gk + e][gr ^ g_sw] = rg[e];
                __syncthreads();
                if (k0 + VK < d) fetch(k0 + VK);
#pragma unroll
                for (int kk = 0; kk < VK; ++kk) {
                    const int swq = (kk >> 3) << 3, swg = (kk >> 4) << 4;
                    const float4 a4 = *reinterpret_cast<const float4 *>(&sQ[kk][(ty * 4) ^ swq]);
                    const ulonglong2 b2 = *reinterpret_cast<const ulonglong2 *>(&sG[kk][(tx * 4) ^ swg]);
                    const float av[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
                    for (int a = 0; a < 4; ++a) {
                        unsigned long long qv;
                        asm("mov.b64 %0, {%1, %1};" : "=l"(qv) : "f"(av[a]));
                        acc[a][0] = canon_step2(acc[a][0], qv, b2.x);
                        acc[a][1] = canon_step2(acc[a][1], qv, b2.y);
                    }
                }
            }
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const float r[4] = {__uint_as_float((uint32_t)acc[a][0]), __uint_as_float((uint32_t)(acc[a][0] >> 32)),
                                    __uint_as_float((uint32_t)acc[a][1]), __uint_as_float((uint32_t)(acc[a][1] >> 32))};
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const int64_t row = mc + tx * 4 + b;
                    if (row < s1) {
                        const uint64_t kk = make_key_asc(r[b], (uint32_t)row);
                        best[a] = kk < best[a] ? kk : best[a];
                    }
                }
            }
        }
#pragma unroll
        for (int a = 0; a < 4; ++a) {  // the 16 lanes tx = 0..15 of a half warp hold one query's partial minima
            uint64_t b = best[a];
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                const uint64_t other = __shfl_xor_sync(0xffffffffu, b, o);
                b = other < b ? other : b;
            }
            const int pr2 = s_pair[ty * 4 + a];
            if (tx == 0 && pr2 >= 0) exact[pr2] = b;
        }
    }
}

// pairs whose best moment is known (pair_one): one canonical score, a thread per pair; one block per query, whose row
// is staged once in shared memory (fp32) and read by broadcast -- the kernel is bound by the L2 -> SM traffic of the
// uncoalesced gallery rows, and the query row was half of it
template <typename T>
__global__ void __launch_bounds__(512) segmax_single_kernel(const SelectParams p, const int32_t *__restrict__ pair_one,
                                                            uint64_t *__restrict__ exact) {
    extern __shared__ __align__(16) float s_q[];
    const int64_t q = blockIdx.x;
    const T *qr = (const T *)p.q + q * p.d;
    for (int k = threadIdx.x; k < p.d; k += blockDim.x) s_q[k] = to_f32<T>(qr[k]);
    __syncthreads();
    if ((int)threadIdx.x >= p.kprime) return;
    const int64_t i = q * p.kprime + threadIdx.x;
    const int row = pair_one[i];
    if (row < 0) return;
    const T *gr = (const T *)p.g + (int64_t)row * p.d;
    float acc = 0.f;
    int k = 0;
    // 32 bytes per load (LDG.256, sm_100): the row reads are uncoalesced -- one sector per lane and request -- and what
    // bounds the kernel is the number of such requests, so a whole sector per request halves them
    if ((((uintptr_t)gr) & 31) == 0 && ((size_t)p.d * sizeof(T)) % 32 == 0) {
        constexpr int E = 32 / (int)sizeof(T);  // elements per load
#pragma unroll 2
        for (; k < p.d; k += E) {
            uint32_t w[8];
            asm volatile("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                         : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                         : "l"(gr + k));
            if (sizeof(T) == 4) {
                const float4 a0 = *reinterpret_cast<const float4 *>(s_q + k), a1 = *reinterpret_cast<const float4 *>(s_q + k + 4);
                const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
                for (int e = 0; e < 8; ++e) acc = canon_step(acc, av[e], __uint_as_float(w[e]));
            } else {
#pragma unroll
                for (int e = 0; e < 8; e += 2) {
                    const float4 a = *reinterpret_cast<const float4 *>(s_q + k + 2 * e);
                    acc = canon_step(acc, a.x, __uint_as_float(w[e] << 16));
                    acc = canon_step(acc, a.y, __uint_as_float(w[e] & 0xffff0000u));
                    acc = canon_step(acc, a.z, __uint_as_float(w[e + 1] << 16));
                    acc = canon_step(acc, a.w, __uint_as_float(w[e + 1] & 0xffff0000u));
                }
            }
        }
    } else if (sizeof(T) == 4 && (p.d & 3) == 0) {
        const float4 *g4 = reinterpret_cast<const float4 *>(gr);
#pragma unroll 4
        for (; k < p.d; k += 4) {
            const float4 b = __ldg(g4 + (k >> 2)), a = *reinterpret_cast<const float4 *>(s_q + k);
            acc = canon_step(acc, a.x, b.x);
            acc = canon_step(acc, a.y, b.y);
            acc = canon_step(acc, a.z, b.z);
            acc = canon_step(acc, a.w, b.w);
        }
    } else if (sizeof(T) == 2 && (p.d & 7) == 0) {  // bf16 rows: 16 bytes = 8 elements per load (a bf16 is the top half of its f32)
        const uint4 *g8 = reinterpret_cast<const uint4 *>(gr);
#pragma unroll 4
        for (; k < p.d; k += 8) {
            const uint4 b = __ldg(g8 + (k >> 3));
            const float4 a0 = *reinterpret_cast<const float4 *>(s_q + k), a1 = *reinterpret_cast<const float4 *>(s_q + k + 4);
            acc = canon_step(acc, a0.x, __uint_as_float(b.x << 16));
            acc = canon_step(acc, a0.y, __uint_as_float(b.x & 0xffff0000u));
            acc = canon_step(acc, a0.z, __uint_as_float(b.y << 16));
            acc = canon_step(acc, a0.w, __uint_as_float(b.y & 0xffff0000u));
            acc = canon_step(acc, a1.x, __uint_as_float(b.z << 16));
            acc = canon_step(acc, a1.y, __uint_as_float(b.z & 0xffff0000u));
            acc = canon_step(acc, a1.z, __uint_as_float(b.w << 16));
            acc = canon_step(acc, a1.w, __uint_as_float(b.w & 0xffff0000u));
        }
    }
    for (; k < p.d; ++k) acc = canon_step(acc, s_q[k], to_f32<T>(gr[k]));
    exact[i] = make_key_asc(acc, (uint32_t)row);
}

// row_eps[q] = the proven bound of |coarse - (|q|^2 - d2)| for query q (error_model), a hair generous in |q|^2
template <typename T>
__global__ void __launch_bounds__(256) segmax_row_eps_kernel(const SelectParams p, float *__restrict__ row_eps) {
    const int64_t q = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (q >= p.nq) return;
    const T *qr = (const T *)p.q + q * p.d;
    float qn = 0.f;
    for (int i = lane; i < p.d; i += 32) {
        const float x = to_f32<T>(qr[i]);
        qn = fmaf(x, x, qn);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) qn += __shfl_xor_sync(0xffffffffu, qn, o);
    qn *= 1.0001f;
    const float gmax2 = __uint_as_float(*p.gmax_norm2_bits);
    if (lane == 0) row_eps[q] = 1.0001f * (p.eps_rel * sqrtf(qn) * sqrtf(gmax2) + p.eps_abs * (qn + gmax2));
}

int launch_segmax_row_eps(const SelectParams &p, float *row_eps, cudaStream_t st) {
    if (p.nq == 0) return VQ_OK;
    if (p.dtype == VQ_F32)
        segmax_row_eps_kernel<float><<<(unsigned)((p.nq + 7) / 8), 256, 0, st>>>(p, row_eps);
    else
        segmax_row_eps_kernel<__nv_bfloat16><<<(unsigned)((p.nq + 7) / 8), 256, 0, st>>>(p, row_eps);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

__global__ void __launch_bounds__(kSelThreads) segmax_finish_kernel(const SelectParams p, const uint64_t *__restrict__ exact) {
    extern __shared__ __align__(16) uint8_t sm[];
    const int ksort = next_pow2(p.kprime);
    uint64_t *ex = (uint64_t *)sm;
    __shared__ float qn_w[kSelThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t q = blockIdx.x;
    const int m = p.pair_m[q];
    const float tau = p.pair_tau[q];
    for (int i = tid; i < ksort; i += kSelThreads) ex[i] = i < m ? exact[q * p.kprime + i] : ~0ull;
    float qn = 0.f;
    if (p.dtype == VQ_F32) {
        const float *qrow = (const float *)p.q + q * p.d;
        for (int i = tid; i < p.d; i += kSelThreads) qn = fmaf(qrow[i], qrow[i], qn);
    } else {
        const __nv_bfloat16 *qrow = (const __nv_bfloat16 *)p.q + q * p.d;
        for (int i = tid; i < p.d; i += kSelThreads) {
            const float x = __bfloat162float(qrow[i]);
            qn = fmaf(x, x, qn);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) qn += __shfl_xor_sync(0xffffffffu, qn, o);
    if (lane == 0) qn_w[warp] = qn;
    __syncthreads();
    block_bitonic_sort(ex, ksort);
    qn = 0.f;
    for (int w = 0; w < kSelThreads / 32; ++w) qn += qn_w[w];
    const int kout = p.k;
    bool proven = true;
    if (tau != -INFINITY) {
        if (m < kout) {
            proven = false;
        } else {
            const float gmax2 = __uint_as_float(*p.gmax_norm2_bits);
            const float eps = p.eps_rel * sqrtf(qn) * sqrtf(gmax2) + p.eps_abs * (qn + gmax2);
            const float d2k = ordered_to_f32((uint32_t)(ex[kout - 1] >> 32));
            proven = qn - tau - eps > d2k;  // every dropped video's best moment has d2 >= qn - tau - eps
        }
    }
    for (int i = tid; i < kout; i += kSelThreads) {
        const uint64_t kk = ex[i];
        const bool ok = (i < m) && kk != ~0ull;
        const uint32_t row = (uint32_t)kk;
        p.out_d2[q * kout + i] = ok ? ordered_to_f32((uint32_t)(kk >> 32)) : INFINITY;
        p.out_idx[q * kout + i] = ok ? (int64_t)row : -1;
        if (p.out_seg) p.out_seg[q * kout + i] = ok ? (int64_t)p.seg_of_row[row] : -1;
    }
    if (!proven && tid == 0) {
        const int slot = atomicAdd(p.unproven_count, 1);
        p.unproven[slot] = (int32_t)q;
    }
}

size_t select_smem_bytes(const SelectParams &p) {
    const size_t ks = (size_t)next_pow2(p.kprime);
    return select_keys_bytes(p.segs_per_group, p.kprime) + ks * 8 + ks * 4 + (size_t)p.d * 4 + 40 * 4 + ((size_t)2 * p.segs_per_group + 2) * 4 +
           (size_t)p.segs_per_group * 8 + 16;
}

template <typename T, bool SEGMAX>
static int launch_sel(const SelectParams &p, cudaStream_t st) {
    const size_t smem = select_smem_bytes(p);
    auto kern = select_kernel<T, SEGMAX>;
    VQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)p.nq, kSelThreads, smem, st>>>(p);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

size_t select_vmajor_scratch_bytes(int64_t nq, int kprime, int64_t nseg) {
    // exact u64 | pair_vid i32 | plist i32 | pair_one i32 (per pair)   pair_m i32 | pair_tau f32 (per query)   vcount | vfill | voff (per video)
    return (size_t)nq * kprime * 20 + (size_t)nq * 8 + (size_t)(nseg + 1) * 12 + 64;
}

// SEGMAX with the video-major re-scoring; scratch: select_vmajor_scratch_bytes(), device memory
int launch_select_segmax_vmajor(SelectParams p, int64_t nseg, void *scratch, cudaStream_t st) {
    if (p.nq == 0) return VQ_OK;
    const int64_t npairs = p.nq * p.kprime;
    uint64_t *exact = (uint64_t *)scratch;
    int32_t *pair_vid = (int32_t *)(exact + npairs);
    int32_t *plist = pair_vid + npairs;
    int32_t *pair_one = plist + npairs;
    int32_t *pair_m = pair_one + npairs;
    float *pair_tau = (float *)(pair_m + p.nq);
    int32_t *vcount = (int32_t *)(pair_tau + p.nq);
    int32_t *vfill = vcount + (nseg + 1);
    int32_t *voff = vfill + (nseg + 1);
    VQ_CUDA(cudaMemsetAsync(vcount, 0, (size_t)(nseg + 1) * 8, st));  // vcount and vfill
    p.pair_vid = pair_vid; p.pair_m = pair_m; p.pair_tau = pair_tau; p.vcount = vcount;
    p.pair_one = p.seg_amb ? pair_one : nullptr;
    VQ_TRY(launch_select(p, st));  // phase A: the k' surviving videos of every query -> pairs
    if (p.seg_amb) {  // pairs with a proven-unique best moment: one canonical score each
        const int threads = (p.kprime + 31) / 32 * 32;  // <= kMaxKPrime = 512
        const size_t qsm = (size_t)((p.d + 3) / 4 * 4) * 4;
        VQ_CHECK(qsm <= 48 * 1024, VQ_ERR_CAPACITY, "segmax: dimension %d exceeds the staged query row", p.d);
        if (p.dtype == VQ_F32)
            segmax_single_kernel<float><<<(unsigned)p.nq, threads, qsm, st>>>(p, pair_one, exact);
        else
            segmax_single_kernel<__nv_bfloat16><<<(unsigned)p.nq, threads, qsm, st>>>(p, pair_one, exact);
    }
    scan_i32_kernel<<<1, 1024, 0, st>>>(vcount, voff, nseg);
    segmax_fill_kernel<<<(unsigned)((npairs + 255) / 256), 256, 0, st>>>(pair_vid, npairs, voff, vfill, plist);
    if (p.dtype == VQ_F32)
        segmax_rescore_kernel<float><<<(unsigned)nseg, VTHREADS, 0, st>>>(p, voff, plist, exact);
    else
        segmax_rescore_kernel<__nv_bfloat16><<<(unsigned)nseg, VTHREADS, 0, st>>>(p, voff, plist, exact);
    const size_t smem = (size_t)next_pow2(p.kprime) * 8;
    segmax_finish_kernel<<<(unsigned)p.nq, kSelThreads, smem, st>>>(p, exact);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

int launch_select(const SelectParams &p, cudaStream_t st) {
    if (p.nq == 0) return VQ_OK;
    VQ_CHECK((int64_t)p.segs_per_group * p.kprime <= (int64_t)kSelThreads * 64, VQ_ERR_CAPACITY,
             "select: %d segments x k'=%d exceeds the per-query gather capacity", p.segs_per_group, p.kprime);
    VQ_CHECK(select_smem_bytes(p) <= 200 * 1024, VQ_ERR_CAPACITY, "select: shared memory budget exceeded");
    const bool seg = p.seg_of_row != nullptr;
    if (p.dtype == VQ_F32) return seg ? launch_sel<float, true>(p, st) : launch_sel<float, false>(p, st);
    return seg ? launch_sel<__nv_bfloat16, true>(p, st) : launch_sel<__nv_bfloat16, false>(p, st);
}

}  // namespace vq
