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
