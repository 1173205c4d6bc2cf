// vq_sort.cu -- the full ascending list of faiss.IndexFlatL2.search(xq, k = ntotal), the only call the reference
// makes (dataloader_baseline.py:1592-1597, :935-940, :1268-1273): every gallery row of every query, ordered by
// (canonical d2 ascending, row ascending).
//
// Per query row of the canonical score tile (l2sq_all_kernel) a STABLE least-significant-digit radix sort on the
// order-preserving 32-bit image of d2 with the gallery row as payload: rows enter in ascending order, a stable sort
// keeps equal scores in that order -- the lowest-index tie-break comes for free and only 32 key bits are sorted.
// Four 8-bit passes; a pass is  histogram (per 4096-element block) -> exclusive scan over (digit, block) ->
// scatter with the stable in-block rank.  Pass 0 reads the float scores (and applies the ignore rule, :1548: ignored
// rows get the all-ones key and end up behind everything, reported as +inf / -1); the last pass writes D / I directly.
// Integer / byte work, HBM-bound: per element and pass 8 B read twice + 8 B written.
#include "vq_blocksel.cuh"
#include "vq_kernels.h"

namespace vq {

static constexpr int kRsThreads = 256;
static constexpr int kRsWarps = kRsThreads / 32;
static constexpr int kRsPerThread = 16;
static constexpr int kRsChunk = kRsThreads * kRsPerThread;  // elements per block
static constexpr uint32_t kRsNoDigit = 0x100u;              // lanes past the end of the row

// Lanes of the warp that hold the same digit (0 .. 256; 256 = "past the end of the row"), as a bit mask: nine ballots,
// one per digit bit.  __match_any_sync gives the same mask but its cost grows with the number of DISTINCT digits in the
// warp -- the passes over the low, random bytes of the scores ran 4x slower than the pass over the top byte (ncu: 1.61 /
// 1.22 / 0.84 / 0.43 ms for the four scatter passes over identical volumes).
__device__ __forceinline__ uint32_t rs_match(uint32_t dg) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 9; ++b) {
        const bool bit = (dg >> b) & 1u;
        const uint32_t vote = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? vote : ~vote;
    }
    return peers;
}

template <bool FIRST>
__device__ __forceinline__ uint64_t rs_load(const RowSortParams &p, int64_t q, int64_t i, const int *ign, int n_ign) {
    if (FIRST) {
        uint32_t key = f32_to_ordered(p.d2[q * p.n + i]);
        if (n_ign) {
            const int v = __ldg(p.col_video + i);
            bool m = false;
            for (int z = 0; z < n_ign; ++z) m |= (v == ign[z]);
            if (m) key = 0xffffffffu;
        }
        return ((uint64_t)key << 32) | (uint64_t)(uint32_t)i;
    }
    return p.in[q * p.n + i];
}

// hist[q][digit][block] = elements of block `block` of row q whose pass digit is `digit`
template <bool FIRST>
__global__ void __launch_bounds__(kRsThreads) rowsort_hist_kernel(const RowSortParams p, int shift, int nblk) {
    __shared__ int h[256];
    __shared__ int ign[kMaxIgnore];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t q = blockIdx.y;
    const int n_ign = (FIRST && p.q_ignore) ? p.n_ign : 0;
    h[tid] = 0;
    if (tid < kMaxIgnore) ign[tid] = tid < n_ign ? p.q_ignore[q * p.n_ign + tid] : -1;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kRsChunk + warp * (kRsChunk / kRsWarps);
#pragma unroll 4
    for (int r = 0; r < kRsPerThread; ++r) {
        const int64_t i = base + r * 32 + lane;
        uint32_t dg = kRsNoDigit;
        if (i < p.n) dg = (uint32_t)(rs_load<FIRST>(p, q, i, ign, n_ign) >> (32 + shift)) & 255u;
        const uint32_t m = rs_match(dg);
        if (dg != kRsNoDigit && lane == __ffs(m) - 1) atomicAdd(&h[dg], __popc(m));
    }
    __syncthreads();
    p.hist[(q * 256 + tid) * nblk + blockIdx.x] = h[tid];
}

// exclusive scan of hist[q] in (digit, block) order: the first output slot of every (digit, block) run
__global__ void __launch_bounds__(256) rowsort_scan_kernel(int32_t *__restrict__ hist, int nblk) {
    __shared__ int tot[256];
    const int tid = threadIdx.x;
    int32_t *h = hist + ((int64_t)blockIdx.x * 256 + tid) * nblk;
    int s = 0;
    for (int b = 0; b < nblk; ++b) s += h[b];
    tot[tid] = s;
    __syncthreads();
    if (tid < 32) {  // 256 totals: 8 per lane, warp scan
        int v[8], run = 0;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            v[e] = tot[tid * 8 + e];
            run += v[e];
        }
        int inc = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (tid >= o) inc += t;
        }
        int ex = inc - run;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            tot[tid * 8 + e] = ex;
            ex += v[e];
        }
    }
    __syncthreads();
    int run = tot[tid];
    for (int b = 0; b < nblk; ++b) {
        const int c = h[b];
        h[b] = run;
        run += c;
    }
}

template <bool FIRST, bool LAST>
__global__ void __launch_bounds__(kRsThreads, 3) rowsort_scatter_kernel(const RowSortParams p, int shift, int nblk) {
    __shared__ int cnt[kRsWarps][256];  // per-warp digit counts, then each warp's running output slot per digit
    __shared__ int ign[kMaxIgnore];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t q = blockIdx.y;
    const int n_ign = (FIRST && p.q_ignore) ? p.n_ign : 0;
#pragma unroll
    for (int w = 0; w < kRsWarps; ++w) cnt[w][tid] = 0;
    if (tid < kMaxIgnore) ign[tid] = tid < n_ign ? p.q_ignore[q * p.n_ign + tid] : -1;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kRsChunk + warp * (kRsChunk / kRsWarps);
    uint64_t e[kRsPerThread];
    int *mine = cnt[warp];
#pragma unroll
    for (int r = 0; r < kRsPerThread; ++r) {
        const int64_t i = base + r * 32 + lane;
        e[r] = i < p.n ? rs_load<FIRST>(p, q, i, ign, n_ign) : 0ull;
    }
    uint32_t mk[kRsPerThread];  // the peers of every element: formed once, used for the counts and for the ranks
#pragma unroll
    for (int r = 0; r < kRsPerThread; ++r) {
        const int64_t i = base + r * 32 + lane;
        const uint32_t dg = i < p.n ? (uint32_t)(e[r] >> (32 + shift)) & 255u : kRsNoDigit;
        const uint32_t m = rs_match(dg);
        mk[r] = m;
        if (dg != kRsNoDigit && lane == __ffs(m) - 1) mine[dg] += __popc(m);
        __syncwarp();
    }
    __syncthreads();
    {  // digit `tid`: first slot of this block's run, handed down the warps in element order
        int run = p.hist[(q * 256 + tid) * nblk + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kRsWarps; ++w) {
            const int c = cnt[w][tid];
            cnt[w][tid] = run;
            run += c;
        }
    }
    __syncthreads();
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kRsPerThread; ++r) {
        const int64_t i = base + r * 32 + lane;
        const uint32_t dg = i < p.n ? (uint32_t)(e[r] >> (32 + shift)) & 255u : kRsNoDigit;
        const uint32_t m = mk[r];
        if (dg != kRsNoDigit) {
            const int64_t slot = mine[dg] + __popc(m & lt);
            if (LAST) {
                if (slot < p.k) {
                    const uint32_t key = (uint32_t)(e[r] >> 32);
                    const bool ok = key != 0xffffffffu;
                    p.out_d2[q * p.k + slot] = ok ? ordered_to_f32(key) : INFINITY;
                    p.out_idx[q * p.k + slot] = ok ? (int64_t)(uint32_t)e[r] : -1;
                }
            } else {
                p.out[q * p.n + slot] = e[r];
            }
        }
        __syncwarp();
        if (dg != kRsNoDigit && lane == __ffs(m) - 1) mine[dg] += __popc(m);
        __syncwarp();
    }
}

size_t rowsort_hist_ints(int64_t nq_tile, int64_t n) {
    return (size_t)nq_tile * 256 * (size_t)((n + kRsChunk - 1) / kRsChunk);
}

int launch_row_sort(RowSortParams p, uint64_t *bufA, uint64_t *bufB, cudaStream_t st) {
    if (p.nq_tile == 0 || p.n == 0) return VQ_OK;
    VQ_CHECK(p.nq_tile <= 65535, VQ_ERR_INVALID, "row_sort: at most 65535 query rows per launch");
    const int nblk = (int)((p.n + kRsChunk - 1) / kRsChunk);
    const dim3 grid((unsigned)nblk, (unsigned)p.nq_tile);
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 8 * pass;
        p.in = (pass & 1) ? bufA : bufB;  // pass 0 reads d2
        p.out = (pass & 1) ? bufB : bufA;
        if (pass == 0)
            rowsort_hist_kernel<true><<<grid, kRsThreads, 0, st>>>(p, shift, nblk);
        else
            rowsort_hist_kernel<false><<<grid, kRsThreads, 0, st>>>(p, shift, nblk);
        rowsort_scan_kernel<<<(unsigned)p.nq_tile, 256, 0, st>>>(p.hist, nblk);
        if (pass == 0)
            rowsort_scatter_kernel<true, false><<<grid, kRsThreads, 0, st>>>(p, shift, nblk);
        else if (pass == 3)
            rowsort_scatter_kernel<false, true><<<grid, kRsThreads, 0, st>>>(p, shift, nblk);
        else
            rowsort_scatter_kernel<false, false><<<grid, kRsThreads, 0, st>>>(p, shift, nblk);
    }
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

}  // namespace vq
