// vq_blocksel.cuh -- block-wide selection / sorting utilities on u64 keys held in shared memory.
// All functions must be called by every thread of the block (they contain __syncthreads()).
#pragma once
#include "vq_common.cuh"

namespace vq {

// block-wide sum of a per-thread int; result returned to all threads.  scratch: int[34] smem.
__device__ __forceinline__ int block_sum(int v, int *scratch) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = __reduce_add_sync(0xffffffffu, v);
    __syncthreads();  // scratch free
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    int t = (lane < nw) ? scratch[lane] : 0;
    t = __reduce_add_sync(0xffffffffu, t);
    return t;
}

// K-th smallest key (1-based K) among keys[0..n).  Requires 1 <= K <= n.  Bit-by-bit bisection:
// 64 rounds of a block-wide count.
__device__ __forceinline__ uint64_t block_kth_smallest(const uint64_t *keys, int n, int K, int *scratch) {
    uint64_t P = 0;
    for (int bit = 63; bit >= 0; --bit) {
        const uint64_t c = P | (1ull << bit);
        int cnt = 0;
        for (int i = threadIdx.x; i < n; i += blockDim.x) cnt += keys[i] < c ? 1 : 0;
        cnt = block_sum(cnt, scratch);
        if (cnt < K) P = c;  // fewer than K keys below c: the K-th smallest has this bit set
    }
    return P;
}

// K-th largest value of the HIGH 32-bit words (1-based), requires 1 <= K <= n.
__device__ __forceinline__ uint32_t block_kth_largest_hi(const uint64_t *keys, int n, int K, int *scratch) {
    uint32_t T = 0;
    for (int bit = 31; bit >= 0; --bit) {
        const uint32_t c = T | (1u << bit);
        int cnt = 0;
        for (int i = threadIdx.x; i < n; i += blockDim.x) cnt += (uint32_t)(keys[i] >> 32) >= c ? 1 : 0;
        cnt = block_sum(cnt, scratch);
        if (cnt >= K) T = c;
    }
    return T;
}

// In-place bitonic sort, ascending, of keys[0..n2) where n2 is a power of two.
__device__ __forceinline__ void block_bitonic_sort(uint64_t *keys, int n2) {
    for (int k = 2; k <= n2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            __syncthreads();
            for (int i = threadIdx.x; i < n2; i += blockDim.x) {
                int ixj = i ^ j;
                if (ixj > i) {
                    uint64_t a = keys[i], b = keys[ixj];
                    bool up = (i & k) == 0;
                    if ((a > b) == up) {
                        keys[i] = b;
                        keys[ixj] = a;
                    }
                }
            }
        }
    }
    __syncthreads();
}

// Packed fp32x2 canonical step (sm_100a FADD2 / FFMA2): two independent chains per instruction, each lane rounded
// exactly as the scalar __fsub_rn / __fmaf_rn pair (no FTZ) -- same bits as canon_step, half the issue slots.
__device__ __forceinline__ unsigned long long canon_step2(unsigned long long acc, unsigned long long q2, unsigned long long g2) {
    unsigned long long df, r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(df) : "l"(q2), "l"(g2));
    asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(r) : "l"(df), "l"(acc));
    return r;
}

__host__ __device__ __forceinline__ int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

// canonical d2 between a query held in shared memory (fp32) and one gallery row in global
// memory; T = float or __nv_bfloat16.  Sequential in i, bit-identical to the oracle.
template <typename T>
__device__ __forceinline__ float canon_d2_row(const float *qs, const T *g, int d);

template <>
__device__ __forceinline__ float canon_d2_row<float>(const float *qs, const float *g, int d) {
    float acc = 0.f;
    int i = 0;
    if ((((uintptr_t)g) & 15) == 0 && (((uintptr_t)qs) & 15) == 0) {
        // eight independent 16-byte loads in flight before the (sequential) chain consumes them: a row gathered at
        // random from HBM costs one memory latency per batch instead of one per load; the query is read as
        // LDS.128 (one shared-memory instruction per four chain steps: the MIO queue was the limiter)
        for (; i + 32 <= d; i += 32) {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldg(reinterpret_cast<const float4 *>(g + i) + u);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float4 qv = *reinterpret_cast<const float4 *>(qs + i + 4 * u);
                acc = canon_step(acc, qv.x, v[u].x);
                acc = canon_step(acc, qv.y, v[u].y);
                acc = canon_step(acc, qv.z, v[u].z);
                acc = canon_step(acc, qv.w, v[u].w);
            }
        }
        for (; i + 4 <= d; i += 4) {
            const float4 v = *reinterpret_cast<const float4 *>(g + i);
            acc = canon_step(acc, qs[i], v.x);
            acc = canon_step(acc, qs[i + 1], v.y);
            acc = canon_step(acc, qs[i + 2], v.z);
            acc = canon_step(acc, qs[i + 3], v.w);
        }
    }
    for (; i < d; ++i) acc = canon_step(acc, qs[i], g[i]);
    return acc;
}

template <>
__device__ __forceinline__ float canon_d2_row<__nv_bfloat16>(const float *qs, const __nv_bfloat16 *g, int d) {
    float acc = 0.f;
    int i = 0;
    if ((((uintptr_t)g) & 15) == 0 && (((uintptr_t)qs) & 15) == 0) {
        for (; i + 64 <= d; i += 64) {  // eight 16-byte loads in flight and LDS.128 query reads, as above
            uint4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldg(reinterpret_cast<const uint4 *>(g + i) + u);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float4 qa = *reinterpret_cast<const float4 *>(qs + i + 8 * u);
                const float4 qb = *reinterpret_cast<const float4 *>(qs + i + 8 * u + 4);
                acc = canon_step(acc, qa.x, __uint_as_float(v[u].x << 16));
                acc = canon_step(acc, qa.y, __uint_as_float(v[u].x & 0xffff0000u));
                acc = canon_step(acc, qa.z, __uint_as_float(v[u].y << 16));
                acc = canon_step(acc, qa.w, __uint_as_float(v[u].y & 0xffff0000u));
                acc = canon_step(acc, qb.x, __uint_as_float(v[u].z << 16));
                acc = canon_step(acc, qb.y, __uint_as_float(v[u].z & 0xffff0000u));
                acc = canon_step(acc, qb.z, __uint_as_float(v[u].w << 16));
                acc = canon_step(acc, qb.w, __uint_as_float(v[u].w & 0xffff0000u));
            }
        }
        for (; i + 8 <= d; i += 8) {
            const uint4 v = *reinterpret_cast<const uint4 *>(g + i);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                acc = canon_step(acc, qs[i + 2 * t], __uint_as_float(w[t] << 16));
                acc = canon_step(acc, qs[i + 2 * t + 1], __uint_as_float(w[t] & 0xffff0000u));
            }
        }
    }
    for (; i < d; ++i) acc = canon_step(acc, qs[i], __bfloat162float(g[i]));
    return acc;
}

// ---------------------------------------------------------------------------------------
// Warp-cooperative canonical re-scoring of 32 gathered rows (one per lane).
// canon_d2_row has every lane walk its own row: a warp-wide 16-byte load then touches 32 different 128-byte lines
// and the L1 tag stage, not HBM, sets the pace (ncu: lsu data-pipe wavefronts 56-81 % with 29 sectors per request).
// Here the warp copies the 32 rows 128 bytes at a time with cp.async -- eight lanes per row, so a request covers four
// full lines -- into a swizzled 4 KB shared-memory tile (two stages), and every lane then reads its own row's 128
// bytes back as eight conflict-free LDS.128.  The chain itself is unchanged (sequential in i, oracle bits).
// Needs d * sizeof(T) % 128 == 0 and 16-byte aligned rows; `tile` = this warp's 2 x 4096 bytes, 128-byte aligned.
// Lanes without a row pass any valid row pointer and ignore the result.
// ---------------------------------------------------------------------------------------
static constexpr int kRescoreTileBytes = 2 * 4096;  // per warp

__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <typename T>
__device__ __forceinline__ float canon_chunk128(float acc, const float *qs, const uint4 (&v)[8]);
template <>
__device__ __forceinline__ float canon_chunk128<float>(float acc, const float *qs, const uint4 (&v)[8]) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {  // 32 floats
        const float4 qv = *reinterpret_cast<const float4 *>(qs + 4 * u);
        acc = canon_step(acc, qv.x, __uint_as_float(v[u].x));
        acc = canon_step(acc, qv.y, __uint_as_float(v[u].y));
        acc = canon_step(acc, qv.z, __uint_as_float(v[u].z));
        acc = canon_step(acc, qv.w, __uint_as_float(v[u].w));
    }
    return acc;
}
template <>
__device__ __forceinline__ float canon_chunk128<__nv_bfloat16>(float acc, const float *qs, const uint4 (&v)[8]) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {  // 64 bf16
        const float4 qa = *reinterpret_cast<const float4 *>(qs + 8 * u);
        const float4 qb = *reinterpret_cast<const float4 *>(qs + 8 * u + 4);
        acc = canon_step(acc, qa.x, __uint_as_float(v[u].x << 16));
        acc = canon_step(acc, qa.y, __uint_as_float(v[u].x & 0xffff0000u));
        acc = canon_step(acc, qa.z, __uint_as_float(v[u].y << 16));
        acc = canon_step(acc, qa.w, __uint_as_float(v[u].y & 0xffff0000u));
        acc = canon_step(acc, qb.x, __uint_as_float(v[u].z << 16));
        acc = canon_step(acc, qb.y, __uint_as_float(v[u].z & 0xffff0000u));
        acc = canon_step(acc, qb.z, __uint_as_float(v[u].w << 16));
        acc = canon_step(acc, qb.w, __uint_as_float(v[u].w & 0xffff0000u));
    }
    return acc;
}

template <typename T>
__device__ __forceinline__ float warp_canon_d2_rows(const float *qs, const T *myrow, int d, uint8_t *tile, int lane) {
    constexpr int EPC = 128 / (int)sizeof(T);  // elements per 128-byte chunk
    const int nch = d / EPC;
    const uint32_t tb = smem_u32(tile);
    const int sub = lane >> 3, pc = lane & 7;
    // step j of a chunk copies piece pc of tile rows 4*j + sub
    const uint8_t *src[8];
    uint32_t dst[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int rt = 4 * j + sub;
        src[j] = (const uint8_t *)__shfl_sync(0xffffffffu, (unsigned long long)myrow, rt) + pc * 16;
        dst[j] = tb + rt * 128 + ((pc ^ (rt & 7)) << 4);
    }
    auto issue = [&](int ch) {
        const uint32_t so = (ch & 1) * 4096;
#pragma unroll
        for (int j = 0; j < 8; ++j) cp_async16(dst[j] + so, src[j] + (size_t)ch * 128);
        cp_async_commit();
    };
    __syncwarp();  // the tile's previous users are done
    issue(0);
    if (nch > 1) issue(1);
    float acc = 0.f;
    const uint8_t *mine = tile + lane * 128;
    for (int ch = 0; ch < nch; ++ch) {
        if (ch + 1 < nch) cp_async_wait<1>(); else cp_async_wait<0>();
        __syncwarp();
        uint4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = *reinterpret_cast<const uint4 *>(mine + (ch & 1) * 4096 + ((u ^ (lane & 7)) << 4));
        __syncwarp();  // everyone has read the stage: it may be refilled
        if (ch + 2 < nch) issue(ch + 2);
        acc = canon_chunk128<T>(acc, qs + ch * EPC, v);
    }
    return acc;
}

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

}  // namespace vq
