// vq_prepare.cu -- index-build side kernels (a1, a2).
//   prepare_rows : embedding rows -> bf16 operand planes for the tensor-core pre-filter
//                  (hi = bf16(x), lo = bf16(x - hi), zero padded to a multiple of 64 columns),
//                  per-row |x|^2, and the augmentation block that carries -|x|^2/2 into the GEMM.
//                  Replaces the list->concat->index.add of dataloader_baseline.py:1502-1510.
//   pool_normalize: F.normalize(p=2,dim=1) per time step (main.py:232) + temporal mean
//                  (dataloader_baseline.py:1481), canonical arithmetic (see oracle).
#include "vq_common.cuh"
#include "vq_kernels.h"

namespace vq {

template <typename T>
__device__ __forceinline__ float ldf(const T *p);
template <>
__device__ __forceinline__ float ldf<float>(const float *p) { return *p; }
template <>
__device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16 *p) { return __bfloat162float(*p); }

// one warp per row at a time, kPrepRows rows per warp; 8 warps per 256-thread block
static constexpr int kPrepRows = 8;
template <typename T>
__global__ void __launch_bounds__(256) prepare_rows_kernel(const T *__restrict__ x, int64_t n, int d, int dp,
                                                           int aug_query, __nv_bfloat16 *__restrict__ hi,
                                                           __nv_bfloat16 *__restrict__ lo,
                                                           float *__restrict__ norm2, uint32_t *max_bits) {
    const int lane = threadIdx.x & 31;
    // kPrepRows rows per warp, one after the other: with one row per warp (8 KB of input per block) the kernel was bound
    // by the rate at which blocks can be launched -- 1.4 TB/s whatever the access width
    float wmax = 0.f;
    for (int rr = 0; rr < kPrepRows; ++rr) {
    const int64_t row = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * kPrepRows + rr;
    if (row >= n) break;
    const int pitch = dp + kAugCols;
    const T *xr = x + row * d;
    __nv_bfloat16 *hr = hi + row * pitch;
    __nv_bfloat16 *lr = lo ? lo + row * pitch : nullptr;
    float acc = 0.f;
    // Eight consecutive columns per lane and round where the row allows it (16-byte loads of bf16 rows, 32 bytes of
    // f32 rows; 16-byte stores of the planes): with two columns per lane the kernel moved 1.4 TB/s, a fifth of HBM.
    const bool wide = (d & 7) == 0 && (((uintptr_t)xr) & (sizeof(T) * 8 - 1)) == 0 && (pitch & 7) == 0 &&
                      (((uintptr_t)hi) & 15) == 0 && (!lo || (((uintptr_t)lo) & 15) == 0);
    for (int c = lane * 8; wide && c < dp; c += 256) {
        float v[8];
        if (c < d) {
            if (sizeof(T) == 4) {
                const float4 a = *reinterpret_cast<const float4 *>(xr + c), b = *reinterpret_cast<const float4 *>(xr + c + 4);
                v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
            } else {
                const uint4 w = *reinterpret_cast<const uint4 *>(xr + c);
                const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    v[2 * e] = __uint_as_float(ww[e] << 16);
                    v[2 * e + 1] = __uint_as_float(ww[e] & 0xffff0000u);
                }
            }
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = 0.f;
        }
        uint32_t hw[4], lw[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const __nv_bfloat16 h0 = __float2bfloat16_rn(v[2 * e]), h1 = __float2bfloat16_rn(v[2 * e + 1]);
            hw[e] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
            const __nv_bfloat16 l0 = __float2bfloat16_rn(v[2 * e] - __bfloat162float(h0));
            const __nv_bfloat16 l1 = __float2bfloat16_rn(v[2 * e + 1] - __bfloat162float(h1));
            lw[e] = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
            acc = fmaf(v[2 * e], v[2 * e], acc);
            acc = fmaf(v[2 * e + 1], v[2 * e + 1], acc);
        }
        *reinterpret_cast<uint4 *>(hr + c) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
        if (lr) *reinterpret_cast<uint4 *>(lr + c) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
    }
    for (int c = lane * 2; !wide && c < dp; c += 64) {
        float v0 = c < d ? ldf(xr + c) : 0.f;
        float v1 = c + 1 < d ? ldf(xr + c + 1) : 0.f;
        __nv_bfloat16 h0 = __float2bfloat16_rn(v0), h1 = __float2bfloat16_rn(v1);
        *reinterpret_cast<__nv_bfloat162 *>(hr + c) = __nv_bfloat162(h0, h1);
        if (lr) {
            __nv_bfloat16 l0 = __float2bfloat16_rn(v0 - __bfloat162float(h0));
            __nv_bfloat16 l1 = __float2bfloat16_rn(v1 - __bfloat162float(h1));
            *reinterpret_cast<__nv_bfloat162 *>(lr + c) = __nv_bfloat162(l0, l1);
        }
        acc = fmaf(v0, v0, acc);
        acc = fmaf(v1, v1, acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    // augmentation block (kernels.h): lanes 0..7 write two columns each
    if (lane < kAugCols / 2) {
        __nv_bfloat16 a0 = __float2bfloat16_rn(0.f), a1 = a0;
        if (lane < 2) {
            if (aug_query) {
                a0 = __float2bfloat16_rn(1.f);
                a1 = lane == 0 ? a0 : a1;
            } else {
                const float b = -0.5f * acc;  // exact scaling
                const __nv_bfloat16 bh = __float2bfloat16_rn(b);
                const float r1 = b - __bfloat162float(bh);  // exact (Sterbenz-like: bh is b's leading bits)
                const __nv_bfloat16 bm = __float2bfloat16_rn(r1);
                const __nv_bfloat16 bl = __float2bfloat16_rn(r1 - __bfloat162float(bm));
                a0 = lane == 0 ? bh : bl;
                a1 = lane == 0 ? bm : a1;
            }
        }
        *reinterpret_cast<__nv_bfloat162 *>(hr + dp + lane * 2) = __nv_bfloat162(a0, a1);
        if (lr) *reinterpret_cast<__nv_bfloat162 *>(lr + dp + lane * 2) = __nv_bfloat162(__float2bfloat16_rn(0.f), __float2bfloat16_rn(0.f));
    }
    if (lane == 0 && norm2) norm2[row] = acc;
    wmax = fmaxf(wmax, acc);
    }
    // one atomic per block: one per row on a single address is what bounded this kernel (4 M same-address atomics =
    // 5.9 ms for 4 M rows whatever the access width or the rows per block)
    if (max_bits) {
        __shared__ float s_max[8];
        if (lane == 0) s_max[threadIdx.x >> 5] = wmax;
        __syncthreads();
        if (threadIdx.x == 0) {
            float m = s_max[0];
#pragma unroll
            for (int w = 1; w < 8; ++w) m = fmaxf(m, s_max[w]);
            atomicMax(max_bits, __float_as_uint(m));  // norms are >= 0: bit order == value order
        }
    }
}

int launch_prepare_rows(const void *x, int dtype, int64_t n, int d, int dp, bool aug_query, __nv_bfloat16 *hi,
                        __nv_bfloat16 *lo, float *norm2, uint32_t *max_bits, cudaStream_t st) {
    if (n == 0) return VQ_OK;
    unsigned grid = (unsigned)((n + 8 * kPrepRows - 1) / (8 * kPrepRows));
    if (dtype == VQ_F32)
        prepare_rows_kernel<float><<<grid, 256, 0, st>>>((const float *)x, n, d, dp, aug_query ? 1 : 0, hi, lo, norm2,
                                                          max_bits);
    else
        prepare_rows_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16 *)x, n, d, dp, aug_query ? 1 : 0,
                                                                  hi, lo, norm2, max_bits);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

// block per batch item; x[b] is [C][T] with T contiguous
__global__ void __launch_bounds__(256) pool_normalize_kernel(const float *__restrict__ x, int C, int T,
                                                             float *__restrict__ out) {
    extern __shared__ float nr[];  // [T]
    const float *xb = x + (int64_t)blockIdx.x * C * T;
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        float acc = 0.f;
        for (int c = 0; c < C; ++c) {
            float v = xb[(int64_t)c * T + t];
            acc = __fmaf_rn(v, v, acc);
        }
        float nrm = __fsqrt_rn(acc);
        nr[t] = nrm > 1e-12f ? nrm : 1e-12f;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float acc = 0.f;
        for (int t = 0; t < T; ++t) acc = __fadd_rn(acc, __fdiv_rn(xb[(int64_t)c * T + t], nr[t]));
        out[(int64_t)blockIdx.x * C + c] = __fdiv_rn(acc, (float)T);
    }
}

int launch_pool_normalize(const float *x, int64_t B, int C, int T, float *out, cudaStream_t st) {
    if (B == 0) return VQ_OK;
    pool_normalize_kernel<<<(unsigned)B, 256, T * sizeof(float), st>>>(x, C, T, out);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

}  // namespace vq
