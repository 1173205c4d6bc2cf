// vq_windows.cu -- SURVEY.md 8-f1: moment / clip window pooling, the step right before the hot path.
// The reference builds every gallery moment as np.mean(feat[:, lo:hi], axis=1) over the per-frame features of
// its untrimmed video (garner_feat, dataloader_baseline.py:1125-1147; clips :809-834): O(windows x length x C)
// on the host.  Here: one block per (video, channel tile) builds the per-channel prefix sums of the video in
// shared memory (float64) and every window of the video is two loads and one divide per channel; the moment
// matrix [n_windows, C] is written straight into HBM, where vq_index_add picks it up.
//
// HBM-bound: algorithmic bytes = 4*C*Ttot (frames read once) + 4*C*n_windows (moments written once).
#include "vq_common.cuh"
#include "vq_kernels.h"

namespace vq {

// CH channels per block; dynamic smem: double P[CH][pitch].  Warp w owns channels w, w+8, ...: the frames are
// staged coalesced, each lane sums its contiguous slice, a shuffle scan gives the slice offsets, and the slice is
// rewritten as running sums -- no block-wide serial step.  Windows: one thread per window, CH consecutive channels
// (one 32-byte sector per window and tile), one reciprocal per window.
static constexpr int kStageLoads = 8;  // (16 in flight measured the same: the staging loads are not what bounds the kernel)

template <int CH>
__global__ void __launch_bounds__(256) window_means_kernel(const float *__restrict__ feat, int C, int64_t Ttot,
                                                           const int64_t *__restrict__ vid_t0,
                                                           const int64_t *__restrict__ win_start,
                                                           const int32_t *__restrict__ lo, const int32_t *__restrict__ hi,
                                                           float *__restrict__ out, int pitch) {
    extern __shared__ __align__(16) uint8_t sm_raw[];
    double *P = (double *)sm_raw;
    // channel tiles of one video are neighbours in launch order: their 32-byte pieces of a window row reach L2 together
    const int ntile = (C + CH - 1) / CH;
    const int v = blockIdx.x / ntile, c0 = (blockIdx.x % ntile) * CH, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t t0 = vid_t0[v];
    const int T = (int)(vid_t0[v + 1] - t0);
    // stage: P[ch][t + 1] = frame value (coalesced along t)
    for (int ch = warp; ch < CH; ch += 8) {
        const int c = c0 + ch;
        double *row = P + ch * pitch;
        if (c < C) {
            // kStageLoads independent loads in flight per lane: a 600-frame channel is three memory latencies, not nineteen
            const float *src = feat + (int64_t)c * Ttot + t0;
            for (int t = lane; t < T; t += 32 * kStageLoads) {
                float v[kStageLoads];
#pragma unroll
                for (int u = 0; u < kStageLoads; ++u) v[u] = t + 32 * u < T ? __ldg(src + t + 32 * u) : 0.f;
#pragma unroll
                for (int u = 0; u < kStageLoads; ++u)
                    if (t + 32 * u < T) row[t + 32 * u + 1] = (double)v[u];
            }
        } else {
            for (int t = lane; t < T; t += 32) row[t + 1] = 0.0;
        }
        if (lane == 0) row[0] = 0.0;
        __syncwarp();
        const int per = (T + 31) / 32, b = min(T, lane * per), e = min(T, b + per);
        double tot = 0.0;
#pragma unroll 4
        for (int t = b; t < e; ++t) tot += row[t + 1];
        double inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        double run = inc - tot;  // sum of the slices before mine
#pragma unroll 4
        for (int t = b; t < e; ++t) {
            run += row[t + 1];
            row[t + 1] = run;
        }
    }
    __syncthreads();
    const int64_t w0 = win_start[v];
    const int nw = (int)(win_start[v + 1] - w0);
    for (int w = tid; w < nw; w += 256) {
        const int a = min(max(lo[w0 + w], 0), T), z = min(max(hi[w0 + w], a), T);  // windows are clipped to the video
        const double rl = z > a ? 1.0 / (double)(z - a) : 0.0;
        float r[CH];
#pragma unroll
        for (int ch = 0; ch < CH; ++ch) r[ch] = (float)((P[ch * pitch + z] - P[ch * pitch + a]) * rl);
        float *dst = out + (w0 + w) * (int64_t)C + c0;
        if (CH == 8 && c0 + CH <= C && (C & 7) == 0 && (((uintptr_t)out) & 31) == 0) {
            // one 256-bit store = the window's whole 32-byte sector of this channel tile (sm_100: STG.256).  As two
            // 16-byte stores every sector went through the L1 data pipe twice (ncu: 180 M store sectors for 2.84 GB
            // written, the pipe 90 % busy).
            asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst), "f"(r[0]), "f"(r[1]), "f"(r[2]),
                         "f"(r[3]), "f"(r[4 % CH]), "f"(r[5 % CH]), "f"(r[6 % CH]), "f"(r[7 % CH])
                         : "memory");
        } else if (CH % 4 == 0 && c0 + CH <= C && (C & 3) == 0 && (((uintptr_t)out) & 15) == 0) {
#pragma unroll
            for (int ch = 0; ch < CH; ch += 4) *reinterpret_cast<float4 *>(dst + ch) = make_float4(r[ch], r[ch + 1], r[ch + 2], r[ch + 3]);
        } else {
#pragma unroll
            for (int ch = 0; ch < CH; ++ch)
                if (c0 + ch < C) dst[ch] = r[ch];
        }
    }
}

int launch_window_means(const float *feat, int C, int64_t Ttot, const int64_t *vid_t0, const int64_t *win_start, int64_t V,
                        int64_t Tmax, const int32_t *lo, const int32_t *hi, float *out, cudaStream_t st) {
    if (V == 0) return VQ_OK;
    const int pitch = ((int)Tmax + 1) | 1;  // odd: the CH prefix rows start in different banks
    auto go = [&](auto kern, int CH) {
        const size_t smem = (size_t)CH * pitch * 8;
        VQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)(V * ((C + CH - 1) / CH)), 256, smem, st>>>(feat, C, Ttot, vid_t0, win_start, lo, hi, out,
                                                                               pitch);
        VQ_CUDA(cudaGetLastError());
        return VQ_OK;
    };
    const size_t budget = 200 * 1024;
    if ((size_t)8 * pitch * 8 <= budget) return go(window_means_kernel<8>, 8);
    if ((size_t)4 * pitch * 8 <= budget) return go(window_means_kernel<4>, 4);
    if ((size_t)2 * pitch * 8 <= budget) return go(window_means_kernel<2>, 2);
    VQ_CHECK((size_t)pitch * 8 <= budget, VQ_ERR_CAPACITY, "a video has %lld frames (max %d)", (long long)Tmax,
             (int)(budget / 8 - 1));
    return go(window_means_kernel<1>, 1);
}

}  // namespace vq
