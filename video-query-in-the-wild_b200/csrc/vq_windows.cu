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

// CH channels per block; dynamic smem: double P[CH][Tmax + 1]
template <int CH>
__global__ void __launch_bounds__(256) window_means_kernel(const float *__restrict__ feat, int C, int64_t Ttot,
                                                           const int64_t *__restrict__ vid_t0,
                                                           const int64_t *__restrict__ win_start,
                                                           const int32_t *__restrict__ lo, const int32_t *__restrict__ hi,
                                                           float *__restrict__ out, int pitch) {
    extern __shared__ __align__(16) uint8_t sm_raw[];
    double *P = (double *)sm_raw;
    __shared__ double part[CH][257];
    const int v = blockIdx.x, c0 = blockIdx.y * CH, tid = threadIdx.x;
    const int64_t t0 = vid_t0[v];
    const int T = (int)(vid_t0[v + 1] - t0);
    const int per = (T + 255) / 256;
    const int b = min(T, tid * per), e = min(T, b + per);
    // local inclusive sums of this thread's slice, then the block-wide offsets
#pragma unroll
    for (int ch = 0; ch < CH; ++ch) {
        const int c = c0 + ch;
        double run = 0.0;
        if (c < C) {
            const float *src = feat + (int64_t)c * Ttot + t0;
            for (int t = b; t < e; ++t) {
                run += (double)src[t];
                P[ch * pitch + t + 1] = run;
            }
        }
        part[ch][tid] = run;
    }
    __syncthreads();
    if (tid < CH) {
        double run = 0.0;
        for (int i = 0; i < 256; ++i) {
            const double x = part[tid][i];
            part[tid][i] = run;
            run += x;
        }
        P[tid * pitch] = 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int ch = 0; ch < CH; ++ch) {
        const double off = part[ch][tid];
        for (int t = b; t < e; ++t) P[ch * pitch + t + 1] += off;
    }
    __syncthreads();
    // windows of this video: thread <-> (window, channel), CH consecutive channels per window (>= one 32-byte sector)
    const int64_t w0 = win_start[v];
    const int nw = (int)(win_start[v + 1] - w0);
    for (int i = tid; i < nw * CH; i += 256) {
        const int w = i / CH, ch = i % CH;
        const int c = c0 + ch;
        if (c >= C) continue;
        const int a = min(max(lo[w0 + w], 0), T), z = min(max(hi[w0 + w], a), T);  // windows are clipped to the video
        const double s = P[ch * pitch + z] - P[ch * pitch + a];
        out[(w0 + w) * (int64_t)C + c] = z > a ? (float)(s / (double)(z - a)) : 0.f;
    }
}

int launch_window_means(const float *feat, int C, int64_t Ttot, const int64_t *vid_t0, const int64_t *win_start, int64_t V,
                        int64_t Tmax, const int32_t *lo, const int32_t *hi, float *out, cudaStream_t st) {
    if (V == 0) return VQ_OK;
    const int pitch = ((int)Tmax + 1) | 1;  // odd: the CH prefix rows start in different banks
    auto go = [&](auto kern, int CH) {
        const size_t smem = (size_t)CH * pitch * 8;
        VQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<dim3((unsigned)V, (unsigned)((C + CH - 1) / CH)), 256, smem, st>>>(feat, C, Ttot, vid_t0, win_start, lo, hi, out,
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
