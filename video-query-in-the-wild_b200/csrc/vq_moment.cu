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

template <int NT, int CAP>
__global__ void __launch_bounds__(256) moment_nms_kernel(const MomentNmsParams p, const int32_t *__restrict__ seg_list,
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
    for (int li = li0; li < li0 + per_group; ++li) {
        if (li >= nlist) break;  // uniform inside a group (NT == 256: one group per block)
        const int v = seg_list[li];
        const int64_t s0 = p.seg_start[v];
        const int S = (int)(p.seg_start[v + 1] - s0);
        if (S == 0) {
            if (t == 0) p.vbest[q * p.nseg + v] = ~0ull;
            continue;
        }
        int n2 = 1;
        while (n2 < S) n2 <<= 1;
        grp_sync<NT>();  // the previous segment's arrays are free
        for (int i = t; i < n2; i += NT) keys[i] = i < S ? make_key_asc(row[s0 + i], (uint32_t)i) : ~0ull;
        for (int k = 2; k <= n2; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1) {
                grp_sync<NT>();
                for (int i = t; i < n2; i += NT) {
                    const int ixj = i ^ j;
                    if (ixj > i) {
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
        for (int a = 0; a < S; ++a) {
            if (dead[a]) continue;
            const double xa1 = (double)x1[a], xa2 = (double)x2[a], la = xa2 - xa1 + 1.0;
            bool wrote = false;
            for (int b = a + 1 + t; b < S; b += NT) {
                if (dead[b]) continue;
                const double xb1 = (double)x1[b], xb2 = (double)x2[b], lb = xb2 - xb1 + 1.0;
                const double xx1 = xa1 > xb1 ? xa1 : xb1, xx2 = xa2 < xb2 ? xa2 : xb2;
                double inter = xx2 - xx1 + 1.0;
                inter = inter < 0.0 ? 0.0 : inter;
                const double ovr = inter / (la + lb - inter);
                if (!(ovr < p.thresh)) {
                    dead[b] = 1;
                    wrote = true;
                }
            }
            (void)wrote;
            grp_sync<NT>();
        }
        for (int i = t; i < S; i += NT) kept[s0 + (uint32_t)keys[i]] = dead[i] ? 0 : 1;
        if (t == 0) p.vbest[q * p.nseg + v] = (keys[0] & 0xffffffff00000000ull) | (uint64_t)(s0 + (uint32_t)keys[0]);
    }
}

int launch_moment_nms(const MomentNmsParams &p, const int32_t *small_list, int n_small, const int32_t *large_list,
                      int n_large, cudaStream_t st) {
    if (p.nq_tile == 0) return VQ_OK;
    if (n_small > 0) {
        constexpr int CAP = kMomentWarpCap, GB = CAP * 17, PER = 4;
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
