// vq_losses.cu -- SURVEY.md 8-f3 / 8-f4: the other pairwise-distance computations of the training loop.
//   f3  memory-distance logits of the VA / VASA heads (models/resnet18_va.py:172-184, resnet18_vasa.py:198-206,
//       :217-229): logits[b,k] = -|normalize(x_b) - M_k|_2 / temperature, forward and backward
//   f4  MarginLoss (metric_learning_loss.py:462-569), NPairLoss (:384-457), ProxyNCALoss (:574-635) on the tuples
//       the host samplers drew ("tuples given", as vq_triplet_given)
// The reference runs these as Python loops over the batch / the tuples (one tiny launch per element).  Batches
// are a few hundred rows, so the kernels are latency-bound: each op is a handful of launches with deterministic
// reductions (sequential or fixed-shape trees, no float atomics).  fp32 arithmetic; parity bar 1e-6 absolute.
#include "vq_common.cuh"
#include "vq_kernels.h"

namespace vq {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_d2(const float *a, const float *b, int d, int lane) {
    float s = 0.f;
    for (int c = lane; c < d; c += 32) {
        const float df = a[c] - b[c];
        s = fmaf(df, df, s);
    }
    return warp_sum(s);
}
__device__ __forceinline__ float warp_dot(const float *a, const float *b, int d, int lane) {
    float s = 0.f;
    for (int c = lane; c < d; c += 32) s = fmaf(a[c], b[c], s);
    return warp_sum(s);
}

// ---- rows scaled to a fixed norm: out = scale * x / max(|x|, 1e-12) (F.normalize), nrm kept for the backward ----
__global__ void __launch_bounds__(256) row_normalize_kernel(const float *__restrict__ x, int64_t rows, int d, float scale,
                                                            float *__restrict__ out, float *__restrict__ nrm) {
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const float *xr = x + r * d;
    const float n = fmaxf(sqrtf(warp_dot(xr, xr, d, lane)), 1e-12f);
    const float k = scale / n;
    for (int c = lane; c < d; c += 32) out[r * d + c] = xr[c] * k;
    if (lane == 0) nrm[r] = n;
}
// dx = scale/|x| * (g - xh (xh . g)), xh = x/|x|
__global__ void __launch_bounds__(256) row_normalize_bwd_kernel(const float *__restrict__ x, const float *__restrict__ nrm,
                                                                const float *__restrict__ g, int64_t rows, int d,
                                                                float scale, float *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const float inv = 1.0f / nrm[r];
    const float dot = warp_dot(x + r * d, g + r * d, d, lane) * inv;  // xh . g
    for (int c = lane; c < d; c += 32) out[r * d + c] = scale * inv * (g[r * d + c] - x[r * d + c] * inv * dot);
}

// ---- all-pairs squared L2, direct differences: out[b,k] = sum (X[b]-M[k])^2; one block per b ----
__global__ void __launch_bounds__(256) pair_d2_kernel(const float *__restrict__ X, const float *__restrict__ M, int K, int d,
                                                      float *__restrict__ out) {
    extern __shared__ float xs[];
    const int b = blockIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int c = threadIdx.x; c < d; c += 256) xs[c] = X[(int64_t)b * d + c];
    __syncthreads();
    for (int k = w; k < K; k += 8) {
        const float s = warp_d2(xs, M + (int64_t)k * d, d, lane);
        if (lane == 0) out[(int64_t)b * K + k] = s;
    }
}
// GX[b] = sum_k W[b,k] (X[b]-M[k])  (side 0, one block per b)   GM[k] = -sum_b W[b,k] (X[b]-M[k])  (side 1, per k)
__global__ void __launch_bounds__(256) pair_grad_kernel(const float *__restrict__ X, const float *__restrict__ M,
                                                        const float *__restrict__ W, int B, int K, int d, int side,
                                                        float *__restrict__ out) {
    const int r = blockIdx.x;
    for (int c = threadIdx.x; c < d; c += 256) {
        float acc = 0.f;
        if (side == 0) {
            const float xv = X[(int64_t)r * d + c];
            for (int k = 0; k < K; ++k) acc = fmaf(W[(int64_t)r * K + k], xv - M[(int64_t)k * d + c], acc);
        } else {
            const float mv = M[(int64_t)r * d + c];
            for (int b = 0; b < B; ++b) acc = fmaf(-W[(int64_t)b * K + r], X[(int64_t)b * d + c] - mv, acc);
        }
        out[(int64_t)r * d + c] = acc;
    }
}

// ---- f3: memory logits ----
__global__ void __launch_bounds__(256) memlogit_fwd_kernel(const float *__restrict__ d2, int64_t n, float inv_t,
                                                           float *__restrict__ logits) {
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) logits[i] = -sqrtf(d2[i]) * inv_t;
}
__global__ void __launch_bounds__(256) memlogit_coef_kernel(const float *__restrict__ d2, const float *__restrict__ G, int64_t n,
                                                            float inv_t, float *__restrict__ W) {
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) {
        const float dist = sqrtf(d2[i]);
        W[i] = dist > 0.f ? -G[i] * inv_t / dist : 0.f;  // torch.norm's subgradient at 0 is 0
    }
}

int launch_memory_logits(const float *x, const float *mem, int B, int K, int d, float temperature, float *logits,
                         const float *G, float *gx, float *gmem, float *scratch, cudaStream_t st) {
    // scratch: xh[B*d] | nrm[B] | d2[B*K] | W[B*K] | gxh[B*d]
    float *xh = scratch, *nrm = xh + (size_t)B * d, *d2 = nrm + B, *W = d2 + (size_t)B * K, *gxh = W + (size_t)B * K;
    row_normalize_kernel<<<(B + 7) / 8, 256, 0, st>>>(x, B, d, 1.0f, xh, nrm);
    pair_d2_kernel<<<B, 256, (size_t)d * 4, st>>>(xh, mem, K, d, d2);
    const int64_t n = (int64_t)B * K;
    if (logits) memlogit_fwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d2, n, 1.0f / temperature, logits);
    if (G) {
        memlogit_coef_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d2, G, n, 1.0f / temperature, W);
        if (gx) {
            pair_grad_kernel<<<B, 256, 0, st>>>(xh, mem, W, B, K, d, 0, gxh);
            row_normalize_bwd_kernel<<<(B + 7) / 8, 256, 0, st>>>(x, nrm, gxh, B, d, 1.0f, gx);
        }
        if (gmem) pair_grad_kernel<<<K, 256, 0, st>>>(xh, mem, W, B, K, d, 1, gmem);
    }
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}
size_t memory_logits_scratch_floats(int B, int K, int d) { return (size_t)2 * B * d + B + (size_t)2 * B * K; }

// ---- f4: ProxyNCA ----
// one warp per batch row: loss_b = d2[b,y] + logsumexp_{c != y}(-d2[b,c]); W[b,c] = 2/B * (c == y ? 1 : -softmax_c)
__global__ void __launch_bounds__(256) proxynca_rows_kernel(const float *__restrict__ d2, const int32_t *__restrict__ y, int B,
                                                            int C, float *__restrict__ lossb, float *__restrict__ W) {
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (b >= B) return;
    const int yb = y[b];
    const float *row = d2 + (int64_t)b * C;
    float m = -INFINITY;
    for (int c = lane; c < C; c += 32)
        if (c != yb) m = fmaxf(m, -row[c]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float s = 0.f;
    for (int c = lane; c < C; c += 32)
        if (c != yb) s += expf(-row[c] - m);
    s = warp_sum(s);
    const float lse = m + logf(s);
    const float k = 2.0f / (float)B;
    for (int c = lane; c < C; c += 32) W[(int64_t)b * C + c] = c == yb ? k : -k * expf(-row[c] - lse);
    if (lane == 0) lossb[b] = row[yb] + lse;
}
__global__ void mean_kernel(const float *__restrict__ v, int n, float *__restrict__ out) {  // sequential, deterministic
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        float s = 0.f;
        for (int i = 0; i < n; ++i) s += v[i];
        *out = n > 0 ? s / (float)n : 0.f;
    }
}

int launch_proxynca(const float *emb, const float *proxies, const int32_t *labels, int B, int C, int d, float *loss,
                    float *gemb, float *gprox, float *scratch, cudaStream_t st) {
    // scratch: bn[B*d] | pn[C*d] | bnrm[B] | pnrm[C] | d2[B*C] | W[B*C] | lossb[B] | gbn[B*d] | gpn[C*d]
    float *bn = scratch, *pn = bn + (size_t)B * d, *bnrm = pn + (size_t)C * d, *pnrm = bnrm + B, *d2 = pnrm + C;
    float *W = d2 + (size_t)B * C, *lossb = W + (size_t)B * C, *gbn = lossb + B, *gpn = gbn + (size_t)B * d;
    row_normalize_kernel<<<(B + 7) / 8, 256, 0, st>>>(emb, B, d, 3.0f, bn, bnrm);
    row_normalize_kernel<<<(C + 7) / 8, 256, 0, st>>>(proxies, C, d, 3.0f, pn, pnrm);
    pair_d2_kernel<<<B, 256, (size_t)d * 4, st>>>(bn, pn, C, d, d2);
    proxynca_rows_kernel<<<(B + 7) / 8, 256, 0, st>>>(d2, labels, B, C, lossb, W);
    mean_kernel<<<1, 32, 0, st>>>(lossb, B, loss);
    if (gemb) {
        pair_grad_kernel<<<B, 256, 0, st>>>(bn, pn, W, B, C, d, 0, gbn);
        row_normalize_bwd_kernel<<<(B + 7) / 8, 256, 0, st>>>(emb, bnrm, gbn, B, d, 3.0f, gemb);
    }
    if (gprox) {
        pair_grad_kernel<<<C, 256, 0, st>>>(bn, pn, W, B, C, d, 1, gpn);
        row_normalize_bwd_kernel<<<(C + 7) / 8, 256, 0, st>>>(proxies, pnrm, gpn, C, d, 3.0f, gprox);
    }
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}
size_t proxynca_scratch_floats(int B, int C, int d) { return (size_t)2 * (B + C) * d + 2 * (size_t)B * C + 2 * B + 2 * C; }

// ---- f4: MarginLoss on given triplets (:520-565) ----
// one warp per triplet: d_ap, d_an = sqrt(sum (a-x)^2 + 1e-8)
__global__ void __launch_bounds__(256) margin_dist_kernel(const float *__restrict__ emb, int d, const int32_t *__restrict__ trip,
                                                          int64_t T, float *__restrict__ dap, float *__restrict__ dan) {
    const int lane = threadIdx.x & 31;
    const int64_t t = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (t >= T) return;
    const float *a = emb + (int64_t)trip[t * 3] * d;
    const float sp = warp_d2(a, emb + (int64_t)trip[t * 3 + 1] * d, d, lane);
    const float sn = warp_d2(a, emb + (int64_t)trip[t * 3 + 2] * d, d, lane);
    if (lane == 0) {
        dap[t] = sqrtf(sp + 1e-8f);
        dan[t] = sqrtf(sn + 1e-8f);
    }
}
// single block: hinge terms, pair count (triplets with an active term, :552-554), loss, per-triplet coefficients
__global__ void __launch_bounds__(256) margin_finish_kernel(const float *__restrict__ dap, const float *__restrict__ dan,
                                                            const float *__restrict__ beta_t, int beta_stride, int64_t T,
                                                            float margin, float *__restrict__ loss, float *__restrict__ kap,
                                                            float *__restrict__ kan, float *__restrict__ gbeta,
                                                            float *__restrict__ pair_count) {
    __shared__ float s_sum[256];
    __shared__ int s_cnt[256];
    float sum = 0.f;
    int cnt = 0;
    const int64_t per = (T + 255) / 256, b0 = threadIdx.x * per, e0 = min(T, b0 + per);
    for (int64_t t = b0; t < e0; ++t) {
        const float be = beta_t[t * beta_stride];
        const float pos = fmaxf(dap[t] - be + margin, 0.f), neg = fmaxf(be - dan[t] + margin, 0.f);
        sum += pos + neg;
        cnt += (pos > 0.f || neg > 0.f) ? 1 : 0;
    }
    s_sum[threadIdx.x] = sum;
    s_cnt[threadIdx.x] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
        int c = 0;
        for (int i = 0; i < 256; ++i) {
            s += s_sum[i];
            c += s_cnt[i];
        }
        s_sum[0] = s;
        s_cnt[0] = c;
        *loss = c == 0 ? s : s / (float)c;
        if (pair_count) *pair_count = (float)c;
    }
    __syncthreads();
    const float den = s_cnt[0] == 0 ? 1.f : (float)s_cnt[0];
    for (int64_t t = threadIdx.x; t < T; t += 256) {
        const float be = beta_t[t * beta_stride];
        const bool pa = dap[t] - be + margin > 0.f, na = be - dan[t] + margin > 0.f;
        kap[t] = pa ? 1.0f / (den * dap[t]) : 0.f;
        kan[t] = na ? 1.0f / (den * dan[t]) : 0.f;
        if (gbeta) gbeta[t] = ((na ? 1.f : 0.f) - (pa ? 1.f : 0.f)) / den;
    }
}
// one warp per gradient row, triplets scanned in order (deterministic gather)
__global__ void __launch_bounds__(256) margin_grad_kernel(const float *__restrict__ emb, int B, int d,
                                                          const int32_t *__restrict__ trip, int64_t T,
                                                          const float *__restrict__ kap, const float *__restrict__ kan,
                                                          float *__restrict__ grad) {
    const int lane = threadIdx.x & 31, r = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= B) return;
    for (int c0 = 0; c0 < d; c0 += 256) {
        float acc[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) acc[u] = 0.f;
        for (int64_t t = 0; t < T; ++t) {
            const int ia = trip[t * 3], ip = trip[t * 3 + 1], in_ = trip[t * 3 + 2];
            if (ia != r && ip != r && in_ != r) continue;
            const float ka = kap[t], kn = kan[t];
            if (ka == 0.f && kn == 0.f) continue;
            const float *a = emb + (int64_t)ia * d, *pp = emb + (int64_t)ip * d, *nn = emb + (int64_t)in_ * d;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int c = c0 + u * 32 + lane;
                if (c < d) {
                    const float dp = a[c] - pp[c], dn = a[c] - nn[c];
                    if (ia == r) acc[u] += ka * dp - kn * dn;
                    if (ip == r) acc[u] -= ka * dp;
                    if (in_ == r) acc[u] += kn * dn;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int c = c0 + u * 32 + lane;
            if (c < d) grad[(int64_t)r * d + c] = acc[u];
        }
    }
}

int launch_margin_given(const float *emb, int64_t B, int d, const int32_t *trip, int64_t T, const float *beta_t,
                        int beta_stride, float margin, float *loss, float *grad, float *gbeta, float *pair_count,
                        float *scratch, cudaStream_t st) {
    float *dap = scratch, *dan = dap + T, *kap = dan + T, *kan = kap + T;  // scratch: 4*T floats
    if (T > 0) margin_dist_kernel<<<(unsigned)((T + 7) / 8), 256, 0, st>>>(emb, d, trip, T, dap, dan);
    margin_finish_kernel<<<1, 256, 0, st>>>(dap, dan, beta_t, beta_stride, T, margin, loss, kap, kan, gbeta, pair_count);
    if (grad) margin_grad_kernel<<<(unsigned)((B + 7) / 8), 256, 0, st>>>(emb, (int)B, d, trip, T, kap, kan, grad);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

// ---- f4: NPairLoss on given tuples [P, 2+N] = (anchor, positive, negatives...) (:397-457) ----
// one block per tuple: s_j = a.(n_j - p), loss_i = log(1 + sum exp(s_j)), pi_j = exp(s_j) / (1 + sum), ws_i = |a|^2 + |p|^2
__global__ void __launch_bounds__(256) npair_tuple_kernel(const float *__restrict__ emb, int d, const int32_t *__restrict__ tup,
                                                          int N, float *__restrict__ pi, float *__restrict__ sig,
                                                          float *__restrict__ lossi, float *__restrict__ ws) {
    extern __shared__ float sh[];  // a[d] | s[N]
    float *as = sh, *s = sh + d;
    __shared__ float s_ap, s_tot;
    const int i = blockIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int32_t *t = tup + (int64_t)i * (2 + N);
    const float *a = emb + (int64_t)t[0] * d, *p = emb + (int64_t)t[1] * d;
    for (int c = threadIdx.x; c < d; c += 256) as[c] = a[c];
    __syncthreads();
    if (w == 0) {
        const float ap = warp_dot(as, p, d, lane), aa = warp_dot(as, as, d, lane), pp = warp_dot(p, p, d, lane);
        if (lane == 0) {
            s_ap = ap;
            ws[i] = aa + pp;
        }
    }
    __syncthreads();
    for (int j = w; j < N; j += 8) {
        const float an = warp_dot(as, emb + (int64_t)t[2 + j] * d, d, lane);
        if (lane == 0) s[j] = expf(an - s_ap);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float tot = 0.f;
        for (int j = 0; j < N; ++j) tot += s[j];
        s_tot = tot;
        lossi[i] = logf(1.0f + tot);
        sig[i] = tot / (1.0f + tot);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < N; j += 256) pi[(int64_t)i * N + j] = s[j] / (1.0f + s_tot);
}
__global__ void npair_finish_kernel(const float *__restrict__ lossi, const float *__restrict__ ws, int P, float l2,
                                    float *__restrict__ loss) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        float a = 0.f, b = 0.f;
        for (int i = 0; i < P; ++i) {
            a += lossi[i];
            b += ws[i];
        }
        *loss = P > 0 ? a / (float)P + l2 * (b / (float)P) : 0.f;
    }
}
// one block per gradient row; tuples scanned in order
__global__ void __launch_bounds__(256) npair_grad_kernel(const float *__restrict__ emb, int d, const int32_t *__restrict__ tup,
                                                         int P, int N, const float *__restrict__ pi,
                                                         const float *__restrict__ sig, float l2, float *__restrict__ grad) {
    const int r = blockIdx.x;
    const float invP = 1.0f / (float)P;
    for (int c = threadIdx.x; c < d; c += 256) {
        float acc = 0.f;
        for (int i = 0; i < P; ++i) {
            const int32_t *t = tup + (int64_t)i * (2 + N);
            const int ia = t[0], ip = t[1];
            const float av = emb[(int64_t)ia * d + c];
            if (ia == r) {
                const float pv = emb[(int64_t)ip * d + c];
                float g = 0.f;
                for (int j = 0; j < N; ++j) g = fmaf(pi[(int64_t)i * N + j], emb[(int64_t)t[2 + j] * d + c] - pv, g);
                acc += invP * g + 2.0f * l2 * invP * av;
            }
            if (ip == r) acc += -sig[i] * invP * av + 2.0f * l2 * invP * emb[(int64_t)ip * d + c];
            for (int j = 0; j < N; ++j)
                if (t[2 + j] == r) acc = fmaf(pi[(int64_t)i * N + j] * invP, av, acc);
        }
        grad[(int64_t)r * d + c] = acc;
    }
}

int launch_npair_given(const float *emb, int64_t B, int d, const int32_t *tup, int P, int N, float l2, float *loss,
                       float *grad, float *scratch, cudaStream_t st) {
    float *pi = scratch, *sig = pi + (size_t)P * N, *lossi = sig + P, *ws = lossi + P;  // scratch: P*N + 3*P floats
    if (P > 0) {
        const size_t smem = ((size_t)d + N) * 4;
        VQ_CUDA(cudaFuncSetAttribute(npair_tuple_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        npair_tuple_kernel<<<P, 256, smem, st>>>(emb, d, tup, N, pi, sig, lossi, ws);
    }
    npair_finish_kernel<<<1, 32, 0, st>>>(lossi, ws, P, l2, loss);
    if (grad) {
        if (P > 0) npair_grad_kernel<<<(unsigned)B, 256, 0, st>>>(emb, d, tup, P, N, pi, sig, l2, grad);
        else VQ_CUDA(cudaMemsetAsync(grad, 0, (size_t)B * d * 4, st));
    }
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

// ---- distance-weighted sampling weights (Sampler.inverse_sphere_distances :291-324), all anchors at once ----
// dist [B,B] = pdist clamped at the lower cutoff; q[i,:] = normalised inverse sphere density, 0 for the anchor's class
__global__ void __launch_bounds__(256) sphere_weights_kernel(const float *__restrict__ dist, const int32_t *__restrict__ labels,
                                                             int B, int dim, float *__restrict__ q) {
    __shared__ float red[256];
    const int i = blockIdx.x, la = labels[i];
    const float *row = dist + (int64_t)i * B;
    float m = -INFINITY;
    for (int j = threadIdx.x; j < B; j += 256) {
        const float dv = row[j];
        const float lq = labels[j] == la ? 0.f
                                         : (2.0f - (float)dim) * logf(dv) - ((float)(dim - 3) / 2) * logf(1.0f - 0.25f * dv * dv);
        m = fmaxf(m, lq);  // the reference takes the max over the row including the zeroed positives (:313-316)
    }
    red[threadIdx.x] = m;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] = fmaxf(red[threadIdx.x], red[threadIdx.x + o]);
        __syncthreads();
    }
    m = red[0];
    __syncthreads();
    float s = 0.f;
    for (int j = threadIdx.x; j < B; j += 256) {
        const float dv = row[j];
        float v = 0.f;
        if (labels[j] != la)
            v = expf((2.0f - (float)dim) * logf(dv) - ((float)(dim - 3) / 2) * logf(1.0f - 0.25f * dv * dv) - m);
        q[(int64_t)i * B + j] = v;
        s += v;
    }
    red[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    const float tot = red[0];
    for (int j = threadIdx.x; j < B; j += 256) q[(int64_t)i * B + j] /= tot;
}
int launch_sphere_weights(const float *dist, const int32_t *labels, int B, int dim, float *q, cudaStream_t st) {
    if (B > 0) sphere_weights_kernel<<<B, 256, 0, st>>>(dist, labels, B, dim, q);
    VQ_CUDA(cudaGetLastError());
    return VQ_OK;
}

}  // namespace vq
