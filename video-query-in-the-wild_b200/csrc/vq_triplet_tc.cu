// vq_triplet_tc.cu -- kernel (3), fused: batch-hard triplet loss forward + backward for groups of
// B <= 256 embeddings, ONE persistent CTA per group, three streaming passes over the group's rows
// (HBM once, then L2), everything else in shared memory / TMEM:
//
//   pass 1  TMA k-blocks [256 rows x 32 fp32] -> tcgen05.mma kind::tf32 Gram E E^T into TMEM
//           (two M=128 x N=256 accumulators = all 512 TMEM columns); the row threads compute the
//           canonical row norms from the same stages.  Only a CTA's first group runs this pass on its own: for
//           later groups its k-blocks are interleaved with pass 3 of the group before (same ring).
//   mine    thread = anchor reads its accumulator row (tcgen05.ld): coarse res = |a|^2+|j|^2-2G.
//           The TF32 Gram is only a PRE-FILTER with a proven error bound eps: every j whose coarse
//           value is within 2*eps of the coarse extreme stays a candidate (Sampler.pdist's clamps,
//           metric_learning_loss.py:276-289, widen the band when they can create ties).
//   pass 2  canonical blocked chains (oracle: orc_dot / orc_d2_blocked) for (anchor, candidate)
//           pairs, 8 lanes x LDS.128 per row block -> exact pdist of the candidates -> final
//           arg-max positive / arg-min negative (ties -> lowest index) and the loss terms
//           relu(|a-p|^2 - |a-n|^2 + m) (TripletLoss.triplet_distance, :342-355).
//   pass 3  backward as a deterministic gather: per gradient row the list of triplets that touch
//           it (in triplet order), acc = fmaf(2/T, x_u - x_v, acc); written once, coalesced.
//
// Results are bit-identical to the CUDA-core path (vq_triplet.cu) and to the oracle whatever the
// tensor cores rounded.  Groups that exceed a fixed capacity (more than 8 positive or 16 negative
// candidates of an anchor, more than 64 extra candidate pairs in a warp) are handed back through `bad_list` and recomputed
// by the CUDA-core path.
#include <stdlib.h>

#include <algorithm>

#include "vq_common.cuh"
#include "vq_kernels.h"

namespace vq {

static constexpr int TB = 256;                    // rows per group tile
static constexpr int TKB = 32;                    // fp32 columns per k-block (128-byte rows)
static constexpr int TSTAGE = TB * TKB * 4;       // 32 KB
static constexpr int TSTAGES = 5;
static constexpr int TROWT = 256;                 // row threads (warps 2..9)
static constexpr int TTHREADS = 64 + TROWT;
static constexpr int TCAND = 8;                   // positive candidates kept per anchor
// negative candidates kept per anchor: with 8, about 1 anchor in 400 000 of random unit rows (252 negatives,
// band 2 eps = 0.0084) overflowed, i.e. ~3 of 4096 groups went to the CUDA-core path, 0.3 ms after the kernel
#ifndef VQ_TCANDN
#define VQ_TCANDN 16  // 9 .. 16; -DVQ_TCANDN=8 rebuilds the former capacity for an A/B
#endif
static constexpr int TCANDN = VQ_TCANDN;
static_assert(TCANDN >= 8 && TCANDN <= 16, "two candidate words");
static constexpr uint32_t kTcSleepNs = 20000;   // suspend-time hint of the waits of the producer / MMA warp
static constexpr int TEXTRA = 64;                 // candidate pairs beyond the first of a role, per warp

struct TcLayout {
    static constexpr int OFF_NRM = TSTAGES * TSTAGE;       // float[256]
    static constexpr int OFF_LAB = OFF_NRM + 1024;         // int[256]
    static constexpr int OFF_TP = OFF_LAB + 1024;          // int[256] mined positive (-1: skipped anchor)
    static constexpr int OFF_TN = OFF_TP + 1024;           // int[256]
    static constexpr int OFF_LT = OFF_TN + 1024;           // float[256] per-anchor loss
    static constexpr int OFF_TL = OFF_LT + 1024;           // uint32[8][TEXTRA] extra candidate pairs (a | j<<8)
    static constexpr int OFF_RES = OFF_TL + 8 * TEXTRA * 4;  // float4[8][32] primaries (d2_p, prod_p, d2_n, prod_n)
    static constexpr int OFF_REX = OFF_RES + 8 * 32 * 16;    // float2[8][TEXTRA] extras (d2, prod)
    static constexpr int OFF_EOFF = OFF_REX + 8 * TEXTRA * 8;  // int[260] CSR offsets of the gradient gather lists
    static constexpr int OFF_ECNT = OFF_EOFF + 1040;        // int[256] counters
    static constexpr int OFF_ENT = OFF_ECNT + 1024;         // uint16[768] entries
    static constexpr int OFF_EN2 = OFF_ENT + 1536;          // uint2[1024] gather entries of pass 3
    static constexpr int OFF_NL = OFF_EN2 + 8192;           // float2[256] (|row|^2, label bits): one LDS.128 per two columns
    static constexpr int OFF_MISC = OFF_NL + 2048;          // int[16]
    static constexpr int OFF_BAR = OFF_MISC + 64;           // full[S] empty[S] tfull tempty
    static constexpr int OFF_TMEM = OFF_BAR + (2 * TSTAGES + 2) * 8;
    static constexpr int TOTAL = OFF_TMEM + 16 + 1024;
};
static_assert(TcLayout::TOTAL <= 227 * 1024, "fused triplet kernel shared memory");

struct TripletTcParams {
    const int32_t *labels;  // [G,B]
    int G, B, d;
    float margin;
    float *loss;     // [G]
    float *grad;     // [G,B,d] or null
    int32_t *trip;   // [G,B,3]
    int32_t *bad_count, *bad_list;
    long long *dbg;  // optional [grid][8] phase clocks of the first group (VQ_DEBUG_TRIPLET_CLOCKS)
};

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        :
        : "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// instruction descriptor: kind::tf32, A=B=tf32 (K-major), D=fp32
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ float tc_tree8_shfl(float t) {  // butterfly over the 8 lanes of a sub-group: c^4, c^2, c^1
    t = __fadd_rn(t, __shfl_xor_sync(0xffffffffu, t, 4));
    t = __fadd_rn(t, __shfl_xor_sync(0xffffffffu, t, 2));
    t = __fadd_rn(t, __shfl_xor_sync(0xffffffffu, t, 1));
    return t;
}
__device__ __forceinline__ float tc_tree8(const float (&t)[8]) {
    const float u0 = __fadd_rn(t[0], t[4]), u1 = __fadd_rn(t[1], t[5]), u2 = __fadd_rn(t[2], t[6]),
                u3 = __fadd_rn(t[3], t[7]);
    return __fadd_rn(__fadd_rn(u0, u2), __fadd_rn(u1, u3));
}
__device__ __forceinline__ float tc_pdist(float ni, float nj, float prod, float eps) {  // Sampler.pdist :286-289
    float res = __fsub_rn(__fadd_rn(ni, nj), __fmul_rn(2.0f, prod));
    res = res < 0.f ? 0.f : res;
    res = res < eps ? eps : res;
    return __fsqrt_rn(res);
}
// 16-byte chunk c of row r inside a 128B-swizzled [256 x 32 fp32] stage
__device__ __forceinline__ float4 tc_ld_chunk(const uint8_t *stage, int r, int c) {
    return *reinterpret_cast<const float4 *>(stage + r * 128 + ((c ^ (r & 7)) << 4));
}
__device__ __forceinline__ float tc_sq4(float4 a, float4 b) {  // chunk chain of (a-b)^2
    float df = __fsub_rn(a.x, b.x), t = __fmaf_rn(df, df, 0.f);
    df = __fsub_rn(a.y, b.y); t = __fmaf_rn(df, df, t);
    df = __fsub_rn(a.z, b.z); t = __fmaf_rn(df, df, t);
    df = __fsub_rn(a.w, b.w); t = __fmaf_rn(df, df, t);
    return t;
}
__device__ __forceinline__ unsigned long long tc_sub2(unsigned long long a, unsigned long long b) {
    unsigned long long r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ unsigned long long tc_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ float tc_dot4(float4 a, float4 b) {
    float t = __fmaf_rn(a.x, b.x, 0.f);
    t = __fmaf_rn(a.y, b.y, t);
    t = __fmaf_rn(a.z, b.z, t);
    t = __fmaf_rn(a.w, b.w, t);
    return t;
}

// ---------------------------------------------------------------------------------------------
// warp 0: TMA producer   warp 1: MMA issuer (pass 1) / stage releaser (passes 2, 3)
// warps 2..9: row threads (thread <-> row of the group; TMEM lane quarter = warp % 4)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TTHREADS, 1)
triplet_tc_kernel(const __grid_constant__ CUtensorMap tm_e, const TripletTcParams p) {
    using L = TcLayout;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    float *s_nrm = (float *)(smem + L::OFF_NRM);
    int *s_lab = (int *)(smem + L::OFF_LAB);
    int *s_tp = (int *)(smem + L::OFF_TP);
    int *s_tn = (int *)(smem + L::OFF_TN);
    float *s_lt = (float *)(smem + L::OFF_LT);
    uint32_t *s_tl = (uint32_t *)(smem + L::OFF_TL);
    float4 *s_res = (float4 *)(smem + L::OFF_RES);
    float2 *s_rex = (float2 *)(smem + L::OFF_REX);
    int *s_eoff = (int *)(smem + L::OFF_EOFF);
    int *s_ecnt = (int *)(smem + L::OFF_ECNT);
    uint16_t *s_ent = (uint16_t *)(smem + L::OFF_ENT);
    float2 *s_nl = (float2 *)(smem + L::OFF_NL);
    uint2 *s_en2 = (uint2 *)(smem + L::OFF_EN2);
    int *s_misc = (int *)(smem + L::OFF_MISC);  // [0] max |row|^2 bits  [1] T  [2] bad flag  [4..11] scan partials
    uint64_t *full_bar = (uint64_t *)(smem + L::OFF_BAR);
    uint64_t *empty_bar = full_bar + TSTAGES;
    uint64_t *tfull_bar = empty_bar + TSTAGES;
    uint64_t *tempty_bar = tfull_bar + 1;
    uint32_t *tmem_slot = (uint32_t *)(smem + L::OFF_TMEM);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KB = (p.d + TKB - 1) / TKB;
    const int B = p.B;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_e);
        for (int s = 0; s < TSTAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 9);  // 8 row warps + the MMA warp (tcgen05.commit in pass 1, plain arrive after)
        }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, 8);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================ TMA producer =================================
        // Stream order per group: [pass 1 of the FIRST group only]  pass 2 (kb = 0..KB-1)  then pass 3 INTERLEAVED with
        // pass 1 of this CTA's next group: P3(g, 0) P1(g', 0) P3(g, 1) P1(g', 1) ...  The next group's Gram (tensor
        // pipe, HBM stream) and row norms are ready when this group's gradient rows are written.
        int stage = 0;
        uint32_t phase = 0;
        auto load = [&](int g, int kb, uint64_t hint) {
            mbar_wait_sleepy(&empty_bar[stage], phase ^ 1u, kTcSleepNs);
            if (elect_one()) {
                mbar_expect_tx(&full_bar[stage], TSTAGE);
                tma_load_2d(&tm_e, &full_bar[stage], smem + stage * TSTAGE, kb * TKB, g * B, hint);
            }
            __syncwarp();
            if (++stage == TSTAGES) {
                stage = 0;
                phase ^= 1u;
            }
        };
        for (int kb = 0; kb < KB; ++kb) load(blockIdx.x, kb, kEvictLast);
        for (int g = blockIdx.x; g < p.G; g += gridDim.x) {
            const int gn = g + gridDim.x;
            for (int kb = 0; kb < KB; ++kb) load(g, kb, kEvictLast);  // the rows are read again in pass 3
            for (int kb = 0; kb < KB; ++kb) {
                load(g, kb, kEvictFirst);
                if (gn < p.G) load(gn, kb, kEvictLast);
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ===================================
        constexpr uint32_t idesc = umma_idesc_tf32(128, 256);
        int stage = 0;
        uint32_t phase = 0, tphase = 0;
        auto advance = [&]() {
            if (++stage == TSTAGES) {
                stage = 0;
                phase ^= 1u;
            }
        };
        auto gram_block = [&](int kb) {  // one k-block of E E^T
            mbar_wait_sleepy(&full_bar[stage], phase, kTcSleepNs);
            tc_fence_after();
            const uint32_t sa = smem_u32(smem + stage * TSTAGE);
            const uint64_t d_top = umma_desc_sw128(sa), d_bot = umma_desc_sw128(sa + 128 * 128);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < TKB / 8; ++k) {  // K = 8 tf32 = 32 bytes per step
                    umma_tf32(tmem_base, d_top + 2 * k, d_top + 2 * k, idesc, (kb | k) != 0);        // rows 0..127 x all
                    umma_tf32(tmem_base + 256, d_bot + 2 * k, d_top + 2 * k, idesc, (kb | k) != 0);  // rows 128..255 x all
                }
                umma_commit(&empty_bar[stage]);
                if (kb == KB - 1) umma_commit(tfull_bar);
            }
            __syncwarp();
            advance();
        };
        auto pass_by = [&]() {  // a stage this warp does not read
            mbar_wait_sleepy(&full_bar[stage], phase, kTcSleepNs);
            if (lane == 0) mbar_arrive(&empty_bar[stage]);
            __syncwarp();
            advance();
        };
        auto gram_begin = [&]() {  // the row threads have finished reading the previous Gram
            mbar_wait_sleepy(tempty_bar, tphase ^ 1u, kTcSleepNs);
            tphase ^= 1u;
            tc_fence_after();
        };
        gram_begin();
        for (int kb = 0; kb < KB; ++kb) gram_block(kb);
        for (int g = blockIdx.x; g < p.G; g += gridDim.x) {
            const bool has_next = g + (int)gridDim.x < p.G;
            for (int kb = 0; kb < KB; ++kb) pass_by();  // pass 2
            if (has_next) gram_begin();
            for (int kb = 0; kb < KB; ++kb) {            // pass 3, interleaved with the next group's pass 1
                pass_by();
                if (has_next) gram_block(kb);
            }
        }
    } else {
        // ================================ row threads ==================================
        const int q = warp & 3, half = (warp - 2) >> 2, rw = warp - 2;
        const int row0 = half * 128 + q * 32, row = row0 + lane;
        const int sw = row & 7;
        const int m = lane >> 3, c = lane & 7;  // sub-group of 8 lanes, 16-byte chunk of a 128-byte row block
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * 256);
        int stage = 0;
        uint32_t phase = 0, tphase = 0;
        auto next_stage = [&]() {
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[stage]);
            if (++stage == TSTAGES) {
                stage = 0;
                phase ^= 1u;
            }
        };
        // canonical |row|^2 partial of one k-block, from the stage the MMA reads
        auto norm_block = [&](float acc) {
            mbar_wait_sleepy(&full_bar[stage], phase, kTcSleepNs);
            const uint8_t *sp = smem + stage * TSTAGE + row * 128;
            float t[8];
#pragma unroll
            for (int cc = 0; cc < 8; ++cc) {
                const float4 x = *reinterpret_cast<const float4 *>(sp + ((cc ^ sw) << 4));
                t[cc] = tc_dot4(x, x);
            }
            acc = __fadd_rn(acc, tc_tree8(t));
            next_stage();
            return acc;
        };
        // ---------------- pass 1 of the first group (later groups: interleaved with the previous pass 3) ----------------
        float na = 0.f;
        for (int kb = 0; kb < KB; ++kb) na = norm_block(na);
        for (int g = blockIdx.x; g < p.G; g += gridDim.x) {
            const bool rvalid = row < B;
            const bool has_next = g + (int)gridDim.x < p.G;
            long long tk[8];
            tk[0] = clock64();
            tk[1] = clock64();
            s_nrm[row] = na;
            s_lab[row] = rvalid ? p.labels[(int64_t)g * B + row] : 0;
            // rows past B carry a NaN norm: their coarse value is NaN and drops out of every extreme / band test of the mine
            s_nl[row] = make_float2(rvalid ? na : __int_as_float(0x7fc00000), __int_as_float(s_lab[row]));
            s_ecnt[row] = 0;
            if (threadIdx.x == 64) {
                s_misc[0] = 0;
                s_misc[1] = 0;
                s_misc[2] = 0;
            }
            named_bar_sync(1, TROWT);
            if (rvalid) atomicMax(&s_misc[0], __float_as_int(na));  // na >= 0: int order == float order
            named_bar_sync(1, TROWT);
            const float nmax = __int_as_float(s_misc[0]);
            const int la = s_lab[row];
            // |coarse - canonical res| <= eps: TF32 operands (2^-10 each, truncation assumed), fp32
            // accumulation inside the tensor core, the canonical chain's own rounding
            const float eps = 4.2e-3f * sqrtf(na * nmax) + 1e-5f * (na + nmax);

            // ---------------- mine: coarse extremes, then candidates inside the band ----------------
            mbar_wait_sleepy(tfull_bar, tphase, kTcSleepNs);
            tphase ^= 1u;
            tc_fence_after();
            tk[2] = clock64();
            float maxp = -INFINITY, minn = INFINITY;
            // the anchor's own column becomes NaN too (the chunk that holds it is warp-uniform: row >> 5 == rw)
            auto blank_diag = [&](uint32_t (&r)[32], int ch) {
                if (ch == (row >> 5)) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) r[i] = i == (row & 31) ? 0x7fc00000u : r[i];
                }
            };
#pragma unroll 1
            for (int ch = 0; ch < TB / 32; ++ch) {
                if (ch * 32 >= B) break;
                uint32_t r[32];
                tmem_ld32(taddr + ch * 32, r);
                tmem_wait_ld();
                blank_diag(r, ch);
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    const float4 nl = *reinterpret_cast<const float4 *>(&s_nl[ch * 32 + i]);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const float cv = __fmaf_rn(-2.0f, __uint_as_float(r[i + h]), __fadd_rn(na, h ? nl.z : nl.x));
                        const bool same = __float_as_int(h ? nl.w : nl.y) == la;
                        maxp = same ? fmaxf(maxp, cv) : maxp;   // fmaxf / fminf drop a NaN operand
                        minn = same ? minn : fminf(minn, cv);
                    }
                }
            }
            const bool avalid = rvalid && maxp > -INFINITY && minn < INFINITY;
            const float qnan = __int_as_float(0x7fc00000);
            const float thr_p = !avalid ? qnan : (maxp < 1e-4f + 2.0f * eps ? -INFINITY : maxp - 2.0f * eps);  // clamp region: every positive ties
            const float thr_n = !avalid ? qnan : fmaxf(minn, 1e-4f) + 2.0f * eps;
            // candidate rows, one byte each, ascending row order: up to TCAND positives, up to TCANDN negatives
            uint64_t cpw = 0, cnw = 0, cnw2 = 0;
            int ncp = 0, ncn = 0;
            if (__any_sync(0xffffffffu, avalid)) {
#pragma unroll 1
                for (int ch = 0; ch < TB / 32; ++ch) {
                    if (ch * 32 >= B) break;
                    uint32_t r[32];
                    tmem_ld32(taddr + ch * 32, r);
                    tmem_wait_ld();
                    blank_diag(r, ch);
                    uint32_t mp = 0, mn = 0;  // band members of this chunk; the (rare) set bits are unpacked below
#pragma unroll
                    for (int i = 0; i < 32; i += 2) {
                        const float4 nl = *reinterpret_cast<const float4 *>(&s_nl[ch * 32 + i]);
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const float cv = __fmaf_rn(-2.0f, __uint_as_float(r[i + h]), __fadd_rn(na, h ? nl.z : nl.x));
                            const bool same = __float_as_int(h ? nl.w : nl.y) == la;
                            mp |= (same && cv >= thr_p) ? (1u << (i + h)) : 0u;
                            mn |= (!same && cv <= thr_n) ? (1u << (i + h)) : 0u;
                        }
                    }
                    while (mp) {
                        const int j = ch * 32 + __ffs(mp) - 1;
                        mp &= mp - 1;
                        if (ncp < TCAND) cpw |= (uint64_t)j << (8 * ncp);
                        ++ncp;
                    }
                    while (mn) {
                        const int j = ch * 32 + __ffs(mn) - 1;
                        mn &= mn - 1;
                        if (ncn < 8) cnw |= (uint64_t)j << (8 * ncn);
                        else if (ncn < TCANDN) cnw2 |= (uint64_t)j << (8 * (ncn - 8));
                        ++ncn;
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar);  // TMEM may take the next group's Gram
            tk[3] = clock64();
            bool bad = ncp > TCAND || ncn > TCANDN;
            ncp = min(ncp, TCAND);
            ncn = min(ncn, TCANDN);
            auto neg_at = [&](int t) { return (uint32_t)(((t < 8 ? cnw : cnw2) >> (8 * (t & 7))) & 255u); };

            // ---------------- work items of pass 2 ----------------
            // primary triple of every anchor (row, first positive candidate, first negative candidate): kept in
            // a register and fetched by shuffle; further candidates go to the warp's list of extra pairs
            const uint32_t prim = avalid ? ((uint32_t)row | ((uint32_t)(cpw & 255u) << 8) | ((uint32_t)(cnw & 255u) << 16))
                                         : ((uint32_t)row * 0x010101u);
            int nex = avalid ? (ncp - 1) + (ncn - 1) : 0;
            int incl = nex;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            int nex_total = __shfl_sync(0xffffffffu, incl, 31);
            const int exoff = incl - nex;
            if (nex_total > TEXTRA) {  // warp-uniform
                bad = true;
                nex_total = 0;
                nex = 0;
            }
            if (__any_sync(0xffffffffu, bad) && lane == 0) s_misc[2] = 1;
            if (nex > 0) {  // extra pair = swizzled chunk-0 byte offsets of the anchor row and the candidate row
                auto base_of = [](uint32_t r) { return (r << 7) | ((r & 7u) << 4); };
                int w = exoff;
                for (int t = 1; t < ncp; ++t)
                    s_tl[rw * TEXTRA + w++] = base_of((uint32_t)row) | (base_of((uint32_t)((cpw >> (8 * t)) & 255u)) << 16);
                for (int t = 1; t < ncn; ++t)
                    s_tl[rw * TEXTRA + w++] = base_of((uint32_t)row) | (base_of(neg_at(t)) << 16);
            }
            const bool wvalid = __any_sync(0xffffffffu, avalid);
            __syncwarp();

            // ---------------- pass 2: canonical chains of the candidate pairs ----------------
            // A sub-group of 8 lanes (lane c = 16-byte chunk c of a 128-byte row block) works on the primaries of lanes
            // 4*i + m, four at a time ("half" h: i = 4h .. 4h+3): 16 partial sums per lane and half (4 primaries x 4
            // chains), each to be summed over the 8 lanes in the oracle's tree order (c^4, c^2, c^1).  A transposing
            // butterfly does that with 14 shuffles instead of 48: at every step a lane keeps the half of its values whose
            // index bit matches its own lane bit and hands the other half to its partner.  Lane c ends up with two sums:
            // primary 4h + (c >> 1), chains 2(c & 1) and 2(c & 1) + 1, which it alone accumulates over the k-blocks.
            // Extra pairs go 16 to a step (4 per sub-group, 2 chains each): 7 shuffles, one sum per lane.
            // The loads of half 0 of the NEXT k-block are issued before half 1 of this one is folded, so that a warp
            // always has twelve 16-byte loads in flight under its arithmetic (the k-block barrier no longer drains it).
            float accp[4] = {0.f, 0.f, 0.f, 0.f}, accx[TEXTRA / 16];
            uint32_t po[8], po2[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t e = __shfl_sync(0xffffffffu, prim, 4 * i + m);
                const uint32_t ra = e & 255u, r1 = (e >> 8) & 255u, r2 = (e >> 16) & 255u;
                auto off = [&](uint32_t r) { return (r << 7) | (((uint32_t)c ^ (r & 7u)) << 4); };
                po[i] = off(ra) | (off(r1) << 16);
                po2[i] = off(r2);
            }
#pragma unroll
            for (int i = 0; i < TEXTRA / 16; ++i) accx[i] = 0.f;
            const bool b4 = (c & 4) != 0, b2 = (c & 2) != 0, b1 = (c & 1) != 0;
            const uint32_t cx2 = (uint32_t)c << 4;
            const int nxs = (nex_total + 15) >> 4;  // steps of 16 extra pairs, warp-uniform
            auto ld16 = [](const uint8_t *q) { return *reinterpret_cast<const float4 *>(q); };
            auto load_half = [&](const uint8_t *sp, const int h, float4 (&xa)[4], float4 (&x1)[4], float4 (&x2)[4]) {
#pragma unroll
                for (int ii = 0; ii < 4; ++ii) {
                    xa[ii] = ld16(sp + (po[4 * h + ii] & 0xffffu));
                    x1[ii] = ld16(sp + (po[4 * h + ii] >> 16));
                    x2[ii] = ld16(sp + po2[4 * h + ii]);
                }
            };
            auto fold_half = [&](const float4 (&xa)[4], const float4 (&x1)[4], const float4 (&x2)[4], float &s0, float &s1) {
                float v[16], w[8], u[4];
#pragma unroll
                for (int ii = 0; ii < 4; ++ii) {
                    v[4 * ii] = tc_sq4(xa[ii], x1[ii]);
                    v[4 * ii + 1] = tc_dot4(xa[ii], x1[ii]);
                    v[4 * ii + 2] = tc_sq4(xa[ii], x2[ii]);
                    v[4 * ii + 3] = tc_dot4(xa[ii], x2[ii]);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    w[j] = __fadd_rn(b4 ? v[8 + j] : v[j], __shfl_xor_sync(0xffffffffu, b4 ? v[j] : v[8 + j], 4));
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    u[j] = __fadd_rn(b2 ? w[4 + j] : w[j], __shfl_xor_sync(0xffffffffu, b2 ? w[j] : w[4 + j], 2));
                s0 = __fadd_rn(s0, __fadd_rn(b1 ? u[2] : u[0], __shfl_xor_sync(0xffffffffu, b1 ? u[0] : u[2], 1)));
                s1 = __fadd_rn(s1, __fadd_rn(b1 ? u[3] : u[1], __shfl_xor_sync(0xffffffffu, b1 ? u[1] : u[3], 1)));
            };
            float4 xa[4], x1[4], x2[4];  // half 0 of the k-block at hand
            if (wvalid) {
                mbar_wait_sleepy(&full_bar[stage], phase, kTcSleepNs);
                load_half(smem + stage * TSTAGE, 0, xa, x1, x2);
            }
            for (int kb = 0; kb < KB; ++kb) {
                mbar_wait_sleepy(&full_bar[stage], phase, kTcSleepNs);
                const uint8_t *sp = smem + stage * TSTAGE;
                if (wvalid) {
                    float4 ya[4], y1[4], y2[4];
                    load_half(sp, 1, ya, y1, y2);
                    fold_half(xa, x1, x2, accp[0], accp[1]);
                    if (kb + 1 < KB) {  // half 0 of the next k-block (the stage after this one in the ring)
                        const int ns = stage + 1 == TSTAGES ? 0 : stage + 1;
                        mbar_wait_sleepy(&full_bar[ns], ns == 0 ? phase ^ 1u : phase, kTcSleepNs);
                        load_half(smem + ns * TSTAGE, 0, xa, x1, x2);
                    }
                    fold_half(ya, y1, y2, accp[2], accp[3]);
#pragma unroll
                    for (int s = 0; s < TEXTRA / 16; ++s) {
                        if (s < nxs) {  // warp-uniform
                            float x[8], w[4], u[2];
#pragma unroll
                            for (int ee = 0; ee < 4; ++ee) {
                                const int idx = 16 * s + 4 * ee + m;
                                const uint32_t e = idx < nex_total ? s_tl[rw * TEXTRA + idx] : 0u;
                                const float4 za = ld16(sp + ((e & 0xffffu) ^ cx2)), zj = ld16(sp + ((e >> 16) ^ cx2));
                                x[2 * ee] = tc_sq4(za, zj);
                                x[2 * ee + 1] = tc_dot4(za, zj);
                            }
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                w[j] = __fadd_rn(b4 ? x[4 + j] : x[j], __shfl_xor_sync(0xffffffffu, b4 ? x[j] : x[4 + j], 4));
#pragma unroll
                            for (int j = 0; j < 2; ++j)
                                u[j] = __fadd_rn(b2 ? w[2 + j] : w[j], __shfl_xor_sync(0xffffffffu, b2 ? w[j] : w[2 + j], 2));
                            accx[s] = __fadd_rn(accx[s], __fadd_rn(b1 ? u[1] : u[0], __shfl_xor_sync(0xffffffffu, b1 ? u[0] : u[1], 1)));
                        }
                    }
                }
                next_stage();
            }
            {   // primary i = 4h + (c >> 1) belongs to lane 4i + m: (d2_p, prod_p | d2_n, prod_n), this lane holds one half
                float *res = reinterpret_cast<float *>(s_res), *rex = reinterpret_cast<float *>(s_rex);
#pragma unroll
                for (int h = 0; h < 2; ++h)
                    *reinterpret_cast<float2 *>(res + (rw * 32 + 4 * (4 * h + (c >> 1)) + m) * 4 + 2 * (c & 1)) =
                        make_float2(accp[2 * h], accp[2 * h + 1]);
#pragma unroll
                for (int s = 0; s < TEXTRA / 16; ++s) {  // extra 16s + 4(c >> 1) + m: (d2, prod), one of the two per lane
                    const int idx = 16 * s + 4 * (c >> 1) + m;
                    if (idx < nex_total) rex[(rw * TEXTRA + idx) * 2 + (c & 1)] = accx[s];
                }
            }
            __syncwarp();
            tk[4] = clock64();

            // ---------------- exact decision among the candidates, per-anchor loss ----------------
            int fp = -1, fn = -1;
            float lterm = 0.f;
            if (avalid) {  // (a group with a capacity overflow is recomputed by the CUDA-core path anyway)
                const float4 pr = s_res[rw * 32 + lane];
                const float2 *rx = s_rex + rw * TEXTRA + exoff;
                // candidate t of a role: t == 0 the primary, t >= 1 the extras in list order
                int bq = 0;
                float d2p = pr.x;
                if (ncp > 1 && nex > 0) {
                    float bd = tc_pdist(na, s_nrm[cpw & 255u], pr.y, 1e-4f);
                    for (int t = 1; t < ncp; ++t) {
                        const float dist = tc_pdist(na, s_nrm[(cpw >> (8 * t)) & 255u], rx[t - 1].y, 1e-4f);
                        if (dist > bd) {  // ascending row order: the first maximum wins
                            bd = dist;
                            bq = t;
                            d2p = rx[t - 1].x;
                        }
                    }
                }
                fp = (int)((cpw >> (8 * bq)) & 255u);
                bq = 0;
                float d2n = pr.z;
                if (ncn > 1 && nex > 0) {
                    const float2 *rn = rx + (ncp - 1);
                    float bd = tc_pdist(na, s_nrm[cnw & 255u], pr.w, 1e-4f);
                    for (int t = 1; t < ncn; ++t) {
                        const float dist = tc_pdist(na, s_nrm[neg_at(t)], rn[t - 1].y, 1e-4f);
                        if (dist < bd) {
                            bd = dist;
                            bq = t;
                            d2n = rn[t - 1].x;
                        }
                    }
                }
                fn = (int)neg_at(bq);
                const float v = __fadd_rn(__fsub_rn(d2p, d2n), p.margin);
                lterm = v > 0.f ? v : 0.f;
            }
            s_tp[row] = fp;
            s_tn[row] = fn;
            s_lt[row] = lterm;
            if (rvalid) {
                int32_t *tr = p.trip + ((int64_t)g * B + row) * 3;
                tr[0] = fp >= 0 ? row : -1;
                tr[1] = fp;
                tr[2] = fn;
            }
            named_bar_sync(1, TROWT);

            // ---------------- loss mean (:379) and the gradient gather lists ----------------
            const bool act = fp >= 0 && lterm > 0.f;
            if (act) {
                atomicAdd(&s_ecnt[row], 1);
                atomicAdd(&s_ecnt[fp], 1);
                atomicAdd(&s_ecnt[fn], 1);
            }
            {
                const unsigned vb = __ballot_sync(0xffffffffu, fp >= 0);
                if (lane == 0 && vb) atomicAdd(&s_misc[1], __popc(vb));
            }
            if (threadIdx.x == 64) {  // sequential sum in triplet order, as the oracle (skipped anchors hold +0)
                float sum = 0.f;
                for (int t = 0; t < TB; t += 4) {
                    const float4 v = *reinterpret_cast<const float4 *>(&s_lt[t]);
                    sum = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(sum, v.x), v.y), v.z), v.w);
                }
                s_misc[3] = __float_as_int(sum);
            }
            named_bar_sync(1, TROWT);
            const int T = s_misc[1];
            if (threadIdx.x == 64) p.loss[g] = T > 0 ? __fdiv_rn(__int_as_float(s_misc[3]), (float)T) : 0.f;
            const int cnt = s_ecnt[row];
            {   // exclusive scan of the per-row list lengths in row order; a row without triplets gets one
                // no-op entry (x_r - x_r) so that every gradient row is written by the same loop
                const int len = cnt > 0 ? cnt : 1;
                int inc2 = len;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, inc2, o);
                    if (lane >= o) inc2 += v;
                }
                if (lane == 31) s_misc[4 + (row >> 5)] = inc2;
                named_bar_sync(1, TROWT);
                int basep = 0;
                for (int b = 0; b < (row >> 5); ++b) basep += s_misc[4 + b];
                s_eoff[row] = basep + inc2 - len;
                if (row == TB - 1) s_eoff[TB] = basep + inc2;
                s_ecnt[row] = 0;
            }
            named_bar_sync(1, TROWT);
            if (act) {  // staging entry = (triplet << 2 | role); sorted by triplet below
                s_ent[s_eoff[row] + atomicAdd(&s_ecnt[row], 1)] = (uint16_t)((row << 2) | 0);
                s_ent[s_eoff[fp] + atomicAdd(&s_ecnt[fp], 1)] = (uint16_t)((row << 2) | 1);
                s_ent[s_eoff[fn] + atomicAdd(&s_ecnt[fn], 1)] = (uint16_t)((row << 2) | 2);
            }
            named_bar_sync(1, TROWT);
            {   // each row: its triplets in order -> entries (byte offsets of rows u and v inside a stage, chunk 0
                // after the swizzle; owner row; last-of-row flag): contribution x_u - x_v
                const int e0 = s_eoff[row];
                auto base_of = [](int r) { return (uint32_t)((r << 7) | ((r & 7) << 4)); };
                if (cnt == 0) {
                    s_en2[e0] = make_uint2(base_of(row) | (base_of(row) << 16), (uint32_t)row | 256u);
                } else {
                    for (int i = e0 + 1; i < e0 + cnt; ++i) {
                        const uint16_t key = s_ent[i];
                        int j = i - 1;
                        while (j >= e0 && s_ent[j] > key) {
                            s_ent[j + 1] = s_ent[j];
                            --j;
                        }
                        s_ent[j + 1] = key;
                    }
                    for (int i = e0; i < e0 + cnt; ++i) {
                        const int t = s_ent[i] >> 2, role = s_ent[i] & 3;
                        int u, v;
                        if (role == 0) { u = s_tn[t]; v = s_tp[t]; }      // anchor:   2(n - p)/T
                        else if (role == 1) { u = row; v = t; }           // positive: 2(p - a)/T
                        else { u = t; v = row; }                          // negative: 2(a - n)/T
                        s_en2[i] = make_uint2(base_of(u) | (base_of(v) << 16),
                                              (uint32_t)row | (i == e0 + cnt - 1 ? 256u : 0u));
                    }
                }
            }
            named_bar_sync(1, TROWT);
            const float scale = T > 0 ? __fdiv_rn(2.0f, (float)T) : 0.f;
            tk[5] = clock64();

            // ---------------- pass 3: gradient rows, 8 lanes x 16 bytes per row block ----------------
            // The entry array is cut at row boundaries into 32 near-equal pieces, one per 8-lane sub-group; a
            // sub-group walks its piece with the chain acc = fmaf(2/T, x_u - x_v, acc) (triplet order inside a row),
            // loads running two entries ahead of the chain, and writes a row when its last entry retires.
            const int E = s_eoff[TB];
            auto piece = [&](int sgi) {
                if (sgi >= TROWT / 8) return E;
                const int x = (sgi * E) / (TROWT / 8);
                if (x == 0) return 0;
                const int r = s_en2[x].y & 255u;
                return s_eoff[r] == x ? x : s_eoff[r + 1];
            };
            const int eb = piece(rw * 4 + m), ee = piece(rw * 4 + m + 1);
            const int nit = __reduce_max_sync(0xffffffffu, ee - eb);
            const uint32_t cx = (uint32_t)c << 4;
            unsigned long long scale2;
            asm("mov.b64 %0, {%1, %1};" : "=l"(scale2) : "r"(__float_as_uint(scale)));
            float na_next = 0.f;
            for (int kb = 0; kb < KB; ++kb) {
                mbar_wait_sleepy(&full_bar[stage], phase, kTcSleepNs);
                if (p.grad) {
                    const uint8_t *sp = smem + stage * TSTAGE;
                    const int col = kb * TKB + c * 4;
                    auto ldx = [&](uint32_t off) { return *reinterpret_cast<const ulonglong2 *>(sp + ((off & 0x7ff0u) ^ cx)); };
                    // E <= 768 of the 1024 slots: reading three entries past a piece's end stays inside shared memory, and
                    // whatever is there only feeds loads inside the stage (offsets are masked) and a chain nobody stores
                    float *gb = p.grad + ((int64_t)g * B) * p.d + col;
                    const uint32_t rlim = col < p.d ? (uint32_t)B : 0u;
                    const uint2 *ep = s_en2 + eb;
                    uint2 en0 = ep[0], en1 = ep[1], en2 = ep[2];
                    ulonglong2 xu0 = ldx(en0.x), xv0 = ldx(en0.x >> 16), xu1 = ldx(en1.x), xv1 = ldx(en1.x >> 16);
                    unsigned long long alo = 0ull, ahi = 0ull;  // packed fp32 pairs: FADD2 / FFMA2, same bits as the scalar chain
#pragma unroll 3
                    for (int it = 0; it < nit; ++it) {
                        const ulonglong2 xu2 = ldx(en2.x), xv2 = ldx(en2.x >> 16);  // loads run two entries ahead of the chain
                        const uint2 en3 = ep[it + 3];
                        alo = tc_fma2(scale2, tc_sub2(xu0.x, xv0.x), alo);
                        ahi = tc_fma2(scale2, tc_sub2(xu0.y, xv0.y), ahi);
                        // straight-line tail: the four sub-groups of a warp finish their rows at different entries
                        const uint32_t r = en0.y & 255u;
                        const bool last = (en0.y & 256u) != 0u;
                        ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(gb + (size_t)(r * (uint32_t)p.d));
                        if (last && eb + it < ee && r < rlim) *dst = make_ulonglong2(alo, ahi);
                        alo = last ? 0ull : alo;
                        ahi = last ? 0ull : ahi;
                        en0 = en1; xu0 = xu1; xv0 = xv1;
                        en1 = en2; xu1 = xu2; xv1 = xv2;
                        en2 = en3;
                    }
                }
                next_stage();
                if (has_next) na_next = norm_block(na_next);  // pass 1 of this CTA's next group
            }
            na = na_next;
            if (threadIdx.x == 64 && s_misc[2]) p.bad_list[atomicAdd(p.bad_count, 1)] = g;
            tk[6] = clock64();
            if (p.dbg && threadIdx.x == 64 && g == (int)blockIdx.x + ((int)(blockIdx.x + gridDim.x) < p.G ? (int)gridDim.x : 0))  // steady state: the second group
                for (int i = 0; i < 7; ++i) p.dbg[blockIdx.x * 8 + i] = tk[i];
            named_bar_sync(1, TROWT);  // shared lists are rebuilt by the next group
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

static int g_tc_grid = 0;

int launch_triplet_batchhard(const float *emb, const int32_t *labels, int64_t G, int64_t B, int d, float margin,
                             float *loss, float *grad, int32_t *trip, float *scratch, int32_t *work, cudaStream_t st) {
    if (G == 0 || B == 0) return VQ_OK;
    const bool fused = B <= TB && d % 4 == 0 && d >= 4 && ((uintptr_t)emb & 15u) == 0 &&
                       (!grad || ((uintptr_t)grad & 15u) == 0) && G * B < (1ll << 31) && !g_hooks.no_tc_triplet;
    if (!fused) return launch_triplet_batchhard_simt(emb, labels, G, B, d, margin, loss, grad, trip, scratch, nullptr, nullptr, st);
    CUtensorMap tm;
    VQ_TRY(make_tensor_map_2d(&tm, emb, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (uint64_t)d, (uint64_t)(G * B), (uint64_t)d * 4,
                              TKB, TB));
    if (!g_tc_grid) {
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        g_tc_grid = sms > 0 ? sms : 148;
    }
    VQ_CUDA(cudaMemsetAsync(work, 0, 4, st));
    TripletTcParams p{};
    p.labels = labels; p.G = (int)G; p.B = (int)B; p.d = d; p.margin = margin;
    p.loss = loss; p.grad = grad; p.trip = trip; p.bad_count = work; p.bad_list = work + 1;
    static long long *dbg_buf = nullptr;
    if (g_hooks.triplet_clocks) {
        if (!dbg_buf) VQ_CUDA(cudaMalloc(&dbg_buf, 148 * 8 * 8));
        p.dbg = dbg_buf;
    }
    VQ_CUDA(cudaFuncSetAttribute(triplet_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TcLayout::TOTAL));
    const int grid = (int)std::min<int64_t>(G, g_tc_grid);
    triplet_tc_kernel<<<grid, TTHREADS, TcLayout::TOTAL, st>>>(tm, p);
    VQ_CUDA(cudaGetLastError());
    if (p.dbg) {
        long long h[8];
        VQ_CUDA(cudaMemcpy(h, p.dbg, 64, cudaMemcpyDeviceToHost));
        fprintf(stderr, "triplet_tc clocks (CTA 0, its second group if it has one): pass1 %lld  wait_gram %lld  mine %lld  pass2 %lld  lists %lld  pass3 %lld\n",
                h[1] - h[0], h[2] - h[1], h[3] - h[2], h[4] - h[3], h[5] - h[4], h[6] - h[5]);
    }
    // groups handed back (capacity overflow): CUDA-core path over the device list, a few resident blocks
    return launch_triplet_batchhard_simt(emb, labels, std::min<int64_t>(G, 8), B, d, margin, loss, grad, trip, scratch,
                                         work + 1, work, st);
}

}  // namespace vq
