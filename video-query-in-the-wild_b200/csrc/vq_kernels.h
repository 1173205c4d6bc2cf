// vq_kernels.h -- internal launch interfaces between the C-ABI layer (vq_api.cu) and the kernels.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace vq {

// Debug / A-B hooks (tests, scripts/ab_*.py): initialised ONCE from the VQ_DEBUG_* environment when the library is
// loaded and changed afterwards only through vq_debug_set() -- never getenv() on a search path.
struct DebugHooks {
    int force_fallback = 0, no_pair = 0, no_chain = 0, dense_rows = -1, no_tc_triplet = 0, triplet_clocks = 0,
        no_tc_eval = 0, no_insert = 0, kprime = 0, trigger = 0, no_vmajor = 0, no_nms_masks = 0, no_seg_unique = 0, walkers = 0, no_l2sq_wide = 0;
};
extern DebugHooks g_hooks;

static constexpr int kTileM = 128;      // queries per work unit (UMMA M)
static constexpr int kTileN = 256;      // gallery rows per tile (UMMA N)
static constexpr int kCandCap = 1024;   // candidate slots per (work unit, query)
static constexpr int kMaxIgnore = 8;    // ignore-list length per query (the reference uses 5)
static constexpr int kMaxKPrime = 512;  // largest pre-filter list (must be <= kCandCap - kTileN)
// Operand planes are [rows, dp + kAugCols] bf16: dp = dim padded to 64, then one 16-column
// "augmentation" block that folds the row-norm bias of the squared-L2 score into the GEMM:
//   gallery row: (h, m, l, 0...) with h + m + l = -|g|^2 / 2 split into three bf16 terms
//   query row  : (1, 1, 1, 0...)
// so the accumulator is <q,g> - |g|^2/2 = (coarse score)/2 and the epilogue needs no per-column
// parameters at all.
static constexpr int kAugCols = 16;

// Work decomposition: the gallery is cut into chunks (row ranges); a SEGMENT is one chunk of one
// query group (a 128-query block, or a pair of them for the CTA-pair kernel) and streams into its own
// candidate lists.  Each persistent CTA (pair) owns a contiguous run of the segment array; the host
// orders that array so that the CTAs sweep a chunk together (L2 sharing of the gallery tiles).
// seg_pred: the segment of the same group on the previous chunk (or -1): a segment starts from its
// predecessor's final threshold when that has already been published, skipping the warm-up.
struct GemmTopkParams {
    int64_t nq, n;
    int kb_count;              // padded dim / 64 (full k-blocks; the augmentation block follows them)
    int n_qblocks;             // 128-query blocks (even for the pair kernel)
    int n_seg;
    const int64_t *seg_begin, *seg_end;  // device [n_seg] gallery rows of each segment
    const int32_t *seg_group;            // device [n_seg] query group of each segment
    const int32_t *cta_seg_start;        // device [n_cta+1] segments [cta_seg_start[c], cta_seg_start[c+1]) belong to CTA (pair) c
    const int32_t *seg_pred;             // device [n_seg] predecessor segment or -1
    const int32_t *col_video;    // device [n] or null
    const int32_t *col_seg;      // device [n] or null (segment id per row, SEGMAX)
    const int32_t *q_ignore;     // device [nq, n_ign] or null
    int n_ign;
    int kprime;
    float *store;       // STORE mode: coarse scores out [nq, store_ld] (or null: top-k pre-filter mode)
    int64_t store_ld;
    int dbg_no_insert;  // debug hook NO_INSERT
    int trigger;        // list fill level that triggers an intermediate compaction (0: 2 k')
    int dense_rows;  // rows of a warp with candidates in one 32-column step from which every thread scans its own row
    // unit = (segment, query block inside the group): u = seg * group_size + member
    uint64_t *cand;   // [units][kTileM][kCandCap] keys (ordered coarse score << 32 | ~row)
    int32_t *cand_cnt;  // [units][kTileM]
    float *cand_tau;    // [units][kTileM] everything not listed has coarse score <= tau (NaN until the unit is done)
    int32_t *cand_cum;  // [units][kTileM] 1: the list continues the predecessor's (it covers all earlier chunks of the row)
    // SEGMAX, optional: which (query, video) candidates have a best moment that is NOT alone at the top -- some other
    // moment of the video may lie within the error band of the coarse maximum (or could not be ruled out).  Every other
    // candidate's exact best moment is the coarse arg-max and needs ONE canonical score instead of the whole video.
    uint32_t *seg_amb;      // [nq][amb_words] bit v: video v is ambiguous for this query (zeroed by the caller) or null
    int amb_words;
    const float *row_eps;   // [nq] proven |coarse - (|q|^2 - d2)| bound of the row's scores
};

int launch_gemm_topk(const CUtensorMap *maps4, const GemmTopkParams &p, int npass, bool segmax,
                     int grid, cudaStream_t st);

// CTA-pair variant (bf16 single pass, padded dim <= 512): maps2 = {queries, gallery}, both with
// boxes of 64 columns x 128 rows; p.n_qblocks must be even (pad with an out-of-range block).
int launch_gemm_topk_pair(const CUtensorMap *maps2, const GemmTopkParams &p, bool segmax, int n_pairs,
                          cudaStream_t st);
int gemm_topk_pair_slots();  // co-resident CTA pairs on this device

// ---- vq_prepare.cu ----
// fp32 or bf16 rows [n,d] -> bf16 planes hi/lo [n, dp + kAugCols] (zero padded; augmentation block
// as described above: aug_query selects the query form), norm2 [n] (may be null); max_norm2
// (ordered-uint atomicMax target, may be null).
int launch_prepare_rows(const void *x, int dtype, int64_t n, int d, int dp, bool aug_query, __nv_bfloat16 *hi,
                        __nv_bfloat16 *lo, float *norm2, uint32_t *max_norm2_bits, cudaStream_t st);
int launch_pool_normalize(const float *x, int64_t B, int C, int T, float *out, cudaStream_t st);

// ---- vq_exact.cu ----
// canonical all-pairs d2 for queries [q_begin, q_begin+nq_tile) -> out[nq_tile, n]
int launch_l2sq_all(const void *q, const void *g, int dtype, int64_t nq_tile, int64_t n, int d,
                    float *out, cudaStream_t st);

struct SelectParams {
    int64_t nq, n;
    int d;
    int dtype;           // element type of q/g (VQ_F32 / VQ_BF16)
    const void *q, *g;   // original rows (canonical re-scoring reads these)
    int group_size;                  // query blocks per group (2 for the pair kernel)
    const int32_t *group_segs;       // device [n_groups][segs_per_group]: segment ids of group g
    int segs_per_group;
    const uint64_t *cand;
    const int32_t *cand_cnt;
    const float *cand_tau;
    const int32_t *cand_cum;     // 1: the unit's list already covers every earlier segment of the row (vq_gemm_topk.cu)
    int kprime, k;
    float eps_rel, eps_abs;      // |coarse - (|q|^2 - d2)| <= eps_rel*|q|*gmax + eps_abs
    const uint32_t *gmax_norm2_bits;
    // segmax
    const int32_t *seg_of_row;   // [n] or null
    const int64_t *seg_start;    // [nseg+1] or null
    // outputs
    float *out_d2;     // [nq,k]
    int64_t *out_idx;  // [nq,k] gallery row (segmax: arg-best row)
    int64_t *out_seg;  // [nq,k] or null (segmax: video id)
    int32_t *unproven;       // [nq] list of query ids that need the exact fallback
    int32_t *unproven_count; // [1]
    // segmax, video-major re-scoring (launch_select_segmax_vmajor sets these): the select kernel only emits the pairs
    int32_t *pair_vid;       // [nq, kprime] surviving video of each slot (-1 = none) or null
    int32_t *pair_m;         // [nq]
    float *pair_tau;         // [nq]
    int32_t *vcount;         // [nseg] pairs per video
    const uint32_t *seg_amb; // [nq][amb_words] from the GEMM (GemmTopkParams::seg_amb) or null: every pair is re-scored in full
    int amb_words;
    int32_t *pair_one;       // [nq, kprime] arg-max row of a pair whose best moment is proven unique (-1 otherwise)
};
int launch_segmax_row_eps(const SelectParams &p, float *row_eps, cudaStream_t st);
int launch_select(const SelectParams &p, cudaStream_t st);
size_t select_vmajor_scratch_bytes(int64_t nq, int kprime, int64_t nseg);
int launch_select_segmax_vmajor(SelectParams p, int64_t nseg, void *scratch, cudaStream_t st);

struct FallbackParams {
    int64_t nq, n;
    int d, dtype;
    const void *q, *g;
    const int32_t *col_video, *q_ignore;
    int n_ign;
    const int32_t *seg_of_row;
    const int64_t *seg_start;
    int64_t nseg;
    int k;
    float *out_d2;
    int64_t *out_idx;
    int64_t *out_seg;
    const int32_t *unproven;        // null => all queries
    const int32_t *unproven_count;  // null => nq
};
int launch_exact_topk(const FallbackParams &p, int grid, cudaStream_t st);

// ---- vq_sort.cu ----
// Full ascending list per query (faiss search with k = ntotal, dataloader_baseline.py:1592-1597): stable LSD radix
// sort of the canonical score tile d2[nq_tile, n]; the first k entries of every sorted row go to out_d2 / out_idx
// ([nq_tile, k]); rows hit by the ignore rule sort last and are reported as +inf / -1.
struct RowSortParams {
    int64_t nq_tile, n, k;
    const float *d2;            // [nq_tile, n] canonical
    const int32_t *col_video;   // [n] or null
    const int32_t *q_ignore;    // [nq_tile, n_ign] or null
    int n_ign;
    int32_t *hist;              // [rowsort_hist_ints(nq_tile, n)]
    float *out_d2;
    int64_t *out_idx;
    const uint64_t *in;         // set per pass by launch_row_sort
    uint64_t *out;
};
size_t rowsort_hist_ints(int64_t nq_tile, int64_t n);
// bufA / bufB: u64[nq_tile * n] each
int launch_row_sort(RowSortParams p, uint64_t *bufA, uint64_t *bufB, cudaStream_t st);

// ---- vq_metrics.cu ----
struct ApRowsParams {
    int64_t nq_tile, n;
    const float *d2;           // [nq_tile, n] canonical
    const int32_t *g_label, *g_video;
    const int32_t *q_label;    // [nq_tile]
    const int32_t *q_ignore;   // [nq_tile, n_ign] or null
    int n_ign;
    double *ap;        // [nq_tile]
    double *recall;    // [nq_tile,3]
    int64_t *top_idx;  // [nq_tile,100] or null
    int32_t *overflow; // [1] set when a query has more positives than kMaxPositives
    // moment mode (kept != null): g_label = closest-hit label of each moment
    const uint8_t *kept;         // [nq_tile, n] survivors of the per-video NMS
    const float *g_iou;          // [n]
    uint64_t *vbest;             // [nq_tile, nseg] key of each video's best moment (ignored videos are blanked)
    const int64_t *seg_start;    // [nseg+1]
    const int32_t *seg_of_row;   // [n]
    int64_t nseg;
    int64_t *n_kept;             // [nq_tile] or null
};
// tensor-core form of the trimmed / clip evaluation (vq_metrics.cu: ap_rows_tc_kernel)
struct ApTcParams {
    int64_t nq_tile, n, ld;
    int d, dtype;
    const float *coarse;         // [nq_tile, ld] coarse scores 2<q,g> - |g|^2 (gemm_topk STORE mode)
    const void *q, *g;           // original rows (canonical re-scoring)
    const int32_t *g_label, *g_video, *q_label, *q_ignore;
    int n_ign;
    float eps_rel, eps_abs;      // |coarse - (|q|^2 - d2)| <= eps_rel*|q|*gmax + eps_abs*(|q|^2 + gmax^2)
    const uint32_t *gmax_norm2_bits;
    double *ap, *recall;
    int32_t *redo, *redo_count;  // queries (tile-relative) the canonical path has to evaluate
};
int launch_ap_rows_tc(const ApTcParams &p, cudaStream_t st);
struct MomentNmsParams {
    int64_t nq_tile, n, nseg;
    const float *d2;             // [nq_tile, n] canonical
    const int64_t *seg_start;    // [nseg+1]
    const float *start, *end;    // [n] seconds
    double thresh;
    uint8_t *kept;               // [nq_tile, n]
    uint64_t *vbest;             // [nq_tile, nseg]
    const unsigned long long *sup;  // [n] suppression masks of the videos of <= 64 moments (launch_nms_masks) or null
};
int launch_nms_masks(const int64_t *seg_start, const int32_t *small_list, int n_small, const float *start, const float *end,
                     double thresh, unsigned long long *sup, cudaStream_t st);
int launch_window_means(const float *feat, int C, int64_t Ttot, const int64_t *vid_t0, const int64_t *win_start, int64_t V,
                        int64_t Tmax, const int32_t *lo, const int32_t *hi, float *out, cudaStream_t st);
// vq_losses.cu (f3 / f4)
int launch_memory_logits(const float *x, const float *mem, int B, int K, int d, float temperature, float *logits,
                         const float *G, float *gx, float *gmem, float *scratch, cudaStream_t st);
size_t memory_logits_scratch_floats(int B, int K, int d);
int launch_proxynca(const float *emb, const float *proxies, const int32_t *labels, int B, int C, int d, float *loss,
                    float *gemb, float *gprox, float *scratch, cudaStream_t st);
size_t proxynca_scratch_floats(int B, int C, int d);
int launch_margin_given(const float *emb, int64_t B, int d, const int32_t *trip, int64_t T, const float *beta_t,
                        int beta_stride, float margin, float *loss, float *grad, float *gbeta, float *pair_count,
                        float *scratch, cudaStream_t st);
int launch_npair_given(const float *emb, int64_t B, int d, const int32_t *tup, int P, int N, float l2, float *loss,
                       float *grad, float *scratch, cudaStream_t st);
int launch_sphere_weights(const float *dist, const int32_t *labels, int B, int dim, float *q, cudaStream_t st);
static constexpr int kMomentWarpCap = 256, kMomentBlockCap = 4096;
int launch_moment_nms(const MomentNmsParams &p, const int32_t *small_list, int n_small, const int32_t *large_list,
                      int n_large, cudaStream_t st);
static constexpr int kMaxPositives = 8192;
// pcap2: power-of-two capacity of the per-query positive list (<= kMaxPositives)
int launch_ap_rows(const ApRowsParams &p, int pcap2, cudaStream_t st);
int launch_average_precision(const int32_t *y_true, const double *y_score, const int64_t *offsets,
                             int64_t nlists, int64_t max_len, double *ap, int32_t *overflow,
                             cudaStream_t st);
int launch_nms1d(const double *dets, const int64_t *offsets, int64_t nseg, double thresh,
                 uint8_t *keep, uint8_t *dead_scratch, cudaStream_t st);
int launch_topk_pack(const float *d2, const int64_t *idx, int64_t count, int64_t row_offset, uint64_t *keys, cudaStream_t st);
int launch_topk_merge_keys(const uint64_t *keys, int G, int64_t nq, int64_t k, float *od, int64_t *oi, cudaStream_t st);
int launch_topk_merge_heads(const uint64_t *keys, int G, int64_t nq, int m, int k, float *od, int64_t *oi, int32_t *incomplete,
                            cudaStream_t st);
int launch_topk_merge(const float *d2, const int64_t *idx, int G, int64_t nq, int64_t k, float *od,
                      int64_t *oi, cudaStream_t st);

// 2-D row-major tensor map (vq_api.cu): box [box_cols x box_rows], 128-byte swizzle, zero OOB fill
int make_tensor_map_2d(CUtensorMap *m, const void *base, CUtensorMapDataType dt, int elem_bytes, uint64_t cols,
                       uint64_t rows, uint64_t pitch_bytes, uint32_t box_cols, uint32_t box_rows);

// ---- vq_triplet.cu ----
int launch_pdist(const float *A, int64_t B, int d, float eps, float *dist, cudaStream_t st);
// dispatcher (vq_triplet_tc.cu): fused tensor-core kernel when B <= 256 and d % 4 == 0, groups it hands
// back (and every other shape) go to the CUDA-core path.  scratch: float[G*B]; work: int32[G+1] device.
int launch_triplet_batchhard(const float *emb, const int32_t *labels, int64_t G, int64_t B, int d,
                             float margin, float *loss, float *grad, int32_t *trip, float *scratch,
                             int32_t *work, cudaStream_t st);
int launch_triplet_batchhard_simt(const float *emb, const int32_t *labels, int64_t G, int64_t B, int d,
                                  float margin, float *loss, float *grad, int32_t *trip, float *scratch,
                                  const int32_t *glist, const int32_t *gcount, cudaStream_t st);
int launch_triplet_given(const float *emb, int64_t B, int d, const int32_t *trip, int64_t T,
                         float margin, float *loss, float *grad, float *scratch, cudaStream_t st);

}  // namespace vq
