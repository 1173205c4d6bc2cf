/*
 * vqwild.h -- C ABI of the B200-native (sm_100a) query->gallery matching path.
 *
 * The reference (dongzhuoyao/video-query-in-the-wild) is pure Python and has no FFI of its
 * own; its hot path sits behind four library call surfaces.  Every entry
 * point below names the reference call site (file:line relative to the reference tree) it
 * replaces.  INTEGRATION.md shows the ctypes binding a maintainer adds on the reference side.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; no torch / C++ types in any signature.
 *   - every function returns 0 on success and a negative vq_status otherwise;
 *     vq_last_error() returns a thread-local message for the last failure.
 *   - a pointer argument documented "host or device" is classified with
 *     cudaPointerGetAttributes: device pointers are used in place (zero copy), host
 *     pointers are staged through pinned memory on `stream` (this is the e2e path).
 *   - `stream` is a cudaStream_t passed as void* (0 = default stream).  Calls are
 *     asynchronous for device outputs and synchronise `stream` before returning when an
 *     output pointer is host memory.
 *   - there is NO CPU fallback: without an sm_100 device every compute entry fails with
 *     VQ_ERR_NO_DEVICE / VQ_ERR_ARCH.
 *   - deterministic: no floating-point atomics in any result-bearing reduction.
 *   - canonical score (the parity target): d2(q,g) = fl32(sum_i (q_i-g_i)^2) accumulated
 *     sequentially in i with fmaf; ranking = d2 ascending, ties -> lowest gallery index.
 */
#ifndef VQWILD_H_
#define VQWILD_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VQ_API __attribute__((visibility("default")))

typedef enum {
    VQ_OK = 0,
    VQ_ERR_INVALID = -1,   /* bad argument (dimension mismatch, k out of range, ...) */
    VQ_ERR_NO_DEVICE = -2, /* no CUDA device / driver */
    VQ_ERR_ARCH = -3,      /* device is not sm_100 */
    VQ_ERR_CUDA = -4,      /* a CUDA call failed; see vq_last_error() */
    VQ_ERR_NOMEM = -5,
    VQ_ERR_CAPACITY = -6,  /* a documented internal capacity was exceeded */
    VQ_ERR_STATE = -7      /* call order (search before add, meta missing, ...) */
} vq_status;

/* element types of embedding inputs */
#define VQ_F32 0
#define VQ_BF16 1

/* tensor-core pre-filter precision (how the coarse scores are formed; results are always
 * the canonical fp32 scores, verified -- see DESIGN.md "Exactness") */
#define VQ_PREC_AUTO 0   /* fp32 input -> SPLIT3, bf16 input -> BF16 */
#define VQ_PREC_BF16 1   /* one bf16 MMA pass */
#define VQ_PREC_SPLIT3 2 /* hi*hi + hi*lo + lo*hi bf16 passes (~2^-16 relative) */
#define VQ_PREC_EXACT 3  /* no tensor-core pre-filter: canonical CUDA-core scores for all pairs */

VQ_API int vq_version(void);
VQ_API const char *vq_last_error(void);
/* Test / A-B hooks ("FORCE_FALLBACK", "NO_PAIR", "NO_CHAIN", "DENSE_ROWS", "NO_TC_TRIPLET", "TRIPLET_CLOCKS",
 * "NO_TC_EVAL", "NO_INSERT", "KPRIME", "TRIGGER", "NO_VMAJOR", "NO_NMS_MASKS", "NO_SEG_UNIQUE", "WALKERS", "NO_L2SQ_WIDE").  Their initial values come from the VQ_DEBUG_<NAME> environment variables, read once when the
 * library is loaded; no entry point reads the environment afterwards. */
VQ_API int vq_debug_set(const char *hook, int value);
/* 0 if device `dev` exists and is sm_100; else VQ_ERR_NO_DEVICE / VQ_ERR_ARCH. */
VQ_API int vq_device_check(int dev);

/* ------------------------------------------------------------------------------------
 * b1. faiss.IndexFlatL2 replacement.
 *   ctor   dataloader_baseline.py:1502 (:872, :1196)  faiss.IndexFlatL2(d)
 *   add    :1507 (:877, :1201)                        index.add(xb float32[N,d])
 *   ntotal :1508
 *   search :1592-1595 (:935-938, :1268-1271)          index.search(xq float32[nq,d], k)
 * ---------------------------------------------------------------------------------- */
typedef struct vq_index vq_index;

VQ_API int vq_index_create(int d, vq_index **out);
VQ_API int vq_index_destroy(vq_index *ix);
/* Append n rows (host or device, C-contiguous [n,d], dtype VQ_F32 or VQ_BF16; all adds to one
 * index must use the same dtype).  Builds the bf16 operand planes and row norms. */
VQ_API int vq_index_add(vq_index *ix, const void *xb, int64_t n, int dtype, void *stream);
VQ_API int64_t vq_index_ntotal(const vq_index *ix);
VQ_API int vq_index_dim(const vq_index *ix);
VQ_API int vq_index_set_precision(vq_index *ix, int prec);
/* Per-row metadata used by the fused labelling (a4): label id and video id per gallery row
 * (dataloader_baseline.py:1540-1556 candidate["label"], candidate["video_id"]).  host or device,
 * int32[ntotal].  video may be NULL (no ignore rule). */
VQ_API int vq_index_set_meta(vq_index *ix, const int32_t *label, const int32_t *video, void *stream);
/* Untrimmed-video segments: seg_start int64[nseg+1] (host), rows [seg_start[v], seg_start[v+1])
 * are the moments/windows of video v (dataloader_baseline.py:1091-1151 builds them video by
 * video, so they are contiguous).  Enables vq_index_search_segmax. */
VQ_API int vq_index_set_segments(vq_index *ix, const int64_t *seg_start, int64_t nseg, void *stream);

/* faiss-shaped search.  xq [nq,d] (host or device, dtype as given); outputs D float32[nq,k]
 * (canonical squared L2, ascending) and I int64[nq,k] (host or device; -1 / +inf padding when
 * fewer than k rows qualify).  k <= ntotal.  k <= 512 takes the tensor-core streaming top-k path
 * (gemm_topk + select); larger k -- the reference's own call is k = ntotal, :1594 -- takes the full-sort path:
 * canonical scores of every pair + a stable per-row radix sort (vq_sort.cu), tiled over the queries.
 * ignore_video: optional int32[nq, n_ignore] (host or device) -- rows whose video id matches
 * any entry are skipped (the reference's ignore rule, :1548); requires vq_index_set_meta. */
VQ_API int vq_index_search(vq_index *ix, const void *xq, int64_t nq, int dtype, int64_t k,
                           const int32_t *ignore_video, int n_ignore, float *D, int64_t *I,
                           void *stream);

/* Workspace of the k <= 512 search path (candidate lists, work plan).  It belongs to the index and is grown on demand,
 * stream-ordered on the caller's stream from the library's own memory pool (no device synchronisation);
 * vq_index_search_workspace_bytes reports what a search of nq queries would need -- persistent_bytes stay with the
 * index, transient_bytes are stream-ordered temporaries of one call -- and vq_index_reserve grows the persistent part
 * ahead of time, after which searches of up to that shape allocate nothing but the temporaries.  Calls on one index
 * must be ordered with respect to each other (same stream, or synchronised): they share this workspace. */
VQ_API int vq_index_search_workspace_bytes(vq_index *ix, int64_t nq, int64_t k, int segmax, int64_t *persistent_bytes,
                                           int64_t *transient_bytes);
VQ_API int vq_index_reserve(vq_index *ix, int64_t nq, int64_t k, int segmax, void *stream);

/* configs[2]: per (query, video) best moment (segment max-pool) and the top-k videos.
 * Outputs [nq,k]: video id (int64), arg-best gallery row (int64), its canonical d2. */
VQ_API int vq_index_search_segmax(vq_index *ix, const void *xq, int64_t nq, int dtype, int64_t k,
                                  int64_t *video, int64_t *row, float *D, void *stream);

/* Statistics of the last search on this index (host ints): [0] queries that needed the exact
 * fallback, [1] tensor-core tiles issued, [2] candidate slots per query, [3] work units,
 * [4] gallery segments (stream-K pieces of the query-group x gallery space), [5] kernels launched.  After
 * vq_index_eval_trimmed: [4] = queries its tensor-core form handed back to the canonical all-pairs path. */
VQ_API int vq_index_last_stats(const vq_index *ix, int64_t *out6);
/* Device time (ms, CUDA events on the caller's stream) of the phases of the last tensor-core
 * search on this index: [0] query operand preparation, [1] the tcgen05 GEMM + streaming
 * top-k kernel, [2] select/verify, [3] exact fallback.  Waits for that search to finish. */
VQ_API int vq_index_last_timing(const vq_index *ix, float *out4);

/* ------------------------------------------------------------------------------------
 * b2. Evaluation: ARV_Retrieval.ranking's per-query loop (dataloader_baseline.py:1526-1610)
 * = labelling (a4) + scoring/ranking (a3) + evaluation_metric.add2dict (a6, :431-496).
 *   q_label int32[nq], ignore_video int32[nq,n_ignore] (host or device)
 *   ap float64[nq]; recall float64[nq,3] (R@30/50/100, :332); top_idx int64[nq,100] (the first
 *   100 survivors in rank order -- the rows behind cm_dict["label"], :441-442); may be NULL.
 * Requires vq_index_set_meta.  Full-list semantics (k = ntotal): AP over all survivors with
 * the last one forced positive (:434).  Scores come from the tensor cores (coarse) and canonical arithmetic is spent only
 * on the query's positives and on survivors within the proven error bound of one of them; the numbers returned are
 * those of the canonical all-pairs evaluation (which queries with > 1022 positives or > 2048 such survivors still take).
 * ---------------------------------------------------------------------------------- */
VQ_API int vq_index_eval_trimmed(vq_index *ix, const void *xq, int64_t nq, int dtype,
                                 const int32_t *q_label, const int32_t *ignore_video,
                                 int n_ignore, double *ap, double *recall, int64_t *top_idx,
                                 void *stream);

/* ------------------------------------------------------------------------------------
 * a8. Moment retrieval evaluation: the per-query body of ARV_Retrieval_Moment.ranking
 * (dataloader_baseline.py:1225-1330) = labelling (:1239-1261) + ranking of ALL moments
 * (:1268-1281) + do_nms (:1283-1312: cluster by video in first-seen order, nms_cpu per cluster)
 * + tp = IoU >= 0.5 (:1315-1319) + multiprocess_calculate (:386-402) on the grouped list.
 *   vq_index_set_moments  per gallery row (host or device, [ntotal]): start_sec, end_sec (f32),
 *                         closest_hit IoU (f32; 0 when there is no hit; tp tests iou >= 0.5f -- the reference
 *                         compares the FLOAT64 IoU, :1315-1319, so callers holding float64 values pass the flag
 *                         (iou64 >= 0.5 ? 1 : 0) to keep boundary cases exact; the Python mirror does) and
 *                         closest_hit label id (int32; -1 when there is none).  Requires vq_index_set_segments (one
 *                         segment per video, <= 4096 moments each) and vq_index_set_meta (video).
 *   vq_index_eval_moment  q_label int32[nq], ignore_video int32[nq,n_ignore];
 *                         ap float64[nq], recall float64[nq,3]; optional n_kept int64[nq] (moments
 *                         surviving NMS, before the ignore rule) and kept_mask uint8[nq,ntotal].
 * ---------------------------------------------------------------------------------- */
VQ_API int vq_index_set_moments(vq_index *ix, const float *start_sec, const float *end_sec,
                                const float *iou, const int32_t *hit_label, void *stream);
VQ_API int vq_index_eval_moment(vq_index *ix, const void *xq, int64_t nq, int dtype,
                                const int32_t *q_label, const int32_t *ignore_video, int n_ignore,
                                double nms_thresh, double *ap, double *recall, int64_t *n_kept,
                                uint8_t *kept_mask, void *stream);

/* sklearn.metrics.average_precision_score for a batch of lists (dataloader_baseline.py:391,
 * :461): lists are concatenated, list s = [offsets[s], offsets[s+1]).  y_true int32, y_score
 * float64 (any order; sorted internally, ties grouped).  host or device; ap float64[nlists]. */
VQ_API int vq_average_precision(const int32_t *y_true, const double *y_score,
                                const int64_t *offsets, int64_t nlists, double *ap, void *stream);

/* ------------------------------------------------------------------------------------
 * b3. nms_cpu (misc_utils/utils_models.py:153-174), segmented: dets float64[total,3] =
 * (x1,x2,score); segment s = rows [offsets[s], offsets[s+1]).  keep uint8[total] (1 = kept).
 * Visiting order inside a segment: score descending, ties lowest row.
 * ---------------------------------------------------------------------------------- */
VQ_API int vq_nms1d_segmented(const double *dets, const int64_t *offsets, int64_t nseg,
                              double thresh, uint8_t *keep, void *stream);

/* vq_nms1d_segmented for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.
 * Same kernel and result as vq_nms1d_segmented; its scratch (one byte per detection) is a file-scope device array
 * of VQ_NMS1D_STATIC_MAX bytes that every GPU loading the library holds until exit (csrc/vq_noalloc.cu).  The call
 * allocates nothing, copies nothing and synchronises nothing: dets, offsets (int64[nseg+1]) and keep must be DEVICE
 * memory; total = offsets[nseg] is passed by the caller (the library does not read device memory on the host) and
 * total > VQ_NMS1D_STATIC_MAX returns VQ_ERR_CAPACITY.  The first call on a device resolves the array's address
 * (cudaGetSymbolAddress; a call with nseg == 0 does only that -- make it before a graph capture).  All calls on a
 * device share the array: calls that overlap on the GPU race and must be ordered by the caller. */
#define VQ_NMS1D_STATIC_MAX 1048576
VQ_API int vq_nms1d_segmented_static(const double *dets, const int64_t *offsets, int64_t nseg, int64_t total,
                                     double thresh, uint8_t *keep, void *stream);

/* ------------------------------------------------------------------------------------
 * b4. Metric-learning loss (misc_utils/metric_learning_loss.py).
 *   vq_pdist            Sampler.pdist :276-289          A f32[B,d] -> dist f32[B,B]
 *   vq_triplet_batchhard  TripletLoss.forward :357-379 with in-kernel batch-hard mining
 *                       (extension; masks as :173-176) + backward.  G independent groups of B
 *                       rows: emb f32[G,B,d], labels int32[G,B] -> loss f32[G], grad f32[G,B,d],
 *                       trip int32[G,B,3] (a,p,n; -1 rows for skipped anchors), may be NULL.
 *                       B <= 256 and d % 4 == 0 run the fused tensor-core kernel (TF32 Gram pre-filter,
 *                       results identical to the CUDA-core path); other shapes the CUDA-core path.
 *   vq_triplet_given    same loss/backward for caller-sampled triplets int32[T,3] (the
 *                       reference's random/semihard/distance samplers run on the host).
 * All pointers host or device.
 * ---------------------------------------------------------------------------------- */
VQ_API int vq_pdist(const float *A, int64_t B, int d, float eps, float *dist, void *stream);
VQ_API int vq_triplet_batchhard(const float *emb, const int32_t *labels, int64_t G, int64_t B,
                                int d, float margin, float *loss, float *grad, int32_t *trip,
                                void *stream);
/* Diagnostics: groups of the last vq_triplet_batchhard call that the fused tensor-core kernel handed
 * back to the CUDA-core path (candidate capacity exceeded); waits for the device. */
VQ_API int vq_triplet_last_handed_back(void);
VQ_API int vq_triplet_given(const float *emb, int64_t B, int d, const int32_t *trip, int64_t T,
                            float margin, float *loss, float *grad, void *stream);

/* ------------------------------------------------------------------------------------
 * f4. The other criteria of misc_utils/metric_learning_loss.py on the tuples the host samplers drew
 * (same forward(batch, labels) modules on the Python side); f3. the memory-distance logits of the
 * VA / VASA heads.  fp32; all pointers host or device; gradients may be NULL (forward only).
 *   vq_margin_given   MarginLoss.forward :520-565 (nu = 0): trip int32[T,3], beta_t f32[T] = the margin
 *                     parameter of each triplet's anchor class (:543-547) or the constant; outputs loss f32[1],
 *                     grad f32[B,d], grad_beta_t f32[T] (scatter to classes on the caller's side), pair_count
 *                     f32[1] = triplets with an active term (:552-554: bool + bool is a logical or).
 *   vq_npair_given    NPairLoss.forward :397-457: tuples int32[P, 2+N] = (anchor, positive, N negatives).
 *   vq_proxynca       ProxyNCALoss.forward :593-635: emb f32[B,d], proxies f32[C,d], labels int32[B].
 *   vq_sphere_weights Sampler.inverse_sphere_distances :291-324 for every anchor: dist f32[B,B] (pdist clamped
 *                     at the lower cutoff, :215), labels int32[B], dim = embedding size -> q f32[B,B], rows sum to 1.
 *   vq_memory_logits  models/resnet18_va.py:172-184, resnet18_vasa.py:198-206, :217-229:
 *                     logits[b,k] = -|normalize(x_b) - mem_k|_2 / temperature (f32[B,K]); with grad_logits
 *                     f32[B,K] also grad_x f32[B,d] (through the normalisation) and/or grad_mem f32[K,d].
 * ---------------------------------------------------------------------------------- */
VQ_API int vq_margin_given(const float *emb, int64_t B, int d, const int32_t *trip, int64_t T,
                           const float *beta_t, float margin, float *loss, float *grad,
                           float *grad_beta_t, float *pair_count, void *stream);
VQ_API int vq_npair_given(const float *emb, int64_t B, int d, const int32_t *tuples, int64_t P, int N,
                          float l2, float *loss, float *grad, void *stream);
VQ_API int vq_proxynca(const float *emb, const float *proxies, const int32_t *labels, int64_t B,
                       int64_t C, int d, float *loss, float *grad_emb, float *grad_proxies, void *stream);
VQ_API int vq_sphere_weights(const float *dist, const int32_t *labels, int64_t B, int dim, float *q,
                             void *stream);
VQ_API int vq_memory_logits(const float *x, const float *mem, int64_t B, int64_t K, int d,
                            float temperature, float *logits, const float *grad_logits, float *grad_x,
                            float *grad_mem, void *stream);

/* a1. Pre-match pooling: F.normalize(p=2, dim=1) per time step (main.py:232) + temporal mean
 * (dataloader_baseline.py:1481).  x f32[B,C,T] -> out f32[B,C]. */
VQ_API int vq_pool_normalize(const float *x, int64_t B, int C, int T, float *out, void *stream);

/* f1. Moment / clip window pooling, the step before the index build (garner_feat,
 * dataloader_baseline.py:1125-1147 and :809-834: feat = np.mean(video_feat[:, lo:hi], axis=1) per window).
 *   feat f32[C, Ttot]: per-frame features of V videos concatenated along time, video v = frames
 *   [vid_t0[v], vid_t0[v+1]) (vid_t0 int64[V+1], HOST); windows grouped by video: video v owns windows
 *   [win_start[v], win_start[v+1]) (int64[V+1], HOST); window w averages frames [lo[w], hi[w]) of its video
 *   (int32, relative to the video's first frame, host or device).  out f32[n_windows, C] (host or device). */
VQ_API int vq_window_means(const float *feat, int C, int64_t Ttot, const int64_t *vid_t0,
                           const int64_t *win_start, int64_t V, const int32_t *lo, const int32_t *hi,
                           float *out, void *stream);

/* 8e. Merge per-shard top-k lists after the all-gather: d2 f32[G,nq,k], idx int64[G,nq,k]
 * (global row ids, -1 = empty) -> out [nq,k] by (d2 asc, idx asc). */
VQ_API int vq_topk_merge(const float *d2, const int64_t *idx, int G, int64_t nq, int64_t k,
                         float *out_d2, int64_t *out_idx, void *stream);

/* 8e, packed form of the same exchange: a rank turns its local list (d2 f32[count], idx int64[count] LOCAL rows,
 * -1 = empty) into one 8-byte key per entry, (order-preserving d2 bits << 32 | row_offset + idx; all ones = empty;
 * global rows must fit 32 bits), all-gathers the keys in ONE collective (8 instead of 12 bytes per entry) and
 * vq_topk_merge_keys sorts keys u64[G,nq,k] into the same [nq,k] answer vq_topk_merge gives. */
VQ_API int vq_topk_pack(const float *d2, const int64_t *idx, int64_t count, int64_t row_offset,
                        uint64_t *keys, void *stream);
VQ_API int vq_topk_merge_keys(const uint64_t *keys, int G, int64_t nq, int64_t k, float *out_d2,
                              int64_t *out_idx, void *stream);

/* 8e, two-round form of the exchange: with G shards a query's global top-k takes ~k/G entries from each, so the ranks
 * first all-gather only the first m < k keys of their (sorted) lists.  keys u64[G,nq,m] -> out [nq,k]; `incomplete`
 * (int32[1], host or device) receives the number of queries whose merge is NOT final -- some shard's head ends before
 * the merged k-th key, so it may hold better entries it has not sent -- in which case the caller repeats the exchange
 * with the full lists (vq_topk_merge_keys).  0 incomplete queries = the result equals the full merge. */
VQ_API int vq_topk_merge_heads(const uint64_t *keys, int G, int64_t nq, int64_t m, int64_t k, float *out_d2,
                               int64_t *out_idx, int32_t *incomplete, void *stream);

/* Canonical all-pairs squared L2 (CUDA cores), out f32[nq,n]; test / small-problem utility. */
VQ_API int vq_l2sq(const float *q, const float *g, int64_t nq, int64_t n, int d, float *out,
                   void *stream);

#ifdef __cplusplus
}
#endif
#endif /* VQWILD_H_ */
