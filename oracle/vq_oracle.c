/*
 * vq_oracle.c -- CPU restatement of the reference's query->gallery matching path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may link or call it.  The product
 * (video-query-in-the-wild_b200/) never routes through this file.
 *
 * Parity status: the reference holds NO golden vectors for this path;
 * faiss (un-vendored, unpinned) is not installable here.  The oracle is pinned instead
 * against outputs of the reference's own Python code run in the authoring container
 * (tests/golden/make_golden.py -> tests/golden/ npz files): evaluation_metric, nms_cpu,
 * ARV_Retrieval.ranking, ARV_Retrieval_Moment.ranking, TripletLoss.  At the faiss
 * boundary ("squared L2, ascending, all N") parity is UNPINNED: the canonical score
 * below is our statement of it.
 *
 * Each function cites the reference file:line (relative to /root/reference) it follows.
 *
 * Build:  gcc -O2 -fPIC -shared -ffp-contract=off -fopenmp -o libvqoracle.so vq_oracle.c -lm
 * (-ffp-contract=off: every fmaf below is explicit, nothing else may be fused.)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------
 * Canonical score.  dataloader_baseline.py:1592-1597 (faiss.IndexFlatL2.search) and
 * :1602 (score = -float(D[j])).  d2 = fl32( sum_{i=0}^{D-1} (q_i-g_i)^2 ), accumulated
 * sequentially in index order with fmaf (round-to-nearest-even, no FTZ).  A CUDA-core
 * kernel using __fsub_rn/__fmaf_rn reproduces these bits exactly.
 * ---------------------------------------------------------------------------------- */
static inline float orc_d2(const float *q, const float *g, int d) {
    float acc = 0.0f;
    for (int i = 0; i < d; ++i) {
        float df = q[i] - g[i];
        acc = fmaf(df, df, acc);
    }
    return acc;
}

ORC_API int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

ORC_API void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* out[nq*n] squared L2, row-major. */
ORC_API void orc_l2sq(const float *q, const float *g, int64_t nq, int64_t n, int d,
                      float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t a = 0; a < nq; ++a)
        for (int64_t b = 0; b < n; ++b) out[a * n + b] = orc_d2(q + a * d, g + b * d, d);
}

/* Gathered variant: d2 of (q[qi[t]], g[gi[t]]) for t<m. */
ORC_API void orc_l2sq_pairs(const float *q, const float *g, const int64_t *qi,
                            const int64_t *gi, int64_t m, int d, float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < m; ++t) out[t] = orc_d2(q + qi[t] * d, g + gi[t] * d, d);
}

/* bf16 (round-to-nearest-even) helpers: the bf16-input mode converts inputs first and
 * then runs the same canonical chain on the widened values. */
static inline float orc_bf16_round(float x) {
    uint32_t u;
    memcpy(&u, &x, 4);
    if ((u & 0x7f800000u) == 0x7f800000u) return x; /* inf/nan unchanged */
    uint32_t lsb = (u >> 16) & 1u;
    u += 0x7fffu + lsb;
    u &= 0xffff0000u;
    float r;
    memcpy(&r, &u, 4);
    return r;
}

ORC_API void orc_bf16_round_array(const float *x, int64_t n, float *out) {
    for (int64_t i = 0; i < n; ++i) out[i] = orc_bf16_round(x[i]);
}

/* ------------------------------------------------------------------------------------
 * Full ranking of one query row.  faiss returns all N ascending in d2
 * (dataloader_baseline.py:1592-1594, k = N); tie order inside faiss is unspecified, the
 * stated tie-break is lowest gallery index first.
 * ---------------------------------------------------------------------------------- */
typedef struct {
    float d2;
    int64_t idx;
} orc_pair;

static int orc_pair_cmp(const void *a, const void *b) {
    const orc_pair *x = (const orc_pair *)a, *y = (const orc_pair *)b;
    if (x->d2 < y->d2) return -1;
    if (x->d2 > y->d2) return 1;
    return (x->idx > y->idx) - (x->idx < y->idx);
}

/* d2row[n] -> order[n] (gallery indices, best first) */
ORC_API void orc_rank_row(const float *d2row, int64_t n, int64_t *order) {
    orc_pair *p = (orc_pair *)malloc(sizeof(orc_pair) * (size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        p[i].d2 = d2row[i];
        p[i].idx = i;
    }
    qsort(p, (size_t)n, sizeof(orc_pair), orc_pair_cmp);
    for (int64_t i = 0; i < n; ++i) order[i] = p[i].idx;
    free(p);
}

/* faiss-shaped search: D[nq*k] ascending squared L2, I[nq*k]. */
ORC_API void orc_search(const float *q, const float *g, int64_t nq, int64_t n, int d,
                        int64_t k, float *D, int64_t *I) {
#pragma omp parallel
    {
        orc_pair *p = (orc_pair *)malloc(sizeof(orc_pair) * (size_t)n);
#pragma omp for schedule(dynamic, 1)
        for (int64_t a = 0; a < nq; ++a) {
            for (int64_t b = 0; b < n; ++b) {
                p[b].d2 = orc_d2(q + a * d, g + b * d, d);
                p[b].idx = b;
            }
            qsort(p, (size_t)n, sizeof(orc_pair), orc_pair_cmp);
            for (int64_t j = 0; j < k && j < n; ++j) {
                D[a * k + j] = p[j].d2;
                I[a * k + j] = p[j].idx;
            }
        }
        free(p);
    }
}

/* ------------------------------------------------------------------------------------
 * sklearn.metrics.average_precision_score (binary), as the reference calls it at
 * dataloader_baseline.py:391 and :461.  Published algorithm (scikit-learn, unpinned by the
 * reference; 1.9.0 in the authoring container): sort by score descending, thresholds at
 * distinct score values, precision = tps/(tps+fps), recall = tps/tps[-1],
 * AP = sum_g (R_g - R_{g-1}) * P_g.  Float64 throughout.
 * Input must already be in descending-score order (ties adjacent).
 * ---------------------------------------------------------------------------------- */
ORC_API double orc_ap_sorted(const int32_t *y_true, const double *y_score, int64_t m) {
    double P = 0.0;
    for (int64_t i = 0; i < m; ++i) P += (double)(y_true[i] != 0);
    if (P == 0.0 || m == 0) return 0.0; /* sklearn warns and returns 0 (never reached: y_true[-1]=1) */
    double ap = 0.0, tps = 0.0, prev_recall = 0.0;
    for (int64_t i = 0; i < m; ++i) {
        tps += (double)(y_true[i] != 0);
        if (i + 1 < m && y_score[i + 1] == y_score[i]) continue; /* inside a tie group */
        double recall = tps / P;
        double precision = tps / (double)(i + 1);
        ap += (recall - prev_recall) * precision;
        prev_recall = recall;
    }
    return ap;
}

typedef struct {
    double s;
    int64_t pos;
    int32_t y;
} orc_sy;

static int orc_sy_cmp(const void *a, const void *b) {
    const orc_sy *x = (const orc_sy *)a, *y = (const orc_sy *)b;
    if (x->s > y->s) return -1;
    if (x->s < y->s) return 1;
    return (x->pos > y->pos) - (x->pos < y->pos);
}

/* Unsorted input: sorts by score descending first (what sklearn does internally). */
ORC_API double orc_ap(const int32_t *y_true, const double *y_score, int64_t m) {
    orc_sy *p = (orc_sy *)malloc(sizeof(orc_sy) * (size_t)(m > 0 ? m : 1));
    for (int64_t i = 0; i < m; ++i) {
        p[i].s = y_score[i];
        p[i].pos = i;
        p[i].y = y_true[i];
    }
    qsort(p, (size_t)m, sizeof(orc_sy), orc_sy_cmp);
    int32_t *yt = (int32_t *)malloc(sizeof(int32_t) * (size_t)(m > 0 ? m : 1));
    double *ys = (double *)malloc(sizeof(double) * (size_t)(m > 0 ? m : 1));
    for (int64_t i = 0; i < m; ++i) {
        yt[i] = p[i].y;
        ys[i] = p[i].s;
    }
    double ap = orc_ap_sorted(yt, ys, m);
    free(p);
    free(yt);
    free(ys);
    return ap;
}

/* ------------------------------------------------------------------------------------
 * One query of evaluation_metric.add2dict / multiprocess_calculate
 * (dataloader_baseline.py:431-435,461,475-481 / :386-402) on an already ranked hit list:
 *   drop ignored; y_true = tp with the LAST survivor forced to 1 (AP only);
 *   AP = sklearn on (y_true, score); R@n = sum_{i<n} tp_i / (sum tp + 1e-10), n in r_at_n.
 * tp/ignore/score are in list order (the order the reference list has when it reaches
 * add2dict: ranked order for trimmed/clip, NMS-grouped order for moments).
 * ---------------------------------------------------------------------------------- */
ORC_API void orc_eval_hit_list(const int32_t *tp, const int32_t *ignore, const double *score,
                               int64_t n, const int32_t *r_at_n, int n_r, double *ap_out,
                               double *recall_out) {
    int32_t *yt = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    double *ys = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
    int64_t m = 0;
    for (int64_t i = 0; i < n; ++i)
        if (!ignore[i]) {
            yt[m] = tp[i];
            ys[m] = score[i];
            ++m;
        }
    double npos = 0.0;
    for (int64_t i = 0; i < m; ++i) npos += (double)yt[i];
    npos += 1e-10;
    for (int r = 0; r < n_r; ++r) {
        double c = 0.0;
        for (int64_t i = 0; i < m && i < r_at_n[r]; ++i) c += (double)yt[i];
        recall_out[r] = c * 1.0 / npos;
    }
    if (m > 0) yt[m - 1] = 1;
    *ap_out = orc_ap(yt, ys, m);
    free(yt);
    free(ys);
}

/* ------------------------------------------------------------------------------------
 * Trimmed / clip retrieval, all queries (ARV_Retrieval.ranking,
 * dataloader_baseline.py:1526-1610; ARV_Retrieval_Clip.ranking :899-953):
 *   tp_j = (gallery_label[j] == query_label)                         (:1553 / :916)
 *   ignore_j = gallery_video[j] in ignore_vids[q, 0..n_ign)          (:1548 / :922)
 *   rank by canonical d2 ascending, ties lowest index; score = -d2   (:1592-1604)
 * Outputs: ap[nq], recall[nq*3] (R@30/50/100), optional top-k survivors' gallery indices.
 * ---------------------------------------------------------------------------------- */
ORC_API void orc_eval_trimmed(const float *q, const float *g, int64_t nq, int64_t n, int d,
                              const int32_t *g_label, const int32_t *g_video,
                              const int32_t *q_label, const int32_t *q_ignore_vids, int n_ign,
                              double *ap, double *recall, int64_t topk, int64_t *top_idx,
                              float *top_d2) {
    static const int32_t r_at_n[3] = {30, 50, 100}; /* dataloader_baseline.py:332 */
#pragma omp parallel
    {
        orc_pair *p = (orc_pair *)malloc(sizeof(orc_pair) * (size_t)n);
        int32_t *tp = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
        int32_t *ig = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
        double *sc = (double *)malloc(sizeof(double) * (size_t)n);
#pragma omp for schedule(dynamic, 1)
        for (int64_t a = 0; a < nq; ++a) {
            for (int64_t b = 0; b < n; ++b) {
                p[b].d2 = orc_d2(q + a * d, g + b * d, d);
                p[b].idx = b;
            }
            qsort(p, (size_t)n, sizeof(orc_pair), orc_pair_cmp);
            int64_t nt = 0;
            for (int64_t r = 0; r < n; ++r) {
                int64_t j = p[r].idx;
                tp[r] = (g_label[j] == q_label[a]);
                int isig = 0;
                for (int t = 0; t < n_ign; ++t)
                    isig |= (g_video[j] == q_ignore_vids[a * n_ign + t]);
                ig[r] = isig;
                sc[r] = -(double)p[r].d2;
                if (!isig && nt < topk) {
                    if (top_idx) top_idx[a * topk + nt] = j;
                    if (top_d2) top_d2[a * topk + nt] = p[r].d2;
                    ++nt;
                }
            }
            for (; nt < topk; ++nt) {
                if (top_idx) top_idx[a * topk + nt] = -1;
                if (top_d2) top_d2[a * topk + nt] = INFINITY;
            }
            orc_eval_hit_list(tp, ig, sc, n, r_at_n, 3, ap + a, recall + a * 3);
        }
        free(p);
        free(tp);
        free(ig);
        free(sc);
    }
}

/* ------------------------------------------------------------------------------------
 * nms_cpu, misc_utils/utils_models.py:153-174.  dets rows = (x1, x2, score), float64 as
 * numpy computes; inclusive length x2-x1+1 (:158); candidates visited in descending score
 * (:159 `argsort()[::-1]`; numpy's order among exactly equal scores is implementation
 * defined -- stated tie-break here: lower row index first); a later candidate survives a
 * kept one iff inter/(len_i+len_j-inter) < thresh (:169-172).
 * keep[] receives row indices in visiting order; returns their count.
 * ---------------------------------------------------------------------------------- */
typedef struct {
    double s;
    int64_t i;
} orc_si;

static int orc_si_cmp(const void *a, const void *b) {
    const orc_si *x = (const orc_si *)a, *y = (const orc_si *)b;
    if (x->s > y->s) return -1;
    if (x->s < y->s) return 1;
    return (x->i > y->i) - (x->i < y->i);
}

ORC_API int64_t orc_nms1d(const double *dets, int64_t n, double thresh, int64_t *keep) {
    orc_si *o = (orc_si *)malloc(sizeof(orc_si) * (size_t)(n > 0 ? n : 1));
    char *dead = (char *)calloc((size_t)(n > 0 ? n : 1), 1);
    for (int64_t i = 0; i < n; ++i) {
        o[i].s = dets[i * 3 + 2];
        o[i].i = i;
    }
    qsort(o, (size_t)n, sizeof(orc_si), orc_si_cmp);
    int64_t nk = 0;
    for (int64_t a = 0; a < n; ++a) {
        if (dead[a]) continue;
        int64_t i = o[a].i;
        keep[nk++] = i;
        double x1i = dets[i * 3], x2i = dets[i * 3 + 1], li = x2i - x1i + 1;
        for (int64_t b = a + 1; b < n; ++b) {
            if (dead[b]) continue;
            int64_t j = o[b].i;
            double x1j = dets[j * 3], x2j = dets[j * 3 + 1], lj = x2j - x1j + 1;
            double xx1 = x1i > x1j ? x1i : x1j;
            double xx2 = x2i < x2j ? x2i : x2j;
            double inter = xx2 - xx1 + 1;
            if (inter < 0.0) inter = 0.0;
            double ovr = inter / (li + lj - inter);
            if (!(ovr < thresh)) dead[b] = 1;
        }
    }
    free(o);
    free(dead);
    return nk;
}

/* ------------------------------------------------------------------------------------
 * Moment retrieval, one query (ARV_Retrieval_Moment.ranking `work`,
 * dataloader_baseline.py:1225-1330):
 *   iou_j = g_iou[j] if g_hit_label[j] == query_label else 0          (:1249-1255)
 *   ignore_j = g_video[j] in ignore_vids                               (:1257)
 *   rank all moments by canonical d2 (:1268-1281); do_nms (:1283-1312): cluster by video in
 *   first-seen ranked order, nms_cpu(thresh) per cluster on (start,end,score), kept items
 *   keep their ranked order inside the cluster, clusters concatenated in first-seen order;
 *   tp = iou >= 0.5 (:1315-1319); multiprocess_calculate (:386-402) on that grouped list.
 * order[n] = ranked gallery indices, d2[n] = their distances (rank order).
 * kept_out (optional, size n) receives the grouped, NMS-kept gallery indices; *n_kept.
 * ---------------------------------------------------------------------------------- */
ORC_API void orc_eval_moment_ranked(const int64_t *order, const float *d2, int64_t n,
                                    const int32_t *g_video, const float *g_start,
                                    const float *g_end, const float *g_iou,
                                    const int32_t *g_hit_label, int32_t q_label,
                                    const int32_t *ignore_vids, int n_ign, double nms_thresh,
                                    double *ap_out, double *recall_out, int64_t *kept_out,
                                    int64_t *n_kept) {
    static const int32_t r_at_n[3] = {30, 50, 100};
    /* cluster ids in first-seen order */
    int32_t maxv = 0;
    for (int64_t i = 0; i < n; ++i)
        if (g_video[i] > maxv) maxv = g_video[i];
    int64_t *cl_of_video = (int64_t *)malloc(sizeof(int64_t) * (size_t)(maxv + 1));
    for (int32_t v = 0; v <= maxv; ++v) cl_of_video[v] = -1;
    int64_t *cl_count = (int64_t *)calloc((size_t)(n > 0 ? n : 1), sizeof(int64_t));
    int64_t *cl_id = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
    int64_t ncl = 0;
    for (int64_t r = 0; r < n; ++r) {
        int32_t v = g_video[order[r]];
        if (cl_of_video[v] < 0) cl_of_video[v] = ncl++;
        cl_id[r] = cl_of_video[v];
        cl_count[cl_id[r]]++;
    }
    int64_t *cl_start = (int64_t *)malloc(sizeof(int64_t) * (size_t)(ncl + 1));
    cl_start[0] = 0;
    for (int64_t c = 0; c < ncl; ++c) cl_start[c + 1] = cl_start[c] + cl_count[c];
    int64_t *fill = (int64_t *)calloc((size_t)(ncl > 0 ? ncl : 1), sizeof(int64_t));
    int64_t *member = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1)); /* rank positions */
    for (int64_t r = 0; r < n; ++r) member[cl_start[cl_id[r]] + fill[cl_id[r]]++] = r;

    int32_t *tp = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    int32_t *ig = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    double *sc = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
    int64_t m = 0;
    for (int64_t c = 0; c < ncl; ++c) {
        int64_t cn = cl_count[c];
        double *dets = (double *)malloc(sizeof(double) * 3 * (size_t)cn);
        int64_t *keep = (int64_t *)malloc(sizeof(int64_t) * (size_t)cn);
        char *is_kept = (char *)calloc((size_t)cn, 1);
        for (int64_t t = 0; t < cn; ++t) {
            int64_t r = member[cl_start[c] + t];
            int64_t j = order[r];
            dets[t * 3] = (double)g_start[j];
            dets[t * 3 + 1] = (double)g_end[j];
            dets[t * 3 + 2] = -(double)d2[r];
        }
        int64_t nk = orc_nms1d(dets, cn, nms_thresh, keep);
        for (int64_t t = 0; t < nk; ++t) is_kept[keep[t]] = 1;
        for (int64_t t = 0; t < cn; ++t) {
            if (!is_kept[t]) continue;
            int64_t r = member[cl_start[c] + t];
            int64_t j = order[r];
            float iou = (g_hit_label[j] == q_label) ? g_iou[j] : 0.0f;
            tp[m] = (iou >= 0.5f);
            int isig = 0;
            for (int u = 0; u < n_ign; ++u) isig |= (g_video[j] == ignore_vids[u]);
            ig[m] = isig;
            sc[m] = -(double)d2[r];
            if (kept_out) kept_out[m] = j;
            ++m;
        }
        free(dets);
        free(keep);
        free(is_kept);
    }
    if (n_kept) *n_kept = m;
    orc_eval_hit_list(tp, ig, sc, m, r_at_n, 3, ap_out, recall_out);
    free(cl_of_video);
    free(cl_count);
    free(cl_id);
    free(cl_start);
    free(fill);
    free(member);
    free(tp);
    free(ig);
    free(sc);
}

/* All queries: rank + the above.  ap[nq], recall[nq*3]. */
ORC_API void orc_eval_moment(const float *q, const float *g, int64_t nq, int64_t n, int d,
                             const int32_t *g_video, const float *g_start, const float *g_end,
                             const float *g_iou, const int32_t *g_hit_label,
                             const int32_t *q_label, const int32_t *q_ignore_vids, int n_ign,
                             double nms_thresh, double *ap, double *recall) {
#pragma omp parallel
    {
        orc_pair *p = (orc_pair *)malloc(sizeof(orc_pair) * (size_t)n);
        int64_t *order = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
        float *d2 = (float *)malloc(sizeof(float) * (size_t)n);
#pragma omp for schedule(dynamic, 1)
        for (int64_t a = 0; a < nq; ++a) {
            for (int64_t b = 0; b < n; ++b) {
                p[b].d2 = orc_d2(q + a * d, g + b * d, d);
                p[b].idx = b;
            }
            qsort(p, (size_t)n, sizeof(orc_pair), orc_pair_cmp);
            for (int64_t r = 0; r < n; ++r) {
                order[r] = p[r].idx;
                d2[r] = p[r].d2;
            }
            orc_eval_moment_ranked(order, d2, n, g_video, g_start, g_end, g_iou, g_hit_label,
                                   q_label[a], q_ignore_vids + a * n_ign, n_ign, nms_thresh,
                                   ap + a, recall + a * 3, NULL, NULL);
        }
        free(p);
        free(order);
        free(d2);
    }
}

/* ------------------------------------------------------------------------------------
 * Segment max-pool + top-k videos (BASELINE.json configs[2]; the implicit per-video max of
 * do_nms's first-seen grouping, dataloader_baseline.py:1285-1311): for each query, each
 * video's best moment (min canonical d2, ties lowest row), videos ranked by it (ties lowest
 * video id).  seg_start[v..v+1] delimits video v's rows.  Outputs [nq*k].
 * ---------------------------------------------------------------------------------- */
ORC_API void orc_segmax_topk(const float *q, const float *g, int64_t nq, int64_t n, int d,
                             const int64_t *seg_start, int64_t nseg, int64_t k,
                             int64_t *top_video, int64_t *top_row, float *top_d2) {
    (void)n;
#pragma omp parallel
    {
        orc_pair *p = (orc_pair *)malloc(sizeof(orc_pair) * (size_t)nseg);
        int64_t *arg = (int64_t *)malloc(sizeof(int64_t) * (size_t)nseg);
#pragma omp for schedule(dynamic, 1)
        for (int64_t a = 0; a < nq; ++a) {
            for (int64_t v = 0; v < nseg; ++v) {
                float best = INFINITY;
                int64_t bi = -1;
                for (int64_t r = seg_start[v]; r < seg_start[v + 1]; ++r) {
                    float x = orc_d2(q + a * d, g + r * d, d);
                    if (x < best || bi < 0) {
                        best = x;
                        bi = r;
                    }
                }
                p[v].d2 = best;
                p[v].idx = v;
                arg[v] = bi;
            }
            qsort(p, (size_t)nseg, sizeof(orc_pair), orc_pair_cmp);
            for (int64_t j = 0; j < k; ++j) {
                if (j < nseg) {
                    top_video[a * k + j] = p[j].idx;
                    top_row[a * k + j] = arg[p[j].idx];
                    top_d2[a * k + j] = p[j].d2;
                } else {
                    top_video[a * k + j] = -1;
                    top_row[a * k + j] = -1;
                    top_d2[a * k + j] = INFINITY;
                }
            }
        }
        free(p);
        free(arg);
    }
}

/* ------------------------------------------------------------------------------------
 * Sampler.pdist, misc_utils/metric_learning_loss.py:276-289:
 *   prod = A A^T; norm = diag(prod); res = (norm_i + norm_j - 2*prod_ij).clamp(min=0);
 *   return res.clamp(min=eps).sqrt()
 * prod is evaluated with the canonical BLOCKED chain below (torch.mm's summation order is a
 * BLAS detail the reference does not pin); the remaining fp32 ops are in the source order.
 *
 * Canonical blocked chain (all dot products and squared distances of the loss path): k runs
 * in blocks of 32; inside a block, chunk c (4 consecutive k) is a sequential fmaf chain t_c
 * starting from 0; the 8 chunk sums are added as a fixed tree (t_c + t_{c^4}, then ^2, then
 * ^1 -- a warp-shuffle butterfly over 8 lanes computes exactly this); the block sum is added
 * to the running accumulator.  Zero padding of the last block does not change any bit.
 * This is the order the CUDA kernels use (vq_triplet*.cu): 8 lanes x LDS.128 per row block.
 * ---------------------------------------------------------------------------------- */
static inline float orc_tree8(const float *t) {
    float u0 = t[0] + t[4], u1 = t[1] + t[5], u2 = t[2] + t[6], u3 = t[3] + t[7];
    float w0 = u0 + u2, w1 = u1 + u3;
    return w0 + w1;
}

static inline float orc_dot(const float *a, const float *b, int d) {
    float acc = 0.0f;
    for (int k0 = 0; k0 < d; k0 += 32) {
        float t[8];
        for (int c = 0; c < 8; ++c) {
            t[c] = 0.0f;
            for (int i = 0; i < 4; ++i) {
                int k = k0 + c * 4 + i;
                if (k < d) t[c] = fmaf(a[k], b[k], t[c]);
            }
        }
        acc = acc + orc_tree8(t);
    }
    return acc;
}

/* squared L2 with the same blocked chain (loss path only; the search path uses orc_d2) */
static inline float orc_d2_blocked(const float *a, const float *b, int d) {
    float acc = 0.0f;
    for (int k0 = 0; k0 < d; k0 += 32) {
        float t[8];
        for (int c = 0; c < 8; ++c) {
            t[c] = 0.0f;
            for (int i = 0; i < 4; ++i) {
                int k = k0 + c * 4 + i;
                if (k < d) {
                    float df = a[k] - b[k];
                    t[c] = fmaf(df, df, t[c]);
                }
            }
        }
        acc = acc + orc_tree8(t);
    }
    return acc;
}

ORC_API void orc_pdist(const float *A, int64_t B, int d, float eps, float *out) {
    float *nrm = (float *)malloc(sizeof(float) * (size_t)B);
    for (int64_t i = 0; i < B; ++i) nrm[i] = orc_dot(A + i * d, A + i * d, d);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < B; ++i)
        for (int64_t j = 0; j < B; ++j) {
            float prod = orc_dot(A + i * d, A + j * d, d);
            float res = (nrm[i] + nrm[j]) - 2.0f * prod;
            if (res < 0.0f) res = 0.0f;
            if (res < eps) res = eps;
            out[i * B + j] = sqrtf(res);
        }
    free(nrm);
}

/* ------------------------------------------------------------------------------------
 * Batch-hard mining -- an EXTENSION (the reference's Sampler has random/semihard/distance/
 * npair only, metric_learning_loss.py:104-111), built from the reference's pieces: masks as
 * semihardsampling :173-176 (pos = same label minus the anchor, neg = different label),
 * p = argmax_{pos} dist, n = argmin_{neg} dist, ties -> lowest index.  Anchors without a
 * positive or without a negative are skipped.  trip[T*3] = (a,p,n); returns T.
 * ---------------------------------------------------------------------------------- */
ORC_API int64_t orc_batchhard_mine(const float *dist, const int64_t *labels, int64_t B,
                                   int32_t *trip) {
    int64_t T = 0;
    for (int64_t i = 0; i < B; ++i) {
        int64_t p = -1, n = -1;
        for (int64_t j = 0; j < B; ++j) {
            if (j == i) continue;
            float x = dist[i * B + j];
            if (labels[j] == labels[i]) {
                if (p < 0 || x > dist[i * B + p]) p = j;
            } else {
                if (n < 0 || x < dist[i * B + n]) n = j;
            }
        }
        if (p < 0 || n < 0) continue;
        trip[T * 3] = (int32_t)i;
        trip[T * 3 + 1] = (int32_t)p;
        trip[T * 3 + 2] = (int32_t)n;
        ++T;
    }
    return T;
}

/* ------------------------------------------------------------------------------------
 * TripletLoss.triplet_distance + forward, metric_learning_loss.py:342-379:
 *   l_t = relu(sum (a-p)^2 - sum (a-n)^2 + margin)   (squared L2 recomputed from the rows)
 *   loss = mean_t l_t
 * and its autograd backward: for active t (l_t > 0)
 *   d/da += 2(n-p)/T, d/dp += -2(a-p)/T, d/dn += 2(a-n)/T.
 * Canonical arithmetic: both sums by the blocked chain (orc_d2_blocked); the mean as a sequential
 * fp32 sum in triplet order divided by T; each gradient row accumulated by scanning the
 * triplets in order with acc = fmaf(2/T, diff, acc) (anchor, then positive, then negative
 * role of the same triplet).  active[T] optional.
 * ---------------------------------------------------------------------------------- */
ORC_API void orc_triplet_fwd_bwd(const float *A, int64_t B, int d, const int32_t *trip,
                                 int64_t T, float margin, float *loss_out, float *grad,
                                 float *per_triplet, int32_t *active) {
    float *lt = (float *)malloc(sizeof(float) * (size_t)(T > 0 ? T : 1));
    float sum = 0.0f;
    for (int64_t t = 0; t < T; ++t) {
        const float *a = A + (int64_t)trip[t * 3] * d;
        const float *p = A + (int64_t)trip[t * 3 + 1] * d;
        const float *n = A + (int64_t)trip[t * 3 + 2] * d;
        float dap = orc_d2_blocked(a, p, d), dan = orc_d2_blocked(a, n, d);
        float v = (dap - dan) + margin;
        lt[t] = v > 0.0f ? v : 0.0f;
        sum += lt[t];
        if (per_triplet) per_triplet[t] = lt[t];
        if (active) active[t] = v > 0.0f;
    }
    *loss_out = T > 0 ? sum / (float)T : 0.0f;
    if (grad) {
        float scale = T > 0 ? 2.0f / (float)T : 0.0f;
#pragma omp parallel for schedule(static)
        for (int64_t r = 0; r < B; ++r) {
            float *gr = grad + r * d;
            for (int c = 0; c < d; ++c) gr[c] = 0.0f;
            for (int64_t t = 0; t < T; ++t) {
                if (!(lt[t] > 0.0f)) continue;
                int32_t ia = trip[t * 3], ip = trip[t * 3 + 1], in = trip[t * 3 + 2];
                const float *a = A + (int64_t)ia * d, *p = A + (int64_t)ip * d,
                            *n = A + (int64_t)in * d;
                if (ia == r)
                    for (int c = 0; c < d; ++c) gr[c] = fmaf(scale, n[c] - p[c], gr[c]);
                if (ip == r)
                    for (int c = 0; c < d; ++c) gr[c] = fmaf(scale, p[c] - a[c], gr[c]);
                if (in == r)
                    for (int c = 0; c < d; ++c) gr[c] = fmaf(scale, a[c] - n[c], gr[c]);
            }
        }
    }
    free(lt);
}

/* ------------------------------------------------------------------------------------
 * Pre-match pooling: feat_func main.py:223-233 (F.normalize(p=2, dim=1) per time step,
 * eps 1e-12) followed by the temporal mean dataloader_baseline.py:1481 (np.mean(axis=-1)).
 * x[B, C, T] -> out[B, C].  Canonical arithmetic: norm^2 by sequential fmaf over c,
 * nrm = max(sqrtf(norm2), 1e-12), v = x / nrm (IEEE division), mean = (sequential sum over
 * t) / T.
 * ---------------------------------------------------------------------------------- */
ORC_API void orc_pool_normalize(const float *x, int64_t B, int C, int T, float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < B; ++b) {
        const float *xb = x + b * (int64_t)C * T;
        float *nr = (float *)malloc(sizeof(float) * (size_t)T);
        for (int t = 0; t < T; ++t) {
            float acc = 0.0f;
            for (int c = 0; c < C; ++c) acc = fmaf(xb[(int64_t)c * T + t], xb[(int64_t)c * T + t], acc);
            float nrm = sqrtf(acc);
            nr[t] = nrm > 1e-12f ? nrm : 1e-12f;
        }
        for (int c = 0; c < C; ++c) {
            float acc = 0.0f;
            for (int t = 0; t < T; ++t) acc += xb[(int64_t)c * T + t] / nr[t];
            out[b * C + c] = acc / (float)T;
        }
        free(nr);
    }
}

/* k-way merge of per-shard top-k lists (no reference counterpart):
 * lists[G][nq][k] of (d2, global idx) ascending -> out [nq][k], order (d2 asc, idx asc). */
ORC_API void orc_topk_merge(const float *d2, const int64_t *idx, int G, int64_t nq, int64_t k,
                            float *out_d2, int64_t *out_idx) {
    for (int64_t a = 0; a < nq; ++a) {
        orc_pair *p = (orc_pair *)malloc(sizeof(orc_pair) * (size_t)(G * k));
        int64_t m = 0;
        for (int g = 0; g < G; ++g)
            for (int64_t j = 0; j < k; ++j) {
                int64_t o = ((int64_t)g * nq + a) * k + j;
                if (idx[o] < 0) continue;
                p[m].d2 = d2[o];
                p[m].idx = idx[o];
                ++m;
            }
        qsort(p, (size_t)m, sizeof(orc_pair), orc_pair_cmp);
        for (int64_t j = 0; j < k; ++j) {
            out_d2[a * k + j] = j < m ? p[j].d2 : INFINITY;
            out_idx[a * k + j] = j < m ? p[j].idx : -1;
        }
        free(p);
    }
}
