// vq_api.cu -- the C ABI (include/vqwild.h): argument checking, host<->device staging, the
// index object and the orchestration of the kernels.  Host logic only; every number a caller
// sees is produced by the kernels in this directory.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "vq_common.cuh"
#include "vq_kernels.h"

namespace vq {

DebugHooks g_hooks;
static int *hook_slot(const char *name) {
    struct { const char *n; int *p; } tab[] = {
        {"FORCE_FALLBACK", &g_hooks.force_fallback}, {"NO_PAIR", &g_hooks.no_pair}, {"NO_CHAIN", &g_hooks.no_chain},
        {"DENSE_ROWS", &g_hooks.dense_rows}, {"NO_TC_TRIPLET", &g_hooks.no_tc_triplet},
        {"TRIPLET_CLOCKS", &g_hooks.triplet_clocks}, {"NO_TC_EVAL", &g_hooks.no_tc_eval},
        {"NO_INSERT", &g_hooks.no_insert}, {"KPRIME", &g_hooks.kprime}, {"TRIGGER", &g_hooks.trigger},
        {"NO_VMAJOR", &g_hooks.no_vmajor}, {"NO_NMS_MASKS", &g_hooks.no_nms_masks},
        {"NO_SEG_UNIQUE", &g_hooks.no_seg_unique}, {"WALKERS", &g_hooks.walkers},
        {"NO_L2SQ_WIDE", &g_hooks.no_l2sq_wide}};
    for (auto &t : tab)
        if (!strcmp(name, t.n)) return t.p;
    return nullptr;
}
static const bool g_hooks_from_env = [] {
    for (const char *n : {"FORCE_FALLBACK", "NO_PAIR", "NO_CHAIN", "DENSE_ROWS", "NO_TC_TRIPLET", "TRIPLET_CLOCKS", "NO_TC_EVAL", "NO_INSERT", "KPRIME", "TRIGGER", "NO_VMAJOR", "NO_NMS_MASKS", "NO_SEG_UNIQUE", "WALKERS", "NO_L2SQ_WIDE"}) {
        char env[64];
        snprintf(env, sizeof(env), "VQ_DEBUG_%s", n);
        if (const char *e = getenv(env)) *hook_slot(n) = atoi(e);
    }
    return true;
}();

static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static int check_device() {
    int dev = 0, n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        set_error("no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
        return VQ_ERR_NO_DEVICE;
    }
    VQ_CUDA(cudaGetDevice(&dev));
    int major = 0;
    VQ_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major != 10) {
        set_error("device %d is sm_%d0; this library is built for sm_100a only", dev, major);
        return VQ_ERR_ARCH;
    }
    return VQ_OK;
}

static int sm_count() {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}

static bool is_device_ptr(const void *p) {
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// The library's own stream-ordered pool, one per device.  Freed blocks stay in the pool (release threshold =
// max): the default pool hands its memory back to the OS at every synchronisation, which put a fresh
// physical allocation of every staging buffer (~10 ms per 100 MB) on the host's critical path of each call.
static cudaMemPool_t g_pool[64] = {};
static int scratch_pool(cudaMemPool_t *out) {
    int dev = 0;
    VQ_CUDA(cudaGetDevice(&dev));
    VQ_CHECK(dev >= 0 && dev < 64, VQ_ERR_INVALID, "device ordinal %d", dev);
    if (!g_pool[dev]) {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        cudaMemPool_t pool = nullptr;
        VQ_CUDA(cudaMemPoolCreate(&pool, &props));
        uint64_t keep = UINT64_MAX;
        VQ_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
        g_pool[dev] = pool;
    }
    *out = g_pool[dev];
    return VQ_OK;
}

// Stream-ordered temporary device buffer.
struct Tmp {
    void *p = nullptr;
    cudaStream_t st = nullptr;
    int alloc(size_t bytes, cudaStream_t s) {
        st = s;
        cudaMemPool_t pool = nullptr;
        VQ_TRY(scratch_pool(&pool));
        VQ_CUDA(cudaMallocFromPoolAsync(&p, bytes ? bytes : 16, pool, s));
        return VQ_OK;
    }
    ~Tmp() {
        if (p) cudaFreeAsync(p, st);
    }
    Tmp() = default;
    Tmp(const Tmp &) = delete;
    Tmp &operator=(const Tmp &) = delete;
};

// Input argument: device pointers pass through, host pointers are copied to a temporary.
struct In {
    Tmp tmp;
    const void *dev = nullptr;
    int bind(const void *src, size_t bytes, cudaStream_t s) {
        if (!src) {
            dev = nullptr;
            return VQ_OK;
        }
        if (is_device_ptr(src)) {
            dev = src;
            return VQ_OK;
        }
        VQ_TRY(tmp.alloc(bytes, s));
        VQ_CUDA(cudaMemcpyAsync(tmp.p, src, bytes, cudaMemcpyHostToDevice, s));
        dev = tmp.p;
        return VQ_OK;
    }
    template <typename T>
    const T *as() const { return (const T *)dev; }
};

// Output argument: device pointers are written in place, host pointers get a temporary that
// finish() copies back.
struct Out {
    Tmp tmp;
    void *dev = nullptr, *host = nullptr;
    size_t bytes = 0;
    int bind(void *dst, size_t nbytes, cudaStream_t s) {
        bytes = nbytes;
        if (!dst) return VQ_OK;
        if (is_device_ptr(dst)) {
            dev = dst;
            return VQ_OK;
        }
        host = dst;
        VQ_TRY(tmp.alloc(nbytes, s));
        dev = tmp.p;
        return VQ_OK;
    }
    int finish(cudaStream_t s) {
        if (host && bytes) VQ_CUDA(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, s));
        return VQ_OK;
    }
    template <typename T>
    T *as() const { return (T *)dev; }
};

// ---------------------------------------------------------------------------------------
// TMA descriptors (driver entry point fetched at run time: no link-time libcuda dependency)
// ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

int make_tensor_map_2d(CUtensorMap *m, const void *base, CUtensorMapDataType dt, int elem_bytes, uint64_t cols,
                       uint64_t rows, uint64_t pitch_bytes, uint32_t box_cols, uint32_t box_rows) {
    EncodeTiledFn fn = encode_fn();
    VQ_CHECK(fn, VQ_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    VQ_CHECK((int)box_cols * elem_bytes <= 128, VQ_ERR_INVALID, "tensor map box wider than the 128-byte swizzle span");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)(rows > 0 ? rows : 1)};
    cuuint64_t strides[1] = {(cuuint64_t)pitch_bytes};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = fn(m, dt, 2, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    VQ_CHECK(r == CUDA_SUCCESS, VQ_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return VQ_OK;
}

// bf16 plane [rows, dp] row-major (dp = plane pitch incl. the augmentation block) -> boxes of 64
// columns x box_rows rows, 128-byte swizzle; the box at column pitch-16 is zero-filled past the edge
static int make_map(CUtensorMap *m, const void *base, int64_t rows, int dp, int box_rows) {
    return make_tensor_map_2d(m, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (uint64_t)dp, (uint64_t)rows, (uint64_t)dp * 2,
                              64u, (uint32_t)box_rows);
}

}  // namespace vq

using namespace vq;

// ---------------------------------------------------------------------------------------
// index object
// ---------------------------------------------------------------------------------------
struct vq_index {
    int d = 0, dp = 0, dpa = 0, dtype = -1, prec = VQ_PREC_AUTO;  // dpa = dp + kAugCols: operand plane pitch
    int64_t n = 0, cap = 0;
    void *raw = nullptr;            // [cap, d] f32 or bf16: the rows as given (canonical re-scoring)
    __nv_bfloat16 *hi = nullptr;    // [cap, dpa]
    __nv_bfloat16 *lo = nullptr;    // [cap, dpa] (f32 input only)
    float *norm2 = nullptr;         // [cap]
    uint32_t *gmax_bits = nullptr;  // [1]
    int32_t *label = nullptr, *video = nullptr;  // [n]
    int32_t *seg_of_row = nullptr;  // [n]
    int64_t *seg_start_dev = nullptr;
    std::vector<int64_t> seg_start;  // host copy
    // moment evaluation (a8): per-row moment metadata, videos split by size class for the NMS kernel
    float *m_start = nullptr, *m_end = nullptr, *m_iou = nullptr;
    int32_t *m_hit = nullptr;
    unsigned long long *m_sup = nullptr;  // NMS suppression masks (videos of <= 64 moments) for threshold m_sup_thresh
    double m_sup_thresh = -1.0;
    bool tc_eval_off = false;  // the tensor-core form of eval_trimmed handed most queries back last time (labels unrelated to the features)
    int32_t *seg_small = nullptr, *seg_large = nullptr;
    int n_seg_small = 0, n_seg_large = 0;
    int64_t seg_max_len = 0;
    // search workspace (grown on demand)
    uint64_t *cand = nullptr;
    size_t cand_bytes = 0;
    int32_t *cand_cnt = nullptr;
    float *cand_tau = nullptr;
    int32_t *cand_cum = nullptr;
    size_t cand_cnt_cap = 0, cand_tau_cap = 0, cand_cum_cap = 0;
    int64_t *plan_rows_dev = nullptr;  // segment plan of the last search (GemmTopkParams)
    size_t plan_rows_cap = 0;
    int32_t *plan_ints_dev = nullptr;
    size_t plan_ints_cap = 0;
    int32_t *unproven = nullptr;
    size_t unproven_cap = 0;
    int32_t *unproven_count = nullptr;
    int32_t *stat_pinned = nullptr;  // pinned host mirror of unproven_count
    int64_t stats[6] = {0, 0, 0, 0, 0, 0};
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // search phases (vq_index_last_timing)
    bool ev_valid = false;
};

static size_t esize(int dtype) { return dtype == VQ_F32 ? 4 : 2; }

extern "C" {

int vq_version(void) { return 200; }

int vq_debug_set(const char *hook, int value) {
    VQ_CHECK(hook, VQ_ERR_INVALID, "NULL hook name");
    int *p = hook_slot(hook);
    VQ_CHECK(p, VQ_ERR_INVALID, "unknown debug hook %s", hook);
    *p = value;
    return VQ_OK;
}
const char *vq_last_error(void) { return g_err; }

int vq_device_check(int dev) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0 || dev >= n) {
        set_error("no CUDA device %d (%s)", dev, cudaGetErrorString(e));
        return VQ_ERR_NO_DEVICE;
    }
    int major = 0;
    VQ_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    VQ_CHECK(major == 10, VQ_ERR_ARCH, "device %d is sm_%d0, need sm_100", dev, major);
    return VQ_OK;
}

int vq_index_create(int d, vq_index **out) {
    VQ_CHECK(out, VQ_ERR_INVALID, "vq_index_create: out is NULL");
    VQ_CHECK(d >= 1 && d <= 4096, VQ_ERR_INVALID, "vq_index_create: d=%d out of range [1,4096]", d);
    VQ_TRY(check_device());
    vq_index *ix = new vq_index();
    ix->d = d;
    ix->dp = (d + 63) / 64 * 64;
    ix->dpa = ix->dp + kAugCols;
    if (cudaMalloc(&ix->gmax_bits, 4) != cudaSuccess || cudaMalloc(&ix->unproven_count, 4) != cudaSuccess ||
        cudaMallocHost(&ix->stat_pinned, 4) != cudaSuccess) {
        set_error("vq_index_create: allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
        delete ix;
        return VQ_ERR_NOMEM;
    }
    for (int i = 0; i < 5; ++i) cudaEventCreate(&ix->ev[i]);
    cudaMemset(ix->gmax_bits, 0, 4);
    cudaMemset(ix->unproven_count, 0, 4);
    *ix->stat_pinned = 0;
    *out = ix;
    return VQ_OK;
}

int vq_index_destroy(vq_index *ix) {
    if (!ix) return VQ_OK;
    cudaDeviceSynchronize();
    void *ptrs[] = {ix->raw,   ix->hi,         ix->lo,           ix->norm2,    ix->gmax_bits,
                    ix->label, ix->video,      ix->seg_of_row,   ix->seg_start_dev, ix->cand,     ix->cand_cnt,
                    ix->cand_tau, ix->cand_cum, ix->plan_rows_dev, ix->plan_ints_dev, ix->unproven, ix->unproven_count,
                    ix->m_start, ix->m_end, ix->m_iou, ix->m_hit, ix->seg_small, ix->seg_large, ix->m_sup};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    if (ix->stat_pinned) cudaFreeHost(ix->stat_pinned);
    for (int i = 0; i < 5; ++i)
        if (ix->ev[i]) cudaEventDestroy(ix->ev[i]);
    delete ix;
    return VQ_OK;
}

int64_t vq_index_ntotal(const vq_index *ix) { return ix ? ix->n : 0; }
int vq_index_dim(const vq_index *ix) { return ix ? ix->d : 0; }

int vq_index_set_precision(vq_index *ix, int prec) {
    VQ_CHECK(ix, VQ_ERR_INVALID, "index is NULL");
    VQ_CHECK(prec >= VQ_PREC_AUTO && prec <= VQ_PREC_EXACT, VQ_ERR_INVALID, "bad precision %d", prec);
    ix->prec = prec;
    return VQ_OK;
}

}  // extern "C"

template <typename T>
static int grow(T **p, int64_t old_elems, int64_t new_elems, cudaStream_t st) {
    T *np_ = nullptr;
    if (cudaMalloc(&np_, (size_t)new_elems * sizeof(T)) != cudaSuccess) {
        set_error("out of device memory growing a buffer to %lld bytes", (long long)(new_elems * sizeof(T)));
        cudaGetLastError();
        return VQ_ERR_NOMEM;
    }
    if (*p && old_elems) VQ_CUDA(cudaMemcpyAsync(np_, *p, (size_t)old_elems * sizeof(T), cudaMemcpyDeviceToDevice, st));
    if (*p) {
        VQ_CUDA(cudaStreamSynchronize(st));
        cudaFree(*p);
    }
    *p = np_;
    return VQ_OK;
}

extern "C" {

int vq_index_add(vq_index *ix, const void *xb, int64_t n, int dtype, void *stream) {
    VQ_CHECK(ix && (xb || n == 0), VQ_ERR_INVALID, "vq_index_add: NULL argument");
    VQ_CHECK(n >= 0, VQ_ERR_INVALID, "vq_index_add: n=%lld", (long long)n);
    VQ_CHECK(dtype == VQ_F32 || dtype == VQ_BF16, VQ_ERR_INVALID, "vq_index_add: dtype %d", dtype);
    VQ_CHECK(ix->dtype < 0 || ix->dtype == dtype, VQ_ERR_INVALID, "vq_index_add: dtype differs from earlier adds");
    VQ_CHECK(ix->n + n < (1ll << 31), VQ_ERR_CAPACITY, "index limited to 2^31-1 rows");
    if (n == 0) return VQ_OK;
    cudaStream_t st = (cudaStream_t)stream;
    ix->dtype = dtype;
    const int64_t nn = ix->n + n;
    if (nn > ix->cap) {
        const int64_t ncap = std::max<int64_t>(nn, ix->cap + ix->cap / 2);
        const size_t es = esize(dtype);
        char *raw = (char *)ix->raw;
        VQ_TRY(grow<char>(&raw, ix->n * ix->d * es, ncap * ix->d * es, st));
        ix->raw = raw;
        VQ_TRY(grow<__nv_bfloat16>(&ix->hi, ix->n * ix->dpa, ncap * ix->dpa, st));
        if (dtype == VQ_F32) VQ_TRY(grow<__nv_bfloat16>(&ix->lo, ix->n * ix->dpa, ncap * ix->dpa, st));
        VQ_TRY(grow<float>(&ix->norm2, ix->n, ncap, st));
        ix->cap = ncap;
    }
    const size_t es = esize(dtype);
    char *dst = (char *)ix->raw + (size_t)ix->n * ix->d * es;
    VQ_CUDA(cudaMemcpyAsync(dst, xb, (size_t)n * ix->d * es, cudaMemcpyDefault, st));
    VQ_TRY(launch_prepare_rows(dst, dtype, n, ix->d, ix->dp, false, ix->hi + ix->n * ix->dpa,
                               ix->lo ? ix->lo + ix->n * ix->dpa : nullptr, ix->norm2 + ix->n, ix->gmax_bits, st));
    ix->n = nn;
    // metadata no longer covers all rows
    if (ix->label) { cudaFree(ix->label); ix->label = nullptr; }
    if (ix->video) { cudaFree(ix->video); ix->video = nullptr; }
    if (ix->seg_of_row) { cudaFree(ix->seg_of_row); ix->seg_of_row = nullptr; }
    ix->seg_start.clear();
    for (void **pp : {(void **)&ix->m_start, (void **)&ix->m_end, (void **)&ix->m_iou, (void **)&ix->m_hit, (void **)&ix->m_sup})
        if (*pp) { cudaFree(*pp); *pp = nullptr; }
    ix->m_sup_thresh = -1.0;
    return VQ_OK;
}

int vq_index_set_meta(vq_index *ix, const int32_t *label, const int32_t *video, void *stream) {
    VQ_CHECK(ix, VQ_ERR_INVALID, "index is NULL");
    VQ_CHECK(ix->n > 0, VQ_ERR_STATE, "vq_index_set_meta before vq_index_add");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t bytes = (size_t)ix->n * 4;
    if (ix->label) { cudaFree(ix->label); ix->label = nullptr; }
    if (ix->video) { cudaFree(ix->video); ix->video = nullptr; }
    ix->tc_eval_off = false;
    if (label) {
        VQ_CUDA(cudaMalloc(&ix->label, bytes));
        VQ_CUDA(cudaMemcpyAsync(ix->label, label, bytes, cudaMemcpyDefault, st));
    }
    if (video) {
        VQ_CUDA(cudaMalloc(&ix->video, bytes));
        VQ_CUDA(cudaMemcpyAsync(ix->video, video, bytes, cudaMemcpyDefault, st));
    }
    return VQ_OK;
}

int vq_index_set_segments(vq_index *ix, const int64_t *seg_start, int64_t nseg, void *stream) {
    VQ_CHECK(ix && seg_start, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(nseg >= 1, VQ_ERR_INVALID, "nseg=%lld", (long long)nseg);
    VQ_CHECK(!is_device_ptr(seg_start), VQ_ERR_INVALID, "seg_start must be host memory");
    VQ_CHECK(seg_start[0] == 0 && seg_start[nseg] == ix->n, VQ_ERR_INVALID,
             "segments must cover rows [0,%lld)", (long long)ix->n);
    for (int64_t v = 0; v < nseg; ++v)
        VQ_CHECK(seg_start[v] <= seg_start[v + 1], VQ_ERR_INVALID, "seg_start not monotone at %lld", (long long)v);
    cudaStream_t st = (cudaStream_t)stream;
    ix->m_sup_thresh = -1.0;  // the suppression masks follow the segmentation
    ix->seg_start.assign(seg_start, seg_start + nseg + 1);
    std::vector<int32_t> sor((size_t)ix->n);
    for (int64_t v = 0; v < nseg; ++v)
        for (int64_t r = seg_start[v]; r < seg_start[v + 1]; ++r) sor[(size_t)r] = (int32_t)v;
    if (ix->seg_of_row) cudaFree(ix->seg_of_row);
    if (ix->seg_start_dev) cudaFree(ix->seg_start_dev);
    VQ_CUDA(cudaMalloc(&ix->seg_of_row, (size_t)ix->n * 4));
    VQ_CUDA(cudaMalloc(&ix->seg_start_dev, (size_t)(nseg + 1) * 8));
    VQ_CUDA(cudaMemcpyAsync(ix->seg_of_row, sor.data(), (size_t)ix->n * 4, cudaMemcpyHostToDevice, st));
    VQ_CUDA(cudaMemcpyAsync(ix->seg_start_dev, seg_start, (size_t)(nseg + 1) * 8, cudaMemcpyHostToDevice, st));
    // size classes for the per-video NMS (vq_moment.cu)
    std::vector<int32_t> small, large;
    ix->seg_max_len = 0;
    for (int64_t v = 0; v < nseg; ++v) {
        const int64_t len = seg_start[v + 1] - seg_start[v];
        ix->seg_max_len = std::max(ix->seg_max_len, len);
        (len <= kMomentWarpCap ? small : large).push_back((int32_t)v);
    }
    if (ix->seg_small) { cudaFree(ix->seg_small); ix->seg_small = nullptr; }
    if (ix->seg_large) { cudaFree(ix->seg_large); ix->seg_large = nullptr; }
    ix->n_seg_small = (int)small.size();
    ix->n_seg_large = (int)large.size();
    if (!small.empty()) {
        VQ_CUDA(cudaMalloc(&ix->seg_small, small.size() * 4));
        VQ_CUDA(cudaMemcpyAsync(ix->seg_small, small.data(), small.size() * 4, cudaMemcpyHostToDevice, st));
    }
    if (!large.empty()) {
        VQ_CUDA(cudaMalloc(&ix->seg_large, large.size() * 4));
        VQ_CUDA(cudaMemcpyAsync(ix->seg_large, large.data(), large.size() * 4, cudaMemcpyHostToDevice, st));
    }
    VQ_CUDA(cudaStreamSynchronize(st));  // sor / small / large are stack-lifetime staging buffers
    return VQ_OK;
}

int vq_index_set_moments(vq_index *ix, const float *start_sec, const float *end_sec, const float *iou,
                         const int32_t *hit_label, void *stream) {
    VQ_CHECK(ix && start_sec && end_sec && iou && hit_label, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(ix->n > 0, VQ_ERR_STATE, "vq_index_set_moments before vq_index_add");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t bytes = (size_t)ix->n * 4;
    ix->m_sup_thresh = -1.0;
    const void *src[4] = {start_sec, end_sec, iou, hit_label};
    void **dst[4] = {(void **)&ix->m_start, (void **)&ix->m_end, (void **)&ix->m_iou, (void **)&ix->m_hit};
    for (int i = 0; i < 4; ++i) {
        if (*dst[i]) { cudaFree(*dst[i]); *dst[i] = nullptr; }
        VQ_CUDA(cudaMalloc(dst[i], bytes));
        VQ_CUDA(cudaMemcpyAsync(*dst[i], src[i], bytes, cudaMemcpyDefault, st));
    }
    return VQ_OK;
}

int vq_index_last_stats(const vq_index *ix, int64_t *out6) {
    VQ_CHECK(ix && out6, VQ_ERR_INVALID, "NULL argument");
    VQ_CUDA(cudaDeviceSynchronize());
    int32_t c = 0;
    VQ_CUDA(cudaMemcpy(&c, ix->unproven_count, 4, cudaMemcpyDeviceToHost));
    for (int i = 0; i < 6; ++i) out6[i] = ix->stats[i];
    out6[0] = c;
    return VQ_OK;
}

int vq_index_last_timing(const vq_index *ix, float *out4) {
    VQ_CHECK(ix && out4, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(ix->ev_valid, VQ_ERR_STATE, "no tensor-core search has run on this index");
    VQ_CUDA(cudaEventSynchronize(ix->ev[4]));
    for (int i = 0; i < 4; ++i) VQ_CUDA(cudaEventElapsedTime(&out4[i], ix->ev[i], ix->ev[i + 1]));
    return VQ_OK;
}

}  // extern "C"

#include "vq_api_search.inl"
