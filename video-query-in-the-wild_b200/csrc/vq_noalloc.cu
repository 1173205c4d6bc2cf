// vq_noalloc.cu -- vq_nms1d_segmented_static: the segmented temporal NMS (vq_nms1d_segmented, nms1d_kernel in
// vq_metrics.cu) for callers that must not allocate: code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc.  No counterpart in the reference.
//
// The kernel's only scratch is one "dead" byte per detection.  Here that scratch is a file-scope device array of
// kNmsStaticMax bytes (1 MiB): every GPU that loads the library holds it until the process exits.  The entry point
// calls no cudaMalloc / cudaFree / pool routine and copies nothing: dets, offsets and keep must already be device
// memory, and nothing is synchronised -- the launch is the only stream operation.
//
// The array's NAME is a host-side stand-in in host code, not a device address, so the address handed to the kernel
// comes from cudaGetSymbolAddress, taken once per device with that device current (scratch_of_current_device).
//
// Every call on a device shares the array.  nms1d_kernel zeroes the bytes of its own rows before it reads them, so
// nothing of the previous call survives into the next one; calls that overlap on the GPU (two streams, two host
// threads) race on it and must be ordered by the caller.
#include <mutex>

#include "../../include/vqwild.h"
#include "vq_common.cuh"
#include "vq_kernels.h"

namespace vq {

static constexpr int64_t kNmsStaticMax = VQ_NMS1D_STATIC_MAX;
static constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_nms_dead[kNmsStaticMax];

static std::mutex g_setup_mu;
static uint8_t *g_dead_addr[kMaxDevices] = {};

// Device address of g_nms_dead on the CURRENT device (each device has its own copy of the array).
static int scratch_of_current_device(uint8_t **out) {
    int dev = -1;
    VQ_CUDA(cudaGetDevice(&dev));
    VQ_CHECK(dev >= 0 && dev < kMaxDevices, VQ_ERR_INVALID, "device ordinal %d", dev);
    std::lock_guard<std::mutex> lock(g_setup_mu);
    if (!g_dead_addr[dev]) {
        int major = 0;
        VQ_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        VQ_CHECK(major == 10, VQ_ERR_ARCH, "device %d is sm_%d0; this library is built for sm_100a only", dev, major);
        void *addr = nullptr;
        VQ_CUDA(cudaGetSymbolAddress(&addr, g_nms_dead));
        VQ_CHECK(addr != nullptr, VQ_ERR_CUDA, "cudaGetSymbolAddress returned a null address on device %d", dev);
        g_dead_addr[dev] = (uint8_t *)addr;
    }
    *out = g_dead_addr[dev];
    return VQ_OK;
}

static bool on_device(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

}  // namespace vq

using namespace vq;

extern "C" VQ_API int vq_nms1d_segmented_static(const double *dets, const int64_t *offsets, int64_t nseg, int64_t total,
                                                double thresh, uint8_t *keep, void *stream) {
    VQ_CHECK(nseg >= 0 && total >= 0, VQ_ERR_INVALID, "bad sizes: nseg %lld, total %lld", (long long)nseg, (long long)total);
    VQ_CHECK(total <= kNmsStaticMax, VQ_ERR_CAPACITY,
             "vq_nms1d_segmented_static: %lld detections, the static scratch holds VQ_NMS1D_STATIC_MAX = %lld; "
             "use vq_nms1d_segmented",
             (long long)total, (long long)kNmsStaticMax);
    uint8_t *dead = nullptr;
    VQ_TRY(scratch_of_current_device(&dead));  // first call on a device: set-up (a call with nseg == 0 does only this)
    if (nseg == 0 || total == 0) return VQ_OK;
    VQ_CHECK(dets && offsets && keep, VQ_ERR_INVALID, "NULL argument");
    VQ_CHECK(on_device(dets) && on_device(offsets) && on_device(keep), VQ_ERR_INVALID,
             "vq_nms1d_segmented_static takes device memory only (staging a host buffer would allocate)");
    return launch_nms1d(dets, offsets, nseg, thresh, keep, dead, (cudaStream_t)stream);
}
