// no_alloc.cu — detbox_nms_static: greedy NMS whose scratch memory is built into the library.
//
// No counterpart in the reference; added for callers that cannot allocate:
// code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc.  Every other device entry
// point of this library takes its scratch from the caller (sized by *_workspace_bytes()); here the scratch of the NMS
// kernels (sorted boxes, suppressor words, kept / suppressed bitsets, segment offsets — nms.cu) is a file-scope
// __device__ array of a fixed, modest size, so the caller passes inputs and outputs and nothing else.
//
//   * Limit: n <= DETBOX_NMS_STATIC_MAX_BOXES (4096) and n_seg <= DETBOX_NMS_STATIC_MAX_SEGMENTS (64); the array is
//     2.25 MiB (detbox_nms_workspace_bytes(4096, 64) = 2,264,832 bytes).  Anything larger is rejected with
//     DETBOX_ERR_UNSUPPORTED ("size beyond a documented limit") before any work is enqueued.
//   * Every GPU that loads the library holds the array until the process exits.
//   * In HOST code the array's name is a host-side stand-in, not a device address: handing `g_nms_scratch` itself to a
//     kernel would fault (or, where the GPU can reach host memory, silently use the wrong memory).  The device address
//     is looked up once per device with cudaGetSymbolAddress, with that device current, and cached.
//   * All calls share the array and it keeps the previous call's contents.  nms_impl (nms.cu) clears what has to
//     start at zero (segment offsets, bitsets, the lower-triangle words it reads) at the start of every call, exactly
//     as it does for a caller-provided workspace that was used before.  Two calls that OVERLAP on the GPU (different
//     streams, or graph branches that run concurrently) race on the array: serialise them.
//   * This file calls no cudaMalloc / cudaFree / allocating library routine; neither does detbox_nms.  The first call on
//     a device performs the symbol look-up (it may trigger lazy module loading): make it once outside graph capture.
#include <mutex>
#include "common.cuh"

using namespace detbox;

namespace {

constexpr size_t kStaticScratchBytes = 2304 * 1024;          // 2.25 MiB >= detbox_nms_workspace_bytes(4096, 64)
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_nms_scratch[kStaticScratchBytes];

std::mutex g_mu;
void *g_scratch_addr[kMaxDevices] = {};                      // device address of g_nms_scratch on device i (0 = not looked up yet)

// set-up, once per device, with that device current
int scratch_on_current_device(void **out)
{
    int dev = -1;
    DETBOX_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return DETBOX_ERR_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[dev]) {
        void *p = nullptr;
        DETBOX_CUDA_TRY(cudaGetSymbolAddress(&p, g_nms_scratch));
        if (!p) return DETBOX_ERR_CUDA;
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return DETBOX_OK;
}

}  // namespace

extern "C" int detbox_nms_static(const float *boxes, int box_stride, const float *scores, long n,
                                 const int32_t *seg_ids, int n_seg, double thresh, int cmp_mode,
                                 int64_t *keep, int32_t *n_keep, void *stream)
{
    if (n < 0) return DETBOX_ERR_INVALID;
    if (n > DETBOX_NMS_STATIC_MAX_BOXES || n_seg > DETBOX_NMS_STATIC_MAX_SEGMENTS) return DETBOX_ERR_UNSUPPORTED;
    const size_t need = detbox_nms_workspace_bytes(n, n_seg);
    if (need > kStaticScratchBytes) return DETBOX_ERR_UNSUPPORTED;
    void *ws = nullptr;
    const int rc = scratch_on_current_device(&ws);
    if (rc != DETBOX_OK) return rc;
    // the existing entry point does all the work: the only difference is where the scratch lives
    return detbox_nms(boxes, box_stride, scores, n, seg_ids, n_seg, thresh, cmp_mode, keep, n_keep,
                      ws, kStaticScratchBytes, stream);
}
