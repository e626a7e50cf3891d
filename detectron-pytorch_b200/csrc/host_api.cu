// host_api.cu — HOST-buffer entry points (the reference-facing, copy-inclusive boundary that
// bench.py's end-to-end leg times).  Scratch memory is cached per device and grown on demand.
#include <mutex>
#include "common.cuh"

namespace detbox {
int host_scratch(int dev, size_t bytes, void **out);
std::mutex &host_mutex();
}
using namespace detbox;

// libs/layers/anchor.py:17 encode() for one image: host arrays in, host arrays out.
extern "C" int detbox_anchor_target_host(const float *anchors, long A, const float *gt, int G,
                                         const detbox_target_cfg *cfg, float *labels, float *label_weights,
                                         float *bbox_targets, float *bbox_inside_weights, int device_id)
{
    if (A < 0 || G < 0 || !cfg) return DETBOX_ERR_INVALID;
    if (A == 0) return DETBOX_OK;
    if (!anchors || !labels || !label_weights || !bbox_targets || !bbox_inside_weights || (G > 0 && !gt)) return DETBOX_ERR_INVALID;
    std::lock_guard<std::mutex> lock(host_mutex());
    int cur = 0;
    DETBOX_CUDA_TRY(cudaGetDevice(&cur));
    if (cur != device_id) DETBOX_CUDA_TRY(cudaSetDevice(device_id));
    const size_t b_anc = align_up((size_t)A * 16), b_gt = align_up((size_t)(G > 0 ? G : 1) * 20), b_off = 256;
    const size_t b_a1 = align_up((size_t)A * 4), b_a4 = align_up((size_t)A * 16);
    const size_t b_ws = detbox_anchor_target_workspace_bytes(A, 1, G);
    void *base = nullptr;
    int rc = host_scratch(device_id, b_anc + b_gt + b_off + 2 * b_a1 + 2 * b_a4 + b_ws, &base);
    if (rc == DETBOX_OK) {
        char *p = (char *)base;
        float *d_anc = (float *)p; p += b_anc;
        float *d_gt = (float *)p; p += b_gt;
        int32_t *d_off = (int32_t *)p; p += b_off;
        float *d_lab = (float *)p; p += b_a1;
        float *d_lw = (float *)p; p += b_a1;
        float *d_bt = (float *)p; p += b_a4;
        float *d_bw = (float *)p; p += b_a4;
        void *ws = p;
        cudaStream_t st = 0;
        auto body = [&]() -> int {
            const int32_t off[2] = {0, G};
            DETBOX_CUDA_TRY(cudaMemcpyAsync(d_anc, anchors, (size_t)A * 16, cudaMemcpyHostToDevice, st));
            if (G > 0) DETBOX_CUDA_TRY(cudaMemcpyAsync(d_gt, gt, (size_t)G * 20, cudaMemcpyHostToDevice, st));
            DETBOX_CUDA_TRY(cudaMemcpyAsync(d_off, off, sizeof(off), cudaMemcpyHostToDevice, st));
            int r = detbox_anchor_target(d_anc, A, d_gt, d_off, 1, G, cfg, d_lab, d_lw, d_bt, d_bw, nullptr, ws, b_ws, st);
            if (r != DETBOX_OK) return r;
            DETBOX_CUDA_TRY(cudaMemcpyAsync(labels, d_lab, (size_t)A * 4, cudaMemcpyDeviceToHost, st));
            DETBOX_CUDA_TRY(cudaMemcpyAsync(label_weights, d_lw, (size_t)A * 4, cudaMemcpyDeviceToHost, st));
            DETBOX_CUDA_TRY(cudaMemcpyAsync(bbox_targets, d_bt, (size_t)A * 16, cudaMemcpyDeviceToHost, st));
            DETBOX_CUDA_TRY(cudaMemcpyAsync(bbox_inside_weights, d_bw, (size_t)A * 16, cudaMemcpyDeviceToHost, st));
            DETBOX_CUDA_TRY(cudaStreamSynchronize(st));
            return DETBOX_OK;
        };
        rc = body();
    }
    if (cur != device_id) cudaSetDevice(cur);
    return rc;
}

// Device-side address of PINNED host memory (cudaHostAlloc / cudaHostRegister): what a kernel dereferences to read
// the buffer over PCIe without a staging copy.  Used for the regression deltas of the top-k gather: only the
// top_n rows per image are ever read, so copying all of rpn_box (16 B x A per image) to the device is wasted link time.
extern "C" int detbox_host_device_pointer(const void *host_ptr, void **dev_ptr_out)
{
    if (!host_ptr || !dev_ptr_out) return DETBOX_ERR_INVALID;
    void *d = nullptr;
    DETBOX_CUDA_TRY(cudaHostGetDevicePointer(&d, const_cast<void *>(host_ptr), 0));
    *dev_ptr_out = d;
    return DETBOX_OK;
}
