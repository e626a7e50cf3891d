// box_transform.cu — standalone encode / decode / clip over [n,4] rows (48 B per row of HBM
// traffic: two float4 loads, one float4 store per thread).
#include "box_math.cuh"

using namespace detbox;

__global__ void __launch_bounds__(256)
encode_kernel(const float4 *__restrict__ anc, const float4 *__restrict__ gt, long n, int enc, float4 *__restrict__ out)
{
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
        out[i] = encode_box(anc[i], gt[i], enc);
}

__global__ void __launch_bounds__(256)
decode_kernel(const float4 *__restrict__ anc, const float4 *__restrict__ del, long n, int enc,
              int clip, float mx, float my, float4 *__restrict__ out)
{
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        float4 b = decode_box(anc[i], del[i], enc);
        if (clip) b = clip_box(b, mx, my);
        out[i] = b;
    }
}

__global__ void __launch_bounds__(256)
clip_kernel(float4 *__restrict__ boxes, long n4, float mx, float my)
{
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x)
        boxes[i] = clip_box(boxes[i], mx, my);
}

// diagnostics: rcp_rn_inrange against __frcp_rn over a range of float bit patterns
__global__ void __launch_bounds__(256)
selftest_rcp_kernel(uint32_t lo, uint32_t hi, unsigned long long *__restrict__ mismatches)
{
    unsigned long long bad = 0;
    for (unsigned long long v = (unsigned long long)lo + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; v <= hi;
         v += (unsigned long long)gridDim.x * blockDim.x) {
        const float b = __uint_as_float((uint32_t)v);
        bad += __float_as_uint(rcp_rn_inrange(b)) != __float_as_uint(__frcp_rn(b));
    }
    if (bad) atomicAdd(mismatches, bad);
}

static unsigned grid_for(long n) {
    long b = (n + 255) / 256;
    const long cap = (long)kNumSMs * 16;
    return (unsigned)(b < cap ? (b > 0 ? b : 1) : cap);
}

extern "C" int detbox_box_encode(const float *anc, const float *gt, long n, int enc, float *out, void *stream)
{
    if (n < 0 || enc < 0 || enc > 2) return DETBOX_ERR_INVALID;
    if (n == 0) return DETBOX_OK;
    if (!anc || !gt || !out) return DETBOX_ERR_INVALID;
    encode_kernel<<<grid_for(n), 256, 0, as_stream(stream)>>>((const float4 *)anc, (const float4 *)gt, n, enc, (float4 *)out);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

extern "C" int detbox_box_decode(const float *anc, const float *del, long n, int enc, int im_h, int im_w,
                                 float *out, void *stream)
{
    if (n < 0 || enc < 0 || enc > 2) return DETBOX_ERR_INVALID;
    if (n == 0) return DETBOX_OK;
    if (!anc || !del || !out) return DETBOX_ERR_INVALID;
    decode_kernel<<<grid_for(n), 256, 0, as_stream(stream)>>>((const float4 *)anc, (const float4 *)del, n, enc,
                                                             im_h > 0, (float)(im_w - 1), (float)(im_h - 1), (float4 *)out);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

extern "C" int detbox_clip_boxes(float *boxes, long n, int k, int im_h, int im_w, void *stream)
{
    if (n < 0 || k <= 0) return DETBOX_ERR_INVALID;
    if (n == 0) return DETBOX_OK;
    if (!boxes) return DETBOX_ERR_INVALID;
    clip_kernel<<<grid_for(n * k), 256, 0, as_stream(stream)>>>((float4 *)boxes, n * k, (float)(im_w - 1), (float)(im_h - 1));
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

extern "C" int detbox_selftest_rcp(uint32_t bits_lo, uint32_t bits_hi, unsigned long long *mismatches_dev, void *stream)
{
    if (!mismatches_dev || bits_hi < bits_lo) return DETBOX_ERR_INVALID;
    DETBOX_CUDA_TRY(cudaMemsetAsync(mismatches_dev, 0, sizeof(unsigned long long), as_stream(stream)));
    selftest_rcp_kernel<<<kNumSMs * 16, 256, 0, as_stream(stream)>>>(bits_lo, bits_hi, mismatches_dev);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}
