// anchors.cu — anchor generation.
//   host : base anchors at one cell   (reference libs/boxes/anchor.py:125-189, :101-103)
//   dev  : expansion over the FPN grids (reference libs/boxes/anchor.py:30-70)
// The expansion is a pure streaming write (16 B per anchor as float32): one thread per anchor,
// one float4 store, all levels in a single launch.
#include <math.h>
#include "common.cuh"

using namespace detbox;

extern "C" int detbox_base_anchors_host(const double *scales, int n_scales, const double *ratios,
                                        int n_ratios, long base_size, double *out)
{
    if (!scales || !ratios || !out || n_scales <= 0 || n_ratios <= 0) return DETBOX_ERR_INVALID;
    // window [-b//2, -b//2, b//2, b//2] (anchor.py:132, Python floor division)
    const double lo = floor(-(double)base_size / 2.0), hi = floor((double)base_size / 2.0);
    const double side = hi - lo + 1.0;
    const double ctr = lo + 0.5 * (side - 1.0);
    const double area = side * side;
    for (int r = 0; r < n_ratios; ++r) {
        // anchor.py:171-176 — np.round is round-half-even == rint()
        const double ws = rint(sqrt(area / ratios[r]));
        const double hs = rint(ws * ratios[r]);
        const double x1 = ctr - 0.5 * (ws - 1.0), x2 = ctr + 0.5 * (ws - 1.0);
        const double y1 = ctr - 0.5 * (hs - 1.0), y2 = ctr + 0.5 * (hs - 1.0);
        const double w = x2 - x1 + 1.0, h = y2 - y1 + 1.0;
        const double cx = x1 + 0.5 * (w - 1.0), cy = y1 + 0.5 * (h - 1.0);
        for (int s = 0; s < n_scales; ++s) {
            const double sc = (double)(float)scales[s];   // scales are float32 constants (anchor.py:103)
            const double sw = w * sc, sh = h * sc;        // anchor.py:185-188
            double *o = out + ((size_t)r * n_scales + s) * 4;
            o[0] = cx - 0.5 * (sw - 1.0);
            o[1] = cy - 0.5 * (sh - 1.0);
            o[2] = cx + 0.5 * (sw - 1.0);
            o[3] = cy + 0.5 * (sh - 1.0);
        }
    }
    return DETBOX_OK;
}

struct PyramidDesc {
    int n_levels;
    int n_cell[DETBOX_MAX_LEVELS];
    int width[DETBOX_MAX_LEVELS];
    int stride[DETBOX_MAX_LEVELS];
    int cell_off[DETBOX_MAX_LEVELS];        // first row of the level in cell_dev
    long long row_off[DETBOX_MAX_LEVELS + 1];  // first output row of the level
};

template <typename T>
__global__ void __launch_bounds__(256)
anchor_pyramid_kernel(const double *__restrict__ cell, PyramidDesc d, T *__restrict__ out)
{
    const long long total = d.row_off[d.n_levels];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        int l = 0;
#pragma unroll
        for (int k = 1; k < DETBOX_MAX_LEVELS; ++k)
            if (k < d.n_levels && i >= d.row_off[k]) l = k;
        const long long r = i - d.row_off[l];
        const int a = (int)(r % d.n_cell[l]);
        const long long loc = r / d.n_cell[l];
        const int x = (int)(loc % d.width[l]);
        const int y = (int)(loc / d.width[l]);
        const double sx = (double)((long long)x * d.stride[l]);
        const double sy = (double)((long long)y * d.stride[l]);
        const double *c = cell + ((size_t)d.cell_off[l] + a) * 4;
        // float64 add, then (optionally) one rounding to float32 == numpy's .astype(float32)
        const double v0 = c[0] + sx, v1 = c[1] + sy, v2 = c[2] + sx, v3 = c[3] + sy;
        if constexpr (sizeof(T) == 4) {
            reinterpret_cast<float4 *>(out)[i] = make_float4((float)v0, (float)v1, (float)v2, (float)v3);
        } else {
            reinterpret_cast<double2 *>(out)[2 * i] = make_double2(v0, v1);
            reinterpret_cast<double2 *>(out)[2 * i + 1] = make_double2(v2, v3);
        }
    }
}

extern "C" int detbox_anchor_pyramid(const double *cell_dev, int n_levels, const int *n_cell,
                                     const int *heights, const int *widths, const int *strides,
                                     void *out_dev, int out_f64, void *stream)
{
    if (!cell_dev || !out_dev || !n_cell || !heights || !widths || !strides) return DETBOX_ERR_INVALID;
    if (n_levels <= 0 || n_levels > DETBOX_MAX_LEVELS) return DETBOX_ERR_INVALID;
    PyramidDesc d;
    d.n_levels = n_levels;
    int coff = 0; long long roff = 0;
    for (int l = 0; l < DETBOX_MAX_LEVELS; ++l) {
        if (l < n_levels) {
            if (n_cell[l] <= 0 || heights[l] < 0 || widths[l] < 0) return DETBOX_ERR_INVALID;
            d.n_cell[l] = n_cell[l]; d.width[l] = widths[l] > 0 ? widths[l] : 1; d.stride[l] = strides[l];
            d.cell_off[l] = coff; d.row_off[l] = roff;
            coff += n_cell[l];
            roff += (long long)heights[l] * widths[l] * n_cell[l];
        } else {
            d.n_cell[l] = 1; d.width[l] = 1; d.stride[l] = 0; d.cell_off[l] = 0; d.row_off[l] = roff;
        }
    }
    d.row_off[n_levels] = roff;
    for (int l = n_levels + 1; l <= DETBOX_MAX_LEVELS; ++l) d.row_off[l] = roff;
    if (roff == 0) return DETBOX_OK;
    long long blocks = (roff + 255) / 256;
    const long long cap = (long long)kNumSMs * 16;
    if (blocks > cap) blocks = cap;
    DETBOX_TRACE_BEGIN(DETBOX_K_ANCHORS, as_stream(stream));
    if (out_f64) anchor_pyramid_kernel<double><<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(cell_dev, d, (double *)out_dev);
    else anchor_pyramid_kernel<float><<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(cell_dev, d, (float *)out_dev);
    DETBOX_TRACE_END(DETBOX_K_ANCHORS, as_stream(stream));
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}
