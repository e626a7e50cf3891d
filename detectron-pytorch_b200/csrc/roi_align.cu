// roi_align.cu — RoIAlign forward / backward over FPN levels in ONE launch.
// Replaces
//   bin-centre : libs/layers/roi_align_tf (roi_align_tf.py:16-42 + crop_and_resize.c:6-268 /
//                crop_and_resize_kernel.cu:10-165)
//   grid-point : libs/layers/roi_align (roi_align_kernel.cu:15-153; RoIAlignAvg/Max fused,
//                modules/roi_align.py:23-46)
//   pyramid    : pyramid_roi_align.py:30-60 (level-assigned: nonzero + gather + kernel + cat +
//                inverse permutation per level) and :82-91 (all levels, channel concat)
//
// Mapping.  One block owns one RoI (x one level in all-levels mode).  Sampling is separable, so
// the block first fills two tiny shared tables — PH row entries and PW column entries, each
// {tap index, tap index, lerp weight, valid} — with the reference's float32 arithmetic evaluated
// ONCE per RoI instead of once per output element; the level is picked in the same prologue
// (pyramid_roi_align.py:33-37) and the result is written straight to the caller's RoI slot, so no
// per-level nonzero/gather/cat/scatter passes exist.
//   NCHW : threads run over (c, ph, pw) flattened = the output's memory order -> fully coalesced
//          stores; the 4 taps of neighbouring pw share sectors, re-use is served by L1/L2.
//   NHWC : threads run over (bin, c/4): every tap is one float4 load that is contiguous across
//          the warp (channel-coalesced), the store is a float4 as well.
// Backward scatters grad * weight into the 4 taps with red.global.add (float4-wide vector
// reductions in NHWC, sm_90+), after an optional cudaMemsetAsync of the dense gradient maps.
#include <stdlib.h>
#include <type_traits>
#include <algorithm>
#include "common.cuh"

using namespace detbox;

namespace {

constexpr int RA_THREADS = 256;
constexpr int RA_MAXP = 33;          // max pooled size + 1 (grid-point pooled variants sample P+1)

struct Tap { int i0, i1; float w; int valid; };

struct RoiArgs {
    detbox_roi_cfg cfg;
    const float *rois; long R;
    const int32_t *levels;
    float *out;                // forward
    const float *grad_out;     // backward
};

// FPN level — pyramid_roi_align.py:33-37, float32, first minimum wins
__device__ __forceinline__ int pick_level(const detbox_roi_cfg &c, const float *roi) {
    const float area = (roi[3] - roi[1] + 1.0f) * (roi[4] - roi[2] + 1.0f);
    int best = 0; float bv = 0.0f;
    for (int l = 0; l < c.n_levels; ++l) {
        const float v = fabsf(logf(sqrtf(area / c.ref_area[l])) / 0.6931472f);
        if (l == 0 || v < bv) { bv = v; best = l; }
    }
    return best;
}

// crop_and_resize taps along one axis from a NORMALISED interval [n0, n1]: crop_and_resize.c:44-72
__device__ __forceinline__ Tap crop_tap(float n0, float n1, int P, int L, int p) {
    const float step = (P > 1) ? (n1 - n0) * (L - 1) / (P - 1) : 0;
    const float in = (P > 1) ? n0 * (L - 1) + p * step : (float)(0.5 * (double)(n0 + n1) * (L - 1));
    Tap t;
    t.valid = !(in < 0 || in > L - 1);
    const int a = (int)floorf(in), b = (int)ceilf(in);
    t.i0 = a; t.i1 = b; t.w = in - a;
    if (!t.valid) { t.i0 = t.i1 = 0; t.w = 0.0f; }
    return t;
}

// bin-centre taps along one axis: roi_align_tf.py:24-40 (normalise), then crop_tap
__device__ __forceinline__ Tap tf_tap(float lo, float hi, float scale, int P, int L, int p) {
    const float b1 = scale * lo, b2 = scale * hi;
    const float sp = (b2 - b1) / (float)P;
    const float n0 = (b1 + sp / 2 - 0.5f) / (float)(L - 1);
    const float nl = sp * (float)(P - 1) / (float)(L - 1);
    return crop_tap(n0, n0 + nl, P, L, p);
}

// grid-point taps along one axis: roi_align_kernel.cu:32-56 (G = number of grid points)
__device__ __forceinline__ Tap gp_tap(float lo, float hi, float scale, int G, int L, int p) {
    const float s = lo * scale, e = hi * scale;
    const float len = fmaxf(e - s + 1.0f, 0.0f);
    const float bin = len / (float)(G - 1);
    const float v = (float)p * bin + s;
    Tap t;
    t.valid = !(v < 0 || v >= L);
    int st = (int)fminf(floorf(v), (float)(L - 2));
    t.w = v - (float)st;
    if (st < 0) st = 0;                 // L == 1: the reference reads out of bounds here
    t.i0 = st; t.i1 = min(st + 1, L - 1);
    if (!t.valid) { t.i0 = t.i1 = 0; t.w = 0.0f; }
    return t;
}

struct Prologue { int level, b, H, W, ok; long r; };   // r: the RoI this block owns (blockIdx.x, or roi_order[blockIdx.x])

__device__ __forceinline__ Prologue roi_prologue(const RoiArgs &a, Tap *ty, Tap *tx, int &gh, int &gw)
{
    __shared__ Prologue s;
    const detbox_roi_cfg &c = a.cfg;
    // Which RoI a block owns is free: the result is written to that RoI's own slot.  roi_order (detbox_roi_locality_order) deals
    // RoIs to consecutive blocks by (level, image, row band, column band), so that overlapping RoIs run close together in
    // time and their shared feature rows are L2 hits instead of DRAM re-reads.
    const long ridx = c.roi_order ? (long)c.roi_order[blockIdx.x] : (long)blockIdx.x;
    const float *roi = a.rois + (size_t)ridx * 5;
    if (threadIdx.x == 0) {
        int l = 0;
        s.r = ridx;
        if (c.all_levels) l = blockIdx.y;
        else if (a.levels) l = a.levels[ridx];
        else if (c.n_levels > 1) l = pick_level(c, roi);
        s.level = l;
        s.b = (int)roi[0];
        s.ok = l >= 0 && l < c.n_levels && s.b >= 0 && s.b < c.batch;
        if (l < 0 || l >= c.n_levels) l = 0;
        s.H = c.height[l]; s.W = c.width[l];
    }
    __syncthreads();
    const int l = min(max(s.level, 0), c.n_levels - 1);
    const int pool = (c.mode == DETBOX_ROI_GRID_POINT && c.pool != DETBOX_POOL_NONE) ? 1 : 0;
    gh = c.pooled_h + pool; gw = c.pooled_w + pool;
    const int t = threadIdx.x;
    if (t < gh) ty[t] = (c.mode == DETBOX_ROI_BIN_CENTRE) ? tf_tap(roi[2], roi[4], c.spatial_scale[l], gh, s.H, t)
                      : (c.mode == DETBOX_ROI_CROP_NORM)  ? crop_tap(roi[1], roi[3], gh, s.H, t)
                                                          : gp_tap(roi[2], roi[4], c.spatial_scale[l], gh, s.H, t);
    else if (t >= 64 && t - 64 < gw)
        tx[t - 64] = (c.mode == DETBOX_ROI_BIN_CENTRE) ? tf_tap(roi[1], roi[3], c.spatial_scale[l], gw, s.W, t - 64)
                   : (c.mode == DETBOX_ROI_CROP_NORM)  ? crop_tap(roi[2], roi[4], gw, s.W, t - 64)
                                                       : gp_tap(roi[1], roi[3], c.spatial_scale[l], gw, s.W, t - 64);
    __syncthreads();
    return s;
}

// one sample from its 4 taps
template <bool GRID>
__device__ __forceinline__ float blend(float tl, float tr, float bl, float br, float wy, float wx) {
    if (GRID) {   // roi_align_kernel.cu:62-65 — double-promoted weights and sum
        const double hr = wy, wr = wx;
        return (float)((double)tl * (1. - hr) * (1. - wr) + (double)tr * (1. - hr) * wr
                       + (double)bl * hr * (1. - wr) + (double)br * hr * wr);
    }
    const float top = tl + (tr - tl) * wx;        // crop_and_resize.c:101-105
    const float bot = bl + (br - bl) * wx;
    return top + (bot - top) * wy;
}

template <bool NHWC>
__device__ __forceinline__ size_t feat_index(int b, int c, int y, int x, int C, int H, int W) {
    return NHWC ? (((size_t)b * H + y) * W + x) * C + c : (((size_t)b * C + c) * H + y) * W + x;
}

template <bool GRID, bool NHWC_IN>
__device__ __forceinline__ float sample_one(const float *f, int b, int c, const Tap &y, const Tap &x, int C, int H, int W, float ext) {
    if (!(y.valid && x.valid)) return ext;
    const float tl = f[feat_index<NHWC_IN>(b, c, y.i0, x.i0, C, H, W)];
    const float tr = f[feat_index<NHWC_IN>(b, c, y.i0, x.i1, C, H, W)];
    const float bl = f[feat_index<NHWC_IN>(b, c, y.i1, x.i0, C, H, W)];
    const float br = f[feat_index<NHWC_IN>(b, c, y.i1, x.i1, C, H, W)];
    return blend<GRID>(tl, tr, bl, br, y.w, x.w);
}

// ---- forward ------------------------------------------------------------------------------------
template <bool GRID, bool NHWC_IN, bool NHWC_OUT, int RA_UNROLL>
__global__ void __launch_bounds__(RA_THREADS)
roi_fwd_kernel(const RoiArgs a)
{
    __shared__ Tap ty[RA_MAXP], tx[RA_MAXP];
    int gh, gw;
    const Prologue s = roi_prologue(a, ty, tx, gh, gw);
    const detbox_roi_cfg &c = a.cfg;
    const int PH = c.pooled_h, PW = c.pooled_w, C = c.channels;
    const int Ctot = c.all_levels ? C * c.n_levels : C;
    const int coff = c.all_levels ? C * blockIdx.y : 0;
    const int bins = PH * PW;
    const long r = s.r;
    const float *f = s.ok ? c.feat[s.level] : nullptr;
    const int pool = (GRID && c.pool != DETBOX_POOL_NONE) ? c.pool : DETBOX_POOL_NONE;
    const float ext = GRID ? 0.0f : c.extrapolation;

    if (NHWC_IN && NHWC_OUT && (C & 3) == 0 && pool == DETBOX_POOL_NONE) {
        // channel-coalesced float4 path
        // Each thread keeps RA_UNROLL items = 4*RA_UNROLL independent 16-byte loads in flight before the
        // first blend: the kernel is a latency-bound gather, bytes in flight are what buys bandwidth.
        const int C4 = C >> 2, total = bins * C4;
        const float4 *p = reinterpret_cast<const float4 *>(f);
        const size_t rowb = (size_t)s.b * s.H;
        float4 *outp = reinterpret_cast<float4 *>(a.out) + ((((size_t)r * bins) * Ctot + coff) >> 2);
        const int Ctot4 = Ctot >> 2;
        for (int base = threadIdx.x; base < total; base += RA_THREADS * RA_UNROLL) {
            float4 tl[RA_UNROLL], tr[RA_UNROLL], bl[RA_UNROLL], br[RA_UNROLL];
            float wy[RA_UNROLL], wx[RA_UNROLL];
            int oidx[RA_UNROLL]; bool live[RA_UNROLL], ok[RA_UNROLL];
#pragma unroll
            for (int u = 0; u < RA_UNROLL; ++u) {
                const int idx = base + u * RA_THREADS;
                live[u] = idx < total;
                const int bin = live[u] ? idx / C4 : 0, c4 = idx - bin * C4;
                const int y = bin / PW, x = bin - y * PW;
                const Tap Y = ty[y], X = tx[x];
                ok[u] = live[u] && f && Y.valid && X.valid;
                wy[u] = Y.w; wx[u] = X.w; oidx[u] = bin * Ctot4 + c4;
                if (ok[u]) {
                    const size_t r0 = (rowb + Y.i0) * s.W, r1 = (rowb + Y.i1) * s.W;
                    tl[u] = __ldg(p + (((r0 + X.i0) * C) >> 2) + c4);
                    tr[u] = __ldg(p + (((r0 + X.i1) * C) >> 2) + c4);
                    bl[u] = __ldg(p + (((r1 + X.i0) * C) >> 2) + c4);
                    br[u] = __ldg(p + (((r1 + X.i1) * C) >> 2) + c4);
                }
            }
#pragma unroll
            for (int u = 0; u < RA_UNROLL; ++u) {
                if (!live[u]) continue;
                float4 v = f ? make_float4(ext, ext, ext, ext) : make_float4(0, 0, 0, 0);
                if (ok[u]) {
                    v.x = blend<GRID>(tl[u].x, tr[u].x, bl[u].x, br[u].x, wy[u], wx[u]);
                    v.y = blend<GRID>(tl[u].y, tr[u].y, bl[u].y, br[u].y, wy[u], wx[u]);
                    v.z = blend<GRID>(tl[u].z, tr[u].z, bl[u].z, br[u].z, wy[u], wx[u]);
                    v.w = blend<GRID>(tl[u].w, tr[u].w, bl[u].w, br[u].w, wy[u], wx[u]);
                }
                __stcs(outp + oidx[u], v);               // streaming store: the output is not re-read here
            }
        }
        return;
    }
    // generic path: iterate in the OUTPUT's memory order
    for (int idx = threadIdx.x; idx < bins * C; idx += RA_THREADS) {
        int ch, bin;
        if (NHWC_OUT) { ch = idx % C; bin = idx / C; } else { ch = idx / bins; bin = idx - ch * bins; }
        const int y = bin / PW, x = bin - y * PW;
        float v = 0.0f;
        if (f) {
            if (pool == DETBOX_POOL_NONE) {
                v = sample_one<GRID, NHWC_IN>(f, s.b, ch, ty[y], tx[x], C, s.H, s.W, ext);
            } else {   // 2x2 / stride 1 pool over the (PH+1)x(PW+1) sample grid (modules/roi_align.py:23-46)
                const float v00 = sample_one<GRID, NHWC_IN>(f, s.b, ch, ty[y], tx[x], C, s.H, s.W, ext);
                const float v01 = sample_one<GRID, NHWC_IN>(f, s.b, ch, ty[y], tx[x + 1], C, s.H, s.W, ext);
                const float v10 = sample_one<GRID, NHWC_IN>(f, s.b, ch, ty[y + 1], tx[x], C, s.H, s.W, ext);
                const float v11 = sample_one<GRID, NHWC_IN>(f, s.b, ch, ty[y + 1], tx[x + 1], C, s.H, s.W, ext);
                v = (pool == DETBOX_POOL_AVG2) ? (((v00 + v01) + v10) + v11) * 0.25f
                                               : fmaxf(fmaxf(v00, v01), fmaxf(v10, v11));
            }
        }
        const size_t o = NHWC_OUT ? ((size_t)r * bins + bin) * Ctot + coff + ch
                                  : ((size_t)r * Ctot + coff + ch) * bins + bin;
        a.out[o] = v;
    }
}

// ---- backward -----------------------------------------------------------------------------------
__device__ __forceinline__ void red_add4(float *p, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <bool GRID>
__device__ __forceinline__ void tap_weights(float g, float wy, float wx, float &w00, float &w01, float &w10, float &w11) {
    if (GRID) {   // roi_align_kernel.cu:147-150
        const double hr = wy, wr = wx;
        w00 = (float)((double)g * (1. - hr) * (1 - wr)); w01 = (float)((double)g * (1. - hr) * wr);
        w10 = (float)((double)g * hr * (1 - wr));        w11 = (float)((double)g * hr * wr);
    } else {      // crop_and_resize.c:258-264
        const float dtop = (1 - wy) * g, dbot = wy * g;
        w00 = (1 - wx) * dtop; w01 = wx * dtop; w10 = (1 - wx) * dbot; w11 = wx * dbot;
    }
}

template <bool GRID, bool NHWC_IN, bool NHWC_OUT>
__global__ void __launch_bounds__(RA_THREADS)
roi_bwd_kernel(const RoiArgs a)
{
    __shared__ Tap ty[RA_MAXP], tx[RA_MAXP];
    int gh, gw;
    const Prologue s = roi_prologue(a, ty, tx, gh, gw);
    if (!s.ok) return;
    const detbox_roi_cfg &c = a.cfg;
    const int PH = c.pooled_h, PW = c.pooled_w, C = c.channels;
    const int Ctot = c.all_levels ? C * c.n_levels : C;
    const int coff = c.all_levels ? C * blockIdx.y : 0;
    const int bins = PH * PW;
    const long r = s.r;
    float *gf = c.grad_feat[s.level];
    const int pool = (GRID && c.pool != DETBOX_POOL_NONE) ? c.pool : DETBOX_POOL_NONE;

    if (pool != DETBOX_POOL_NONE) {
        // RoIAlignAvg / RoIAlignMax (modules/roi_align.py:23-46): the 2x2 / stride-1 pool over the
        // (PH+1)x(PW+1) sample grid is differentiated here, fused with the 4-tap scatter.
        const float *f = c.feat[s.level];
        for (int idx = threadIdx.x; idx < bins * C; idx += RA_THREADS) {
            int ch, bin;
            if (NHWC_OUT) { ch = idx % C; bin = idx / C; } else { ch = idx / bins; bin = idx - ch * bins; }
            const int y = bin / PW, x = bin - y * PW;
            const size_t o = NHWC_OUT ? ((size_t)r * bins + bin) * Ctot + coff + ch
                                      : ((size_t)r * Ctot + coff + ch) * bins + bin;
            const float g = a.grad_out[o];
            int pick = -1;
            if (pool == DETBOX_POOL_MAX2) {          // first maximum in window order, like max_pool2d
                float best = 0.0f;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float v = sample_one<GRID, NHWC_IN>(f, s.b, ch, ty[y + (k >> 1)], tx[x + (k & 1)], C, s.H, s.W, 0.0f);
                    if (k == 0 || v > best) { best = v; pick = k; }
                }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (pick >= 0 && k != pick) continue;
                const Tap Y = ty[y + (k >> 1)], X = tx[x + (k & 1)];
                if (!(Y.valid && X.valid)) continue;
                float w00, w01, w10, w11;
                tap_weights<GRID>(pick >= 0 ? g : 0.25f * g, Y.w, X.w, w00, w01, w10, w11);
                atomicAdd(gf + feat_index<NHWC_IN>(s.b, ch, Y.i0, X.i0, C, s.H, s.W), w00);
                atomicAdd(gf + feat_index<NHWC_IN>(s.b, ch, Y.i0, X.i1, C, s.H, s.W), w01);
                atomicAdd(gf + feat_index<NHWC_IN>(s.b, ch, Y.i1, X.i0, C, s.H, s.W), w10);
                atomicAdd(gf + feat_index<NHWC_IN>(s.b, ch, Y.i1, X.i1, C, s.H, s.W), w11);
            }
        }
        return;
    }

    if (NHWC_IN && NHWC_OUT && (C & 3) == 0) {
        const int C4 = C >> 2;
        for (int idx = threadIdx.x; idx < bins * C4; idx += RA_THREADS) {
            const int c4 = idx % C4, bin = idx / C4;
            const int y = bin / PW, x = bin - y * PW;
            const Tap Y = ty[y], X = tx[x];
            if (!(Y.valid && X.valid)) continue;
            const float4 g = reinterpret_cast<const float4 *>(a.grad_out)[((((size_t)r * bins + bin) * Ctot + coff) >> 2) + c4];
            float4 w00, w01, w10, w11;
            tap_weights<GRID>(g.x, Y.w, X.w, w00.x, w01.x, w10.x, w11.x);
            tap_weights<GRID>(g.y, Y.w, X.w, w00.y, w01.y, w10.y, w11.y);
            tap_weights<GRID>(g.z, Y.w, X.w, w00.z, w01.z, w10.z, w11.z);
            tap_weights<GRID>(g.w, Y.w, X.w, w00.w, w01.w, w10.w, w11.w);
            const size_t rowb = (size_t)s.b * s.H;
            red_add4(gf + ((rowb + Y.i0) * s.W + X.i0) * C + 4 * c4, w00);
            red_add4(gf + ((rowb + Y.i0) * s.W + X.i1) * C + 4 * c4, w01);
            red_add4(gf + ((rowb + Y.i1) * s.W + X.i0) * C + 4 * c4, w10);
            red_add4(gf + ((rowb + Y.i1) * s.W + X.i1) * C + 4 * c4, w11);
        }
        return;
    }
    for (int idx = threadIdx.x; idx < bins * C; idx += RA_THREADS) {
        int ch, bin;
        if (NHWC_OUT) { ch = idx % C; bin = idx / C; } else { ch = idx / bins; bin = idx - ch * bins; }
        const int y = bin / PW, x = bin - y * PW;
        const Tap Y = ty[y], X = tx[x];
        if (!(Y.valid && X.valid)) continue;
        const size_t o = NHWC_OUT ? ((size_t)r * bins + bin) * Ctot + coff + ch
                                  : ((size_t)r * Ctot + coff + ch) * bins + bin;
        const float g = a.grad_out[o];
        float w00, w01, w10, w11;
        tap_weights<GRID>(g, Y.w, X.w, w00, w01, w10, w11);
        atomicAdd(gf + feat_index<NHWC_IN>(s.b, ch, Y.i0, X.i0, C, s.H, s.W), w00);
        atomicAdd(gf + feat_index<NHWC_IN>(s.b, ch, Y.i0, X.i1, C, s.H, s.W), w01);
        atomicAdd(gf + feat_index<NHWC_IN>(s.b, ch, Y.i1, X.i0, C, s.H, s.W), w10);
        atomicAdd(gf + feat_index<NHWC_IN>(s.b, ch, Y.i1, X.i1, C, s.H, s.W), w11);
    }
}

// ---- NCHW, shared-memory staged (the reference's own layout: [B,C,H,W] features, [R,C,PH,PW] output) ---------------
// In NCHW the 4 taps of an output element lie in 2 rows of ONE channel plane, and a RoI's taps of one channel cover a
// small window of that plane (about 11 x 11 pixels after level assignment).  The generic path gathers them one scalar
// at a time: every 32-byte sector is requested ~3 times through L1 and the lanes of a request scatter over 7 rows.
// Here ONE block owns a RoI (the tap tables and the window are computed once) and walks the channels in chunks: the
// window of every channel of the chunk — whole 16-byte pieces of each row, so a sector is requested exactly once — is
// copied into shared memory with cp.async (no registers; the next chunk is in flight while this one is sampled), then
// every thread samples ITS bin (tap offsets and weights held in registers) for one channel after the other and the
// [C,PH,PW] output leaves in memory order.  Backward is the mirror image: contributions are summed in a zeroed shared
// window (shared-memory atomics) and every touched 16-byte piece leaves through ONE red.global.add.v4.f32 —
// block-aggregated atomics instead of 4 per output element.  A window too large for the buffer (RoIs far larger than
// their level) falls back to the direct gather / scatter.
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_CHMAX = 32;                // channels per chunk (at most)
constexpr int RS_HALF = 5 * 1024;           // floats per staging buffer; forward uses two (40 KB: 4-5 blocks per SM)
constexpr int RS_FLOATS = 2 * RS_HALF;
constexpr int RS_KREG = 3;                  // 16-byte pieces per lane whose offsets live in registers (96 per channel)

struct Window { int y0, x0, wh, ww, vec; }; // ww: row pitch in floats (multiple of 4 when vec)

__device__ __forceinline__ Window roi_window(const Tap *ty, const Tap *tx, int gh, int gw, int W, const float *plane0, size_t plane)
{
    __shared__ Window s;
    if (threadIdx.x < 32) {
        const int l = threadIdx.x;
        int ymin = 1 << 30, ymax = -1, xmin = 1 << 30, xmax = -1;
        for (int i = l; i < gh; i += 32) if (ty[i].valid) { ymin = min(ymin, min(ty[i].i0, ty[i].i1)); ymax = max(ymax, max(ty[i].i0, ty[i].i1)); }
        for (int i = l; i < gw; i += 32) if (tx[i].valid) { xmin = min(xmin, min(tx[i].i0, tx[i].i1)); xmax = max(xmax, max(tx[i].i0, tx[i].i1)); }
        ymin = __reduce_min_sync(0xffffffffu, ymin); ymax = __reduce_max_sync(0xffffffffu, ymax);
        xmin = __reduce_min_sync(0xffffffffu, xmin); xmax = __reduce_max_sync(0xffffffffu, xmax);
        if (l == 0) {
            Window w;
            // 16-byte pieces need every row of every plane 16-byte aligned: W % 4 == 0 (then plane % 4 == 0 too) and an aligned base
            w.vec = ((W & 3) == 0) && ((plane & 3) == 0) && (((size_t)plane0 & 15) == 0);
            if (ymax < 0 || xmax < 0) { w.y0 = w.x0 = 0; w.wh = w.ww = 0; }
            else {
                w.y0 = ymin; w.wh = ymax - ymin + 1;
                w.x0 = w.vec ? (xmin & ~3) : xmin;
                w.ww = w.vec ? (((xmax + 4) & ~3) - w.x0) : (xmax - xmin + 1);
            }
            s = w;
        }
    }
    __syncthreads();
    return s;
}

__device__ __forceinline__ void cp_async16(float *smem_dst, const float *gmem_src) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async4(float *smem_dst, const float *gmem_src) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// piece k of a channel's window -> offsets in the staged window and in the feature plane (unit: floats)
__device__ __forceinline__ void piece_offsets(const Window &w, int W, int k, int &so, int &go) {
    if (w.vec) { const int rq = w.ww >> 2, y = k / rq, q = k - y * rq; so = y * w.ww + 4 * q; go = y * W + 4 * q; }
    else { const int y = k / w.ww, x = k - y * w.ww; so = k; go = y * W + x; }
}

// One bin's four taps as offsets into a channel's staged window, with the lerp weights
struct BinTaps { int o00, o01, o10, o11; float wy, wx; bool valid; };
__device__ __forceinline__ BinTaps bin_taps(const Window &w, const Tap &Y, const Tap &X) {
    BinTaps b;
    b.valid = Y.valid && X.valid;
    const int r0 = (Y.i0 - w.y0) * w.ww - w.x0, r1 = (Y.i1 - w.y0) * w.ww - w.x0;
    b.o00 = r0 + X.i0; b.o01 = r0 + X.i1; b.o10 = r1 + X.i0; b.o11 = r1 + X.i1;
    b.wy = Y.w; b.wx = X.w;
    if (!b.valid) b.o00 = b.o01 = b.o10 = b.o11 = 0;
    return b;
}
template <bool GRID>
__device__ __forceinline__ float sample_taps(const float *win, const BinTaps &b, float ext) {
    if (!b.valid) return ext;
    return blend<GRID>(win[b.o00], win[b.o01], win[b.o10], win[b.o11], b.wy, b.wx);
}

template <bool GRID>
__global__ void __launch_bounds__(RS_THREADS)
roi_fwd_nchw_staged_kernel(const RoiArgs a)
{
    __shared__ Tap ty[RA_MAXP], tx[RA_MAXP];
    extern __shared__ __align__(16) float s_win[];
    int gh, gw;
    const Prologue s = roi_prologue(a, ty, tx, gh, gw);
    const detbox_roi_cfg &c = a.cfg;
    const int PH = c.pooled_h, PW = c.pooled_w, C = c.channels;
    const int Ctot = c.all_levels ? C * c.n_levels : C;
    const int coff = c.all_levels ? C * blockIdx.y : 0;
    const int bins = PH * PW;
    const long r = s.r;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const float *f = s.ok ? c.feat[s.level] : nullptr;
    const int pool = (GRID && c.pool != DETBOX_POOL_NONE) ? c.pool : DETBOX_POOL_NONE;
    const float ext = GRID ? 0.0f : c.extrapolation;
    float *out = a.out + ((size_t)r * Ctot + coff) * bins;
    if (!f) {                                                   // bad level / batch index: zeros (as the generic path)
        for (int i = tid; i < C * bins; i += RS_THREADS) out[i] = 0.0f;
        return;
    }
    const size_t plane = (size_t)s.H * s.W;
    const float *f0 = f + (size_t)s.b * C * plane;
    const Window w = roi_window(ty, tx, gh, gw, s.W, f0, plane);
    const int cs = w.wh * w.ww;                                 // floats per staged channel
    if (cs == 0) {                                              // every sample outside the map
        for (int i = tid; i < C * bins; i += RS_THREADS) out[i] = ext;
        return;
    }
    if (cs > RS_HALF) {                                         // a channel's window does not fit: direct gather
        for (int i = tid; i < C * bins; i += RS_THREADS) {
            const int ch = i / bins, bin = i - ch * bins, y = bin / PW, x = bin - y * PW;
            float v;
            if (pool == DETBOX_POOL_NONE) v = sample_one<GRID, false>(f, s.b, ch, ty[y], tx[x], C, s.H, s.W, ext);
            else {
                const float v00 = sample_one<GRID, false>(f, s.b, ch, ty[y], tx[x], C, s.H, s.W, ext);
                const float v01 = sample_one<GRID, false>(f, s.b, ch, ty[y], tx[x + 1], C, s.H, s.W, ext);
                const float v10 = sample_one<GRID, false>(f, s.b, ch, ty[y + 1], tx[x], C, s.H, s.W, ext);
                const float v11 = sample_one<GRID, false>(f, s.b, ch, ty[y + 1], tx[x + 1], C, s.H, s.W, ext);
                v = (pool == DETBOX_POOL_AVG2) ? (((v00 + v01) + v10) + v11) * 0.25f : fmaxf(fmaxf(v00, v01), fmaxf(v10, v11));
            }
            out[i] = v;
        }
        return;
    }
    const int CH = min(RS_CHMAX, RS_HALF / cs);                 // channels per chunk
    const int per = w.vec ? w.wh * (w.ww >> 2) : cs;            // copies per channel (16-byte pieces, or floats)
    const float *src0 = f0 + (size_t)w.y0 * s.W + w.x0;
    // this lane's first RS_KREG pieces of every channel: offsets computed once, kept in registers
    int so[RS_KREG], go[RS_KREG];
#pragma unroll
    for (int j = 0; j < RS_KREG; ++j) { so[j] = go[j] = 0; if (lane + 32 * j < per) piece_offsets(w, s.W, lane + 32 * j, so[j], go[j]); }
    auto stage = [&](int c0, float *buf) {                      // warp = channel, lanes = pieces
        const int n = min(CH, C - c0);
        for (int ch = wid; ch < n; ch += RS_WARPS) {
            float *d = buf + ch * cs;
            const float *g = src0 + (size_t)(c0 + ch) * plane;
#pragma unroll
            for (int j = 0; j < RS_KREG; ++j)
                if (lane + 32 * j < per) { if (w.vec) cp_async16(d + so[j], g + go[j]); else cp_async4(d + so[j], g + go[j]); }
            for (int k = lane + 32 * RS_KREG; k < per; k += 32) {
                int o, q; piece_offsets(w, s.W, k, o, q);
                if (w.vec) cp_async16(d + o, g + q); else cp_async4(d + o, g + q);
            }
        }
        cp_async_commit();
    };
    // this thread's bin (threads 0..63 of every group of 64 own bins 0..63; 4 channel groups)
    const int bl = tid & 63, cg = tid >> 6;
    stage(0, s_win);
    int it = 0;
    for (int c0 = 0; c0 < C; c0 += CH, ++it) {
        float *buf = s_win + (it & 1) * RS_HALF;
        const bool more = c0 + CH < C;
        if (more) stage(c0 + CH, s_win + ((it + 1) & 1) * RS_HALF);
        if (more) cp_async_wait<1>(); else cp_async_wait<0>();
        __syncthreads();
        const int n = min(CH, C - c0);
        float *o = out + (size_t)c0 * bins;
        for (int bin = bl; bin < bins; bin += 64) {
            const int y = bin / PW, x = bin - y * PW;
            if (pool == DETBOX_POOL_NONE) {
                const BinTaps t = bin_taps(w, ty[y], tx[x]);
                for (int ch = cg; ch < n; ch += RS_THREADS / 64)
                    __stcs(o + ch * bins + bin, sample_taps<GRID>(buf + ch * cs, t, ext));   // streaming store: not re-read here
            } else {   // 2x2 / stride 1 pool over the (PH+1)x(PW+1) sample grid (modules/roi_align.py:23-46)
                const BinTaps t00 = bin_taps(w, ty[y], tx[x]), t01 = bin_taps(w, ty[y], tx[x + 1]);
                const BinTaps t10 = bin_taps(w, ty[y + 1], tx[x]), t11 = bin_taps(w, ty[y + 1], tx[x + 1]);
                for (int ch = cg; ch < n; ch += RS_THREADS / 64) {
                    const float *win = buf + ch * cs;
                    const float v00 = sample_taps<GRID>(win, t00, ext), v01 = sample_taps<GRID>(win, t01, ext);
                    const float v10 = sample_taps<GRID>(win, t10, ext), v11 = sample_taps<GRID>(win, t11, ext);
                    __stcs(o + ch * bins + bin, (pool == DETBOX_POOL_AVG2) ? (((v00 + v01) + v10) + v11) * 0.25f
                                                                            : fmaxf(fmaxf(v00, v01), fmaxf(v10, v11)));
                }
            }
        }
        __syncthreads();                                        // this buffer is refilled two iterations from now
    }
}

template <bool GRID>
__global__ void __launch_bounds__(RS_THREADS)
roi_bwd_nchw_staged_kernel(const RoiArgs a)
{
    __shared__ Tap ty[RA_MAXP], tx[RA_MAXP];
    extern __shared__ __align__(16) float s_win[];
    int gh, gw;
    const Prologue s = roi_prologue(a, ty, tx, gh, gw);
    if (!s.ok) return;
    const detbox_roi_cfg &c = a.cfg;
    const int PH = c.pooled_h, PW = c.pooled_w, C = c.channels;
    const int Ctot = c.all_levels ? C * c.n_levels : C;
    const int coff = c.all_levels ? C * blockIdx.y : 0;
    const int bins = PH * PW;
    const long r = s.r;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const size_t plane = (size_t)s.H * s.W;
    float *g0 = c.grad_feat[s.level] + (size_t)s.b * C * plane;
    const float *go_ = a.grad_out + ((size_t)r * Ctot + coff) * bins;
    const Window w = roi_window(ty, tx, gh, gw, s.W, g0, plane);
    const int cs = w.wh * w.ww;
    if (cs == 0) return;                                        // no sample inside the map: no gradient
    if (cs > RS_FLOATS) {                                       // direct scatter
        for (int i = tid; i < C * bins; i += RS_THREADS) {
            const int ch = i / bins, bin = i - ch * bins, y = bin / PW, x = bin - y * PW;
            const Tap Y = ty[y], X = tx[x];
            if (!(Y.valid && X.valid)) continue;
            float w00, w01, w10, w11;
            tap_weights<GRID>(go_[i], Y.w, X.w, w00, w01, w10, w11);
            float *gp = g0 + (size_t)ch * plane;
            atomicAdd(gp + (size_t)Y.i0 * s.W + X.i0, w00);
            atomicAdd(gp + (size_t)Y.i0 * s.W + X.i1, w01);
            atomicAdd(gp + (size_t)Y.i1 * s.W + X.i0, w10);
            atomicAdd(gp + (size_t)Y.i1 * s.W + X.i1, w11);
        }
        return;
    }
    const int CH = min(RS_CHMAX, RS_FLOATS / cs);               // one buffer: nothing is loaded ahead
    const int per = w.vec ? w.wh * (w.ww >> 2) : cs;
    float *dst0 = g0 + (size_t)w.y0 * s.W + w.x0;
    int so[RS_KREG], go[RS_KREG];
#pragma unroll
    for (int j = 0; j < RS_KREG; ++j) { so[j] = go[j] = 0; if (lane + 32 * j < per) piece_offsets(w, s.W, lane + 32 * j, so[j], go[j]); }
    for (int i = tid; i < min(C, CH) * cs; i += RS_THREADS) s_win[i] = 0.0f;
    __syncthreads();
    const int bl = tid & 63, cg = tid >> 6;
    auto flush_piece = [&](float *d, float *g, int o, int q) {  // one 16-byte piece (or one float) out, zero behind it
        if (w.vec) {
            const float4 v = *reinterpret_cast<const float4 *>(d + o);
            if (v.x != 0.0f || v.y != 0.0f || v.z != 0.0f || v.w != 0.0f) {
                red_add4(g + q, v);
                *reinterpret_cast<float4 *>(d + o) = make_float4(0, 0, 0, 0);
            }
        } else {
            const float v = d[o];
            if (v != 0.0f) { atomicAdd(g + q, v); d[o] = 0.0f; }
        }
    };
    for (int c0 = 0; c0 < C; c0 += CH) {
        const int n = min(CH, C - c0);
        const float *g = go_ + (size_t)c0 * bins;
        for (int bin = bl; bin < bins; bin += 64) {
            const int y = bin / PW, x = bin - y * PW;
            const BinTaps t = bin_taps(w, ty[y], tx[x]);
            if (!t.valid) continue;
            for (int ch = cg; ch < n; ch += RS_THREADS / 64) {
                float w00, w01, w10, w11;
                tap_weights<GRID>(__ldcs(g + ch * bins + bin), t.wy, t.wx, w00, w01, w10, w11);
                float *win = s_win + ch * cs;
                atomicAdd(win + t.o00, w00); atomicAdd(win + t.o01, w01);
                atomicAdd(win + t.o10, w10); atomicAdd(win + t.o11, w11);
            }
        }
        __syncthreads();
        for (int ch = wid; ch < n; ch += RS_WARPS) {            // warp = channel, lanes = pieces
            float *d = s_win + ch * cs;
            float *gp = dst0 + (size_t)(c0 + ch) * plane;
#pragma unroll
            for (int j = 0; j < RS_KREG; ++j)
                if (lane + 32 * j < per) flush_piece(d, gp, so[j], go[j]);
            for (int k = lane + 32 * RS_KREG; k < per; k += 32) {
                int o, q; piece_offsets(w, s.W, k, o, q);
                flush_piece(d, gp, o, q);
            }
        }
        __syncthreads();
    }
}

// ---- NCHW backward as a GATHER ------------------------------------------------------------------------------------------
// The direct NCHW backward issues 4 scalar red.global.add per output element (205 M at BASELINE configs[2]) and is bound by
// the L2's atomic rate (ncu: 27 % issue-active, profiles/r02a_ncu_roi.txt); summing in a shared-memory window first
// (roi_bwd_nchw_staged_kernel) lost to its shared-memory atomics.  Here the two tap tables are INVERTED once per RoI: for
// every feature row a tap touches, the (bin row, weight) pairs that land on it; for every 16-byte piece of a feature row
// a tap touches, the (pixel in piece, bin column, weight) triples.  Then, per channel, one thread owns one (touched row,
// touched piece): it sums its few terms in registers from the RoI's grad_out values (staged per channel chunk with
// coalesced loads) and emits ONE red.global.add.v4.f32 — the block-aggregated atomic north_star (4) asks for; no
// shared-memory atomics, no window buffer, work bounded by touched rows x touched pieces whatever the RoI's size.
// Every term is the reference's own product (crop_and_resize.c:258-264: (1-wx) * ((1-wy) * g); roi_align_kernel.cu:147-150
// in double), only the order of the float32 additions differs (tests: the summation bound of tests/test_gpu_roi.py).
constexpr int RG_THREADS = 256;
constexpr int RG_MAXE = 2 * RA_MAXP;        // a bin row / column contributes to two feature rows / columns
constexpr int RG_GFLOATS = 8192;            // staged grad_out floats per chunk (32 KB)

// One axis inverted.  Entries sorted by feature pixel (stable: bins ascending); a GROUP is a feature row (shift 0) or a
// 16-byte piece of a row (shift 2); sub[g * 4 + p] .. sub[g * 4 + p + 1] are the entries of pixel p of group g, so the
// consumer's accumulator index is a compile-time constant.  w / wd: the weight this entry applies, resolved once —
// (hi ? w : 1 - w) in float32 (crop_and_resize.c:258-264) or in double (roi_align_kernel.cu:147-150).
struct InvAxis {
    short bin[RG_MAXE]; float w[RG_MAXE]; double wd[RG_MAXE];
    short key[RG_MAXE]; short sub[RG_MAXE * 4 + 1];
    short pix[RG_MAXE];                     // builder's scratch: the feature pixel of every entry
    int n_groups;
};

// Serial (one thread): at most 2 * 33 entries.
template <bool GRID>
__device__ __forceinline__ void invert_axis(const Tap *t, const int n_bins, const int shift, InvAxis &ax)
{
    short *pix = ax.pix;
    int n = 0;
    for (int b = 0; b < n_bins; ++b) {
        if (!t[b].valid) continue;
        for (int hi = 0; hi < 2; ++hi) {
            const short px = (short)(hi ? t[b].i1 : t[b].i0);
            const float wf = hi ? t[b].w : 1 - t[b].w;
            const double wdd = hi ? (double)t[b].w : 1. - (double)t[b].w;
            int k = n++;
            while (k > 0 && pix[k - 1] > px) { pix[k] = pix[k - 1]; ax.bin[k] = ax.bin[k - 1]; ax.w[k] = ax.w[k - 1]; if (GRID) ax.wd[k] = ax.wd[k - 1]; --k; }
            pix[k] = px; ax.bin[k] = (short)b; ax.w[k] = wf; if (GRID) ax.wd[k] = wdd;
        }
    }
    int g = 0, k = 0;
    while (k < n) {
        const int key = pix[k] >> shift;
        ax.key[g] = (short)key;
        for (int p = 0; p < 4; ++p) {
            ax.sub[g * 4 + p] = (short)k;
            while (k < n && (pix[k] >> shift) == key && pix[k] - (key << shift) == p) ++k;
        }
        ++g;
    }
    ax.sub[g * 4] = (short)n;
    ax.n_groups = g;
}

template <bool GRID>
__global__ void __launch_bounds__(RG_THREADS, 4)
roi_bwd_nchw_gather_kernel(const RoiArgs a)
{
    __shared__ Tap ty[RA_MAXP], tx[RA_MAXP];
    __shared__ InvAxis s_r, s_c;
    extern __shared__ __align__(16) float s_g[];
    int gh, gw;
    const Prologue s = roi_prologue(a, ty, tx, gh, gw);
    if (!s.ok) return;
    const detbox_roi_cfg &c = a.cfg;
    const int PH = c.pooled_h, PW = c.pooled_w, C = c.channels;
    const int Ctot = c.all_levels ? C * c.n_levels : C;
    const int coff = c.all_levels ? C * blockIdx.y : 0;
    const int bins = PH * PW;
    const long r = s.r;
    const int tid = threadIdx.x;
    const size_t plane = (size_t)s.H * s.W;
    float *g0 = c.grad_feat[s.level] + (size_t)s.b * C * plane;
    const float *go = a.grad_out + ((size_t)r * Ctot + coff) * bins;
    // 16-byte pieces need every row of every plane 16-byte aligned; otherwise pieces are single pixels (scalar atomics)
    const bool vec = ((s.W & 3) == 0) && ((plane & 3) == 0) && (((size_t)g0 & 15) == 0);
    if (tid == 0) invert_axis<GRID>(ty, gh, 0, s_r);
    if (tid == 32) invert_axis<GRID>(tx, gw, vec ? 2 : 0, s_c);
    __syncthreads();
    const int nr = s_r.n_groups, nq = s_c.n_groups;
    if (nr == 0 || nq == 0) return;                             // no sample inside the map: no gradient
    const unsigned per = (unsigned)(nr * nq);                   // (row, piece) items per channel
    const unsigned m_per = 0xffffffffu / per, m_nq = 0xffffffffu / (unsigned)nq;     // floor(2^32 / d) up to one: fixed below
    const int CH = max(1, min(C, RG_GFLOATS / bins));
    for (int c0 = 0; c0 < C; c0 += CH) {
        const int n = min(CH, C - c0);
        if (c0) __syncthreads();                                // the previous chunk has been consumed
        for (int i = tid; i < n * bins; i += RG_THREADS) s_g[i] = __ldcs(go + (size_t)c0 * bins + i);
        __syncthreads();
        for (unsigned i = tid; i < (unsigned)n * per; i += RG_THREADS) {
            unsigned ch = __umulhi(i, m_per), k = i - ch * per;
            if (k >= per) { ++ch; k -= per; }
            unsigned ri = __umulhi(k, m_nq), qi = k - ri * (unsigned)nq;
            if (qi >= (unsigned)nq) { ++ri; qi -= (unsigned)nq; }
            const float *g = s_g + ch * bins;
            float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            const int e0 = s_r.sub[ri * 4], e1 = s_r.sub[ri * 4 + 4];          // a row group holds one pixel row: all its entries
            const short *cs = s_c.sub + qi * 4;
            for (int er = e0; er < e1; ++er) {
                const float *grow = g + s_r.bin[er] * PW;
                const float rw = s_r.w[er];
                const double rwd = GRID ? s_r.wd[er] : 0.0;
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    for (int ec = cs[p]; ec < cs[p + 1]; ++ec) {
                        const float gv = grow[s_c.bin[ec]];
                        if (GRID) acc[p] += (float)((double)gv * rwd * s_c.wd[ec]);   // roi_align_kernel.cu:147-150
                        else acc[p] += s_c.w[ec] * (rw * gv);                          // crop_and_resize.c:258-264: wx-factor * (wy-factor * g)
                    }
                }
            }
            float *dst = g0 + (size_t)(c0 + ch) * plane + (size_t)s_r.key[ri] * s.W;
            if (vec) {
                if (acc[0] != 0.0f || acc[1] != 0.0f || acc[2] != 0.0f || acc[3] != 0.0f)
                    red_add4(dst + 4 * (int)s_c.key[qi], make_float4(acc[0], acc[1], acc[2], acc[3]));
            } else if (acc[0] != 0.0f) atomicAdd(dst + (int)s_c.key[qi], acc[0]);
        }
    }
}

// ---- NCHW backward, separable (bin-centre mode, PW <= 8) -------------------------------------------------------------------------
// The gather kernel above evaluates every product colw * (roww * g) on its own: ~375 products per (RoI, channel) at BASELINE
// configs[2], in 26 (row, piece) items of five short loops each — ncu: 87 % issue-active, L2 at 13 %.  Bilinear weights are
// separable, so the same sums factor: T[r][bx] = sum over the bin rows landing on feature row r of roww * g[by][bx]  (7 partial
// sums per touched row), then out[r][x] = sum over the bin columns landing on pixel x of colw * T[r][bx] — ~270 multiply-adds in
// 2.5 x fewer items, one thread per (channel, touched row), one red.global.add.v4.f32 per touched 16-byte piece.  Each output is
// colw * (sum roww * g) instead of sum colw * (roww * g): the same weights, one rounding apart — inside north_star's 1e-5 for the
// RoIAlign backward and inside the summation bound tests/test_gpu_roi.py asserts (its n + 8 slack covers the extra roundings).
constexpr int RS2_PW = 8;

__global__ void __launch_bounds__(RG_THREADS, 4)
roi_bwd_nchw_sep_kernel(const RoiArgs a)
{
    __shared__ Tap ty[RA_MAXP], tx[RA_MAXP];
    __shared__ InvAxis s_r, s_c;
    __shared__ float s_t[RS2_PW * RG_THREADS];          // T of every thread's current row: [bx][tid], conflict-free
    extern __shared__ __align__(16) float s_g[];
    int gh, gw;
    const Prologue s = roi_prologue(a, ty, tx, gh, gw);
    if (!s.ok) return;
    const detbox_roi_cfg &c = a.cfg;
    const int PH = c.pooled_h, PW = c.pooled_w, C = c.channels;
    const int Ctot = c.all_levels ? C * c.n_levels : C;
    const int coff = c.all_levels ? C * blockIdx.y : 0;
    const int bins = PH * PW;
    const long r = s.r;
    const int tid = threadIdx.x;
    const size_t plane = (size_t)s.H * s.W;
    float *g0 = c.grad_feat[s.level] + (size_t)s.b * C * plane;
    const float *go = a.grad_out + ((size_t)r * Ctot + coff) * bins;
    const bool vec = ((s.W & 3) == 0) && ((plane & 3) == 0) && (((size_t)g0 & 15) == 0);
    if (tid == 0) invert_axis<false>(ty, gh, 0, s_r);
    if (tid == 32) invert_axis<false>(tx, gw, vec ? 2 : 0, s_c);
    __syncthreads();
    const int nr = s_r.n_groups, nq = s_c.n_groups;
    if (nr == 0 || nq == 0) return;                             // no sample inside the map: no gradient
    const unsigned m_nr = 0xffffffffu / (unsigned)nr;
    const int CH = max(1, min(C, RG_GFLOATS / bins));
    float *T = s_t + tid;
    for (int c0 = 0; c0 < C; c0 += CH) {
        const int n = min(CH, C - c0);
        if (c0) __syncthreads();                                // the previous chunk has been consumed
        for (int i = tid; i < n * bins; i += RG_THREADS) s_g[i] = __ldcs(go + (size_t)c0 * bins + i);
        __syncthreads();
        for (unsigned i = tid; i < (unsigned)(n * nr); i += RG_THREADS) {
            unsigned ch = __umulhi(i, m_nr), ri = i - ch * (unsigned)nr;
            if (ri >= (unsigned)nr) { ++ch; ri -= (unsigned)nr; }
            const float *g = s_g + ch * bins;
            // stage 1: the bin rows landing on this feature row
            float t[RS2_PW];
#pragma unroll
            for (int bx = 0; bx < RS2_PW; ++bx) t[bx] = 0.0f;
            for (int er = s_r.sub[ri * 4], e1 = s_r.sub[ri * 4 + 4]; er < e1; ++er) {
                const float *grow = g + s_r.bin[er] * PW;
                const float rw = s_r.w[er];
#pragma unroll
                for (int bx = 0; bx < RS2_PW; ++bx) if (bx < PW) t[bx] += rw * grow[bx];
            }
#pragma unroll
            for (int bx = 0; bx < RS2_PW; ++bx) T[bx * RG_THREADS] = t[bx];
            // stage 2: the bin columns landing on every touched piece of this row (only this thread reads its T: no barrier)
            float *dst = g0 + (size_t)(c0 + ch) * plane + (size_t)s_r.key[ri] * s.W;
            for (int qi = 0; qi < nq; ++qi) {
                const short *cs = s_c.sub + qi * 4;
                float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
                for (int p = 0; p < 4; ++p)
                    for (int ec = cs[p]; ec < cs[p + 1]; ++ec) acc[p] += s_c.w[ec] * T[s_c.bin[ec] * RG_THREADS];
                if (vec) {
                    if (acc[0] != 0.0f || acc[1] != 0.0f || acc[2] != 0.0f || acc[3] != 0.0f)
                        red_add4(dst + 4 * (int)s_c.key[qi], make_float4(acc[0], acc[1], acc[2], acc[3]));
                } else if (acc[0] != 0.0f) atomicAdd(dst + (int)s_c.key[qi], acc[0]);
            }
        }
    }
}

// ---- NCHW backward, tile-centric gather: the dense gradient written ONCE (round 2, late) -----------------------------------------
// Every formulation above is RoI-centric: zero-fill the dense gradient pyramid (1.43 GB at BASELINE configs[2]), then read-modify-
// write the scattered 16-byte pieces the RoIs touch (0.95 GB read + 0.49 GB written): 2.9 GB for a result of 1.43 GB.  Here the
// OUTPUT is the loop: a block owns a tile of RT_TH x RT_TW pixels of one (level, image) plane and, chunk of RT_CH channels by chunk,
// sums in registers what every RoI touching the tile contributes to its pixel, then stores the tile — zeros included — exactly once,
// 128 contiguous bytes per warp and channel.  No memset, no atomics, no read of the gradient; grad_out (205 MB) is re-read from L2
// by the 2-4 tiles a RoI touches.
//   prep kernel: per RoI its level, image, touched pixel window and the PH + PW taps (the SAME tf_tap / crop_tap as the forward).
//   tile kernel: the block scans the R window records (16 bytes each), keeps the RoIs that meet its tile in index order (ordered
//   compaction, so the summation order is fixed: RoI ascending, bin row ascending, bin column ascending — deterministic, unlike the
//   reference's atomics), then per channel chunk stages grad_out of up to NB listed RoIs and accumulates.
// A pixel's term is the reference's own product, colw * (roww * g) (crop_and_resize.c:258-264); the weights of a feature row are
// warp-uniform (a warp is one tile row), the at most two bin columns landing on a pixel are found once per RoI.
constexpr int RT_TH = 8, RT_TW = 32, RT_THREADS = RT_TH * RT_TW, RT_CH = 32, RT_CAP = 256, RT_NBMAX = 4, RT_J = 16;
constexpr int RT_NC = 4;                         // listed RoIs whose weight tables are kept for the whole round
constexpr int RT_NOTAP = -30000;
constexpr int RT_PMAX = 8;                       // pooled size covered (7 x 7, 8 x 8); larger outputs keep the scatter kernels
constexpr int RT_APITCH = RT_CH + 1;             // accumulator pitch: [pixel][33]
struct RtTap { short i0, i1; float w; };          // i0 < 0: the bin samples nothing inside the map
struct TileGeom { int off[DETBOX_MAX_LEVELS + 1]; int nty[DETBOX_MAX_LEVELS], ntx[DETBOX_MAX_LEVELS]; };

__global__ void __launch_bounds__(256)
roi_tile_prep_kernel(const RoiArgs a, int4 *__restrict__ hdr, RtTap *__restrict__ taps)
{
    const detbox_roi_cfg &c = a.cfg;
    const int PH = c.pooled_h, PW = c.pooled_w;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < a.R; i += (long)gridDim.x * blockDim.x) {
        const float *roi = a.rois + (size_t)i * 5;
        int l = 0;
        if (a.levels) l = a.levels[i];
        else if (c.n_levels > 1) l = pick_level(c, roi);
        const int b = (int)roi[0];
        const bool ok = l >= 0 && l < c.n_levels && b >= 0 && b < c.batch;
        if (l < 0 || l >= c.n_levels) l = 0;
        const int H = c.height[l], W = c.width[l];
        RtTap *t = taps + (size_t)i * (PH + PW);
        int y0 = 1 << 30, y1 = -1, x0 = 1 << 30, x1 = -1;
        for (int p = 0; p < PH; ++p) {
            const Tap v = (c.mode == DETBOX_ROI_BIN_CENTRE) ? tf_tap(roi[2], roi[4], c.spatial_scale[l], PH, H, p) : crop_tap(roi[1], roi[3], PH, H, p);
            RtTap o; o.i0 = (short)(v.valid ? v.i0 : -1); o.i1 = (short)v.i1; o.w = v.w;
            t[p] = o;
            if (v.valid) { y0 = min(y0, v.i0); y1 = max(y1, v.i1); }
        }
        for (int p = 0; p < PW; ++p) {
            const Tap v = (c.mode == DETBOX_ROI_BIN_CENTRE) ? tf_tap(roi[1], roi[3], c.spatial_scale[l], PW, W, p) : crop_tap(roi[2], roi[4], PW, W, p);
            RtTap o; o.i0 = (short)(v.valid ? v.i0 : -1); o.i1 = (short)v.i1; o.w = v.w;
            t[PH + p] = o;
            if (v.valid) { x0 = min(x0, v.i0); x1 = max(x1, v.i1); }
        }
        const bool none = !ok || y1 < 0 || x1 < 0;
        hdr[i] = none ? make_int4(-1, 0, 0, 0) : make_int4((l << 24) | b, y0 | (y1 << 16), x0 | (x1 << 16), 0);
    }
}

__global__ void __launch_bounds__(RT_THREADS, 3)
roi_bwd_nchw_tile_kernel(const RoiArgs a, const int4 *__restrict__ hdr, const RtTap *__restrict__ taps, const TileGeom geo, const int nb_max)
{
    extern __shared__ __align__(16) float s_dynf[];               // accumulators [256 pixels][33], then grad_out [nb][RT_CH][bins]
    float *s_acc = s_dynf, *s_g = s_dynf + RT_THREADS * RT_APITCH;
    // weight table of a listed RoI against this tile: s_rw[slot][row][bin row] and, per row, the bit mask of the bin rows with a
    // non-zero weight (the loops below walk set bits, not table entries); the bin columns are walked tap by tap (s_ct).  Slots
    // 0 .. RT_NC-1 hold the first RT_NC RoIs of the list for the whole round (built once, used by every channel chunk);
    // later RoIs are rebuilt per chunk in slots RT_NC .. RT_NC+RT_NBMAX-1.
    __shared__ float s_rw[RT_NC + RT_NBMAX][RT_TH][RT_PMAX];
    __shared__ unsigned char s_rm[RT_NC + RT_NBMAX][RT_TH];
    __shared__ RtTap s_ct[RT_NC + RT_NBMAX][RT_PMAX];           // the bin columns' taps, pixel indices relative to the tile (i0 = RT_NOTAP: none)
    __shared__ __align__(8) unsigned long long s_bar;            // completion of the bulk copies that stage grad_out
    __shared__ int s_list[RT_CAP];
    __shared__ __align__(16) unsigned short s_wc[RT_J][RT_THREADS / 32];
    __shared__ __align__(16) unsigned short s_wany[RT_THREADS / 32];
    const detbox_roi_cfg &c = a.cfg;
    const int PH = c.pooled_h, PW = c.pooled_w, C = c.channels, bins = PH * PW;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    int l = 0;
#pragma unroll
    for (int k = 1; k < DETBOX_MAX_LEVELS; ++k) if (k < c.n_levels && (int)blockIdx.x >= geo.off[k]) l = k;
    const int rem = (int)blockIdx.x - geo.off[l], per = geo.nty[l] * geo.ntx[l];
    const int b = rem / per, t2 = rem - b * per, tyi = t2 / geo.ntx[l], txi = t2 - tyi * geo.ntx[l];
    const int H = c.height[l], W = c.width[l];
    const int ty0 = tyi * RT_TH, tx0 = txi * RT_TW, ty1 = min(ty0 + RT_TH, H) - 1, tx1 = min(tx0 + RT_TW, W) - 1;
    const int y = ty0 + w, x = tx0 + lane;
    const bool inside = y < H && x < W;
    const size_t plane = (size_t)H * W;
    float *gbase = c.grad_feat[l] + (size_t)b * C * plane + (size_t)y * W + x;
    const int key = (l << 24) | b;
    // grad_out rows of a (RoI, chunk) are contiguous; 16-byte copies need every such block aligned and a whole number of pieces
    const bool vec16 = (((size_t)a.grad_out & 15) == 0) && (((size_t)C * bins) & 3) == 0 && ((bins * RT_CH) & 3) == 0 && (C % RT_CH == 0 || ((bins * (C % RT_CH)) & 3) == 0);
    const bool vecst = (W & 3) == 0 && (plane & 3) == 0 && (((size_t)c.grad_feat[l]) & 15) == 0;    // whole 16-byte quads
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"((unsigned)__cvta_generic_to_shared(&s_bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    unsigned phase = 0;                                           // (the init is published by the barrier at the top of the first round)
    long base = 0;
    int round = 0;
    do {
        // ---- the RoIs meeting this tile, in index order (at most RT_CAP per round; a round ends before the 256 records that
        //      would overflow the list, so rounds after the first — thousands of RoIs on one tile — accumulate into the stored tile)
        // RT_J records per thread are tested with their loads in flight together (one L2 latency, one barrier), then compacted in
        // index order: record i = base + j * 256 + tid, so the order is j, then warp, then lane.
        unsigned hitmask = 0, wany = 0;
#pragma unroll
        for (int j0 = 0; j0 < RT_J; j0 += 8) {
            int4 h[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const long i = base + (long)(j0 + j) * RT_THREADS + tid;
                h[j] = i < a.R ? __ldg(&hdr[i]) : make_int4(-1, 0, 0, 0);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const bool hit = h[j].x == key && (h[j].y & 0xffff) <= ty1 && (h[j].y >> 16) >= ty0 && (h[j].z & 0xffff) <= tx1 && (h[j].z >> 16) >= tx0;
                const unsigned ball = __ballot_sync(0xffffffffu, hit);
                if (lane == 0) s_wc[j0 + j][w] = (unsigned short)__popc(ball);
                hitmask |= hit ? 1u << (j0 + j) : 0u;
                wany |= ball ? 1u << (j0 + j) : 0u;
            }
        }
        if (lane == 0) s_wany[w] = (unsigned short)wany;
        __syncthreads();
        // the record groups j in which ANY warp had a hit (typically one to three of the 16): only those are compacted
        const uint4 wa = *reinterpret_cast<const uint4 *>(s_wany);
        unsigned jmask = wa.x | wa.y | wa.z | wa.w;
        jmask = (jmask | (jmask >> 16)) & 0xffffu;
        int n = 0, J = RT_J;
        for (; jmask; jmask &= jmask - 1) {
            const int j = __ffs((int)jmask) - 1;
            const uint4 wc = *reinterpret_cast<const uint4 *>(&s_wc[j][0]);      // eight 16-bit counts
            const unsigned cnt8[8] = {wc.x & 0xffffu, wc.x >> 16, wc.y & 0xffffu, wc.y >> 16, wc.z & 0xffffu, wc.z >> 16, wc.w & 0xffffu, wc.w >> 16};
            int tot = 0, off = 0;
#pragma unroll
            for (int ww = 0; ww < RT_THREADS / 32; ++ww) { off += ww < w ? (int)cnt8[ww] : 0; tot += (int)cnt8[ww]; }
            if (n + tot > RT_CAP) { J = j; break; }               // block-uniform; tot <= 256 = RT_CAP, so J >= 1
            const bool hit = (hitmask >> j) & 1u;
            const unsigned ball = __ballot_sync(0xffffffffu, hit);
            if (hit) s_list[n + off + __popc(ball & ((1u << lane) - 1u))] = (int)(base + (long)j * RT_THREADS + tid);
            n += tot;
        }
        const long next = min(a.R, base + (long)J * RT_THREADS);
        __syncthreads();
        // weight tables of list entries [e_lo, e_hi) into slots slot0 ..  (all threads; followed by a barrier at the caller)
        auto build_tables = [&](const int e_lo, const int e_hi, const int slot0) {
            const int per = RT_TH * PH + PW;
            for (int idx = tid; idx < (e_hi - e_lo) * per; idx += RT_THREADS) {
                const int k = idx / per, q = idx - k * per;
                const RtTap *tp = taps + (size_t)s_list[e_lo + k] * (PH + PW);
                if (q < RT_TH * PH) {
                    const int yy = q / PH, py = q - yy * PH;
                    const RtTap t = tp[py];
                    const int yg = ty0 + yy;
                    s_rw[slot0 + k][yy][py] = t.i0 < 0 ? 0.0f : (t.i0 == yg ? 1 - t.w : 0.0f) + ((t.i1 == yg && t.i1 != t.i0) ? t.w : 0.0f);
                } else {
                    const int px = q - RT_TH * PH;
                    RtTap t = tp[PH + px];
                    if (t.i0 < 0) t.i0 = (short)RT_NOTAP; else { t.i0 = (short)(t.i0 - tx0); t.i1 = (short)(t.i1 - tx0); }
                    s_ct[slot0 + k][px] = t;
                }
            }
            __syncthreads();
            for (int idx = tid; idx < (e_hi - e_lo) * RT_TH; idx += RT_THREADS) {
                const int k = idx / RT_TH, q = idx - k * RT_TH;
                unsigned m = 0;
                for (int py = 0; py < PH; ++py) m |= s_rw[slot0 + k][q][py] != 0.0f ? 1u << py : 0u;
                s_rm[slot0 + k][q] = (unsigned char)m;
            }
        };
        if (n == 0 && round == 0 && next >= a.R) {
            // no RoI meets this tile (the memset's share of the work): zeros straight from registers, no accumulator pass
            if (vecst) {
                const int chs = lane >> 3, x4 = (lane & 7) * 4;
                if (y < H && tx0 + x4 < W) {
                    float *gq = c.grad_feat[l] + (size_t)b * C * plane + (size_t)y * W + tx0 + x4;
                    for (int k = chs; k < C; k += 4) __stcs(reinterpret_cast<float4 *>(gq + (size_t)k * plane), make_float4(0, 0, 0, 0));
                }
            } else if (inside) {
                for (int k = 0; k < C; ++k) __stcs(gbase + (size_t)k * plane, 0.0f);
            }
            return;
        }
        build_tables(0, min(n, RT_NC), 0);
        // ---- channel chunk by channel chunk.  Lanes are CHANNELS here: a warp owns one tile row and walks the few pixels of
        //      it a listed RoI touches; the weights of a pixel are warp-uniform (the tables above), every lane sums its
        //      channel's terms colw * (roww * g) (crop_and_resize.c:258-264) and adds them to the pixel's accumulator in
        //      shared memory ([pixel][33]: conflict-free for lane = channel here and for lane = pixel in the store below).
        //      (First version: lanes = pixels, accumulators in registers — 12.7 of 32 lanes active on pedestrian-shaped RoIs
        //      4-5 pixels wide, 2.0 G warp instructions, 4.1 ms; second: dense weight tables scanned per pixel, 0.85 G, 1.7 ms.)
        // grad_out of list entries [e0, e0 + nb) for the chunk at c0 -> s_g: one bulk copy (TMA, UBLKCP) per RoI, its chunk is contiguous
        auto issue = [&](const int c0x, const int e0x) {
            if (tid != 0) return;
            const int bytes = min(RT_CH, C - c0x) * bins * 4, nbx = min(nb_max, n - e0x);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");         // earlier reads of s_g before the async writes
            const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"((unsigned)(nbx * bytes)) : "memory");
            for (int k = 0; k < nbx; ++k) {
                const float *src = a.grad_out + ((size_t)s_list[e0x + k] * C + c0x) * bins;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             :: "r"((unsigned)__cvta_generic_to_shared(s_g + k * (RT_CH * bins))), "l"(src), "r"((unsigned)bytes), "r"(bar) : "memory");
            }
        };
        bool pre = false;                                         // the first batch of the coming chunk is already in flight
        for (int c0 = 0; c0 < C; c0 += RT_CH) {
            const int nch = min(RT_CH, C - c0);
            __syncthreads();                                      // s_acc of the previous chunk has been stored; tables are built
            for (int q = tid; q < RT_THREADS * RT_APITCH / 4; q += RT_THREADS) reinterpret_cast<float4 *>(s_acc)[q] = make_float4(0, 0, 0, 0);
            for (int e0 = 0; e0 < n; e0 += nb_max) {
                const int nb = min(nb_max, n - e0);
                const int slot0 = e0 < RT_NC ? e0 : RT_NC;        // (RT_NC is a multiple of the batch size)
                __syncthreads();                                  // the previous batch has been consumed (and s_acc is zero)
                const int per_roi = nch * bins;
                if (vec16) {
                    if (!(pre && e0 == 0)) issue(c0, e0);
                    pre = false;
                } else {
                    for (int k = 0; k < nb; ++k) {
                        const float *src = a.grad_out + ((size_t)s_list[e0 + k] * C + c0) * bins;
                        float *dst = s_g + k * (RT_CH * bins);
                        for (int q = tid; q < per_roi; q += RT_THREADS) dst[q] = __ldg(src + q);
                    }
                }
                if (e0 >= RT_NC) build_tables(e0, e0 + nb, RT_NC);
                if (vec16) {
                    const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar);
                    unsigned done;
                    do {
                        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                                     : "=r"(done) : "r"(bar), "r"(phase) : "memory");
                    } while (!done);
                    phase ^= 1u;
                }
                __syncthreads();
                for (int k = 0; k < nb; ++k) {
                    const int sl = slot0 + k;
                    const unsigned rm = s_rm[sl][w];
                    if (rm == 0u || lane >= nch) continue;        // (the first test is warp-uniform)
                    const float *rwp = s_rw[sl][w];
                    const float *gk = s_g + k * (RT_CH * bins) + lane * bins;
                    // separable, like roi_bwd_nchw_sep_kernel: T[bx] = sum over the bin rows landing on this feature row of
                    // roww * g[by][bx] once per (row, RoI, channel), then colw * T per pixel — the same weights, one rounding apart
                    // from the exact-term sum (inside the bound the tests derive)
                    float T[RT_PMAX];
#pragma unroll
                    for (int px = 0; px < RT_PMAX; ++px) T[px] = 0.0f;
                    for (unsigned mr = rm; mr; mr &= mr - 1) {
                        const int py = __ffs((int)mr) - 1;
                        const float rw = rwp[py];
                        const float *gp = gk + py * PW;
#pragma unroll
                        for (int px = 0; px < RT_PMAX; ++px) if (px < PW) T[px] += rw * gp[px];
                    }
                    // bin columns left to right: a bin adds (1 - w) * T to pixel i0 and w * T to pixel i1 = i0 + 1; a0 / a1 collect the
                    // sums of pixels pc and pc + 1 and go to the accumulator when the walk leaves them (all of it warp-uniform)
                    float *arow = s_acc + (w * RT_TW) * RT_APITCH + lane;
                    int pc = RT_NOTAP; float a0 = 0.0f, a1 = 0.0f;
                    auto flush = [&](const int p, const float v) { if (p >= 0 && p < RT_TW) arow[p * RT_APITCH] += v; };
#pragma unroll
                    for (int px = 0; px < RT_PMAX; ++px) {
                        if (px >= PW) continue;
                        const RtTap t = s_ct[sl][px];
                        if (t.i0 == RT_NOTAP) continue;
                        if (t.i0 != pc) {
                            flush(pc, a0);
                            if (t.i0 == pc + 1) { a0 = a1; }
                            else { flush(pc + 1, a1); a0 = 0.0f; }
                            a1 = 0.0f;
                            pc = t.i0;
                        }
                        a0 += (1 - t.w) * T[px];
                        if (t.i1 != t.i0) a1 += t.w * T[px];
                    }
                    flush(pc, a0);
                    flush(pc + 1, a1);
                }
            }
            __syncthreads();
            // the next chunk's first batch is requested now: its L2 latency passes behind this chunk's store and the zeroing
            if (vec16 && n > 0 && c0 + RT_CH < C) { issue(c0 + RT_CH, 0); pre = true; }
            if (round == 0 && vecst) {
                // four pixels x one channel per store: lane = (channel offset 0..3, pixel quad 0..7); the four shared-memory reads
                // of a quad hit banks (x + ch) % 32, distinct across the warp
                const int chs = lane >> 3, x4 = (lane & 7) * 4;
                if (y < H && tx0 + x4 < W) {
                    const float *ap = s_acc + (w * RT_TW + x4) * RT_APITCH;
                    float *gq = c.grad_feat[l] + (size_t)b * C * plane + (size_t)y * W + tx0 + x4;
                    for (int k = chs; k < nch; k += 4)
                        __stcs(reinterpret_cast<float4 *>(gq + (size_t)(c0 + k) * plane),
                               make_float4(ap[k], ap[RT_APITCH + k], ap[2 * RT_APITCH + k], ap[3 * RT_APITCH + k]));
                }
            } else if (inside) {
                const float *ap = s_acc + (w * RT_TW + lane) * RT_APITCH;
                if (round > 0) { for (int k = 0; k < nch; ++k) gbase[(size_t)(c0 + k) * plane] += ap[k]; }
                else { for (int k = 0; k < nch; ++k) __stcs(gbase + (size_t)(c0 + k) * plane, ap[k]); }
            }
        }
        base = next;
        ++round;
    } while (base < a.R);
}

__global__ void __launch_bounds__(256)
assign_levels_kernel(const float *__restrict__ boxes, long R, detbox_roi_cfg c, int32_t *__restrict__ levels)
{
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < R; i += (long)gridDim.x * blockDim.x) {
        const float *b = boxes + i * 4;
        const float roi[5] = {0.0f, b[0], b[1], b[2], b[3]};
        levels[i] = pick_level(c, roi);
    }
}

int check_cfg(const detbox_roi_cfg *c, bool backward)
{
    if (!c || c->n_levels < 1 || c->n_levels > DETBOX_MAX_LEVELS) return DETBOX_ERR_INVALID;
    if (c->batch < 0 || c->channels <= 0 || c->pooled_h <= 0 || c->pooled_w <= 0) return DETBOX_ERR_INVALID;
    if (c->mode != DETBOX_ROI_BIN_CENTRE && c->mode != DETBOX_ROI_GRID_POINT && c->mode != DETBOX_ROI_CROP_NORM) return DETBOX_ERR_INVALID;
    if (c->pool < 0 || c->pool > 2 || (c->pool != 0 && c->mode != DETBOX_ROI_GRID_POINT)) return DETBOX_ERR_INVALID;
    if (c->pooled_h + 1 > RA_MAXP || c->pooled_w + 1 > RA_MAXP) return DETBOX_ERR_UNSUPPORTED;
    if ((c->layout != 0 && c->layout != 1) || (c->out_layout != 0 && c->out_layout != 1)) return DETBOX_ERR_INVALID;
    for (int l = 0; l < c->n_levels; ++l) {
        if (c->height[l] <= 0 || c->width[l] <= 0) return DETBOX_ERR_INVALID;
        if (backward ? !c->grad_feat[l] : !c->feat[l]) return DETBOX_ERR_INVALID;
        if (backward && c->pool == DETBOX_POOL_MAX2 && !c->feat[l]) return DETBOX_ERR_INVALID;   // arg-max needs the features
    }
    return DETBOX_OK;
}

template <typename F> int dispatch(const detbox_roi_cfg &c, F &&f)
{
    const bool grid = c.mode == DETBOX_ROI_GRID_POINT, in = c.layout == DETBOX_LAYOUT_NHWC, out = c.out_layout == DETBOX_LAYOUT_NHWC;
    const int k = (grid ? 4 : 0) | (in ? 2 : 0) | (out ? 1 : 0);
    switch (k) {
        case 0: return f(std::integral_constant<int, 0>());
        case 1: return f(std::integral_constant<int, 1>());
        case 2: return f(std::integral_constant<int, 2>());
        case 3: return f(std::integral_constant<int, 3>());
        case 4: return f(std::integral_constant<int, 4>());
        case 5: return f(std::integral_constant<int, 5>());
        case 6: return f(std::integral_constant<int, 6>());
        default: return f(std::integral_constant<int, 7>());
    }
}

}  // namespace

extern "C" int detbox_roi_align_forward(const detbox_roi_cfg *cfg, const float *rois, long R,
                                        const int32_t *levels, float *out, void *stream)
{
    int rc = check_cfg(cfg, false);
    if (rc != DETBOX_OK) return rc;
    if (R < 0) return DETBOX_ERR_INVALID;
    if (R == 0) return DETBOX_OK;
    if (!rois || !out) return DETBOX_ERR_INVALID;
    RoiArgs a; a.cfg = *cfg; a.rois = rois; a.R = R; a.levels = levels; a.out = out; a.grad_out = nullptr;
    dim3 grid((unsigned)R, cfg->all_levels ? cfg->n_levels : 1);
    cudaStream_t st = as_stream(stream);
    static const bool no_staged = getenv("DETBOX_ROI_NOSTAGED") != nullptr;
    if (cfg->layout == DETBOX_LAYOUT_NCHW && cfg->out_layout == DETBOX_LAYOUT_NCHW && !no_staged) {
        DETBOX_TRACE_BEGIN(DETBOX_K_ROI_FWD, st);
        if (cfg->mode == DETBOX_ROI_GRID_POINT) roi_fwd_nchw_staged_kernel<true><<<grid, RS_THREADS, RS_FLOATS * sizeof(float), st>>>(a);
        else roi_fwd_nchw_staged_kernel<false><<<grid, RS_THREADS, RS_FLOATS * sizeof(float), st>>>(a);
        DETBOX_TRACE_END(DETBOX_K_ROI_FWD, st);
        DETBOX_LAUNCH_CHECK();
        return DETBOX_OK;
    }
    return dispatch(*cfg, [&](auto K) -> int {
        constexpr int k = decltype(K)::value;
        DETBOX_TRACE_BEGIN(DETBOX_K_ROI_FWD, st);
        static const int unroll = getenv("DETBOX_ROI_UNROLL") ? atoi(getenv("DETBOX_ROI_UNROLL")) : 1;
        if ((k & 3) == 3 && unroll == 4) roi_fwd_kernel<(k & 4) != 0, (k & 2) != 0, (k & 1) != 0, 4><<<grid, RA_THREADS, 0, st>>>(a);
        else if ((k & 3) == 3 && unroll == 2) roi_fwd_kernel<(k & 4) != 0, (k & 2) != 0, (k & 1) != 0, 2><<<grid, RA_THREADS, 0, st>>>(a);
        else roi_fwd_kernel<(k & 4) != 0, (k & 2) != 0, (k & 1) != 0, 1><<<grid, RA_THREADS, 0, st>>>(a);
        DETBOX_TRACE_END(DETBOX_K_ROI_FWD, st);
        DETBOX_LAUNCH_CHECK();
        return DETBOX_OK;
    });
}

extern "C" int detbox_roi_align_backward(const detbox_roi_cfg *cfg, const float *rois, long R,
                                         const int32_t *levels, const float *grad_out, int zero_grad, void *stream)
{
    int rc = check_cfg(cfg, true);
    if (rc != DETBOX_OK) return rc;
    if (R < 0) return DETBOX_ERR_INVALID;
    cudaStream_t st = as_stream(stream);
    if (zero_grad) {
        DETBOX_TRACE_BEGIN(DETBOX_K_ROI_ZERO, st);
        for (int l = 0; l < cfg->n_levels; ++l)
            DETBOX_CUDA_TRY(cudaMemsetAsync(cfg->grad_feat[l], 0,
                                            (size_t)cfg->batch * cfg->channels * cfg->height[l] * cfg->width[l] * sizeof(float), st));
        DETBOX_TRACE_END(DETBOX_K_ROI_ZERO, st);
    }
    if (R == 0) return DETBOX_OK;
    if (!rois || !grad_out) return DETBOX_ERR_INVALID;
    RoiArgs a; a.cfg = *cfg; a.rois = rois; a.R = R; a.levels = levels; a.out = nullptr; a.grad_out = grad_out;
    dim3 grid((unsigned)R, cfg->all_levels ? cfg->n_levels : 1);
    // Measured on B200 at BASELINE configs[2] (profiles/r02_roi_staged_vs_generic.txt): the staged backward scatters in
    // 0.94 ms against 0.73 ms for the direct red.global.add path (shared-memory float atomics on taps that neighbouring
    // bins share, plus a zero and a flush pass per chunk), so it is opt-in; the staged FORWARD is the default (0.24 vs 0.37 ms).
    static const bool bwd_direct = getenv("DETBOX_ROI_BWD_DIRECT") != nullptr;
    if (cfg->layout == DETBOX_LAYOUT_NCHW && cfg->out_layout == DETBOX_LAYOUT_NCHW && cfg->pool == DETBOX_POOL_NONE && !bwd_direct &&
        getenv("DETBOX_ROI_BWD_STAGED") == nullptr && (long)cfg->pooled_h * cfg->pooled_w <= RG_GFLOATS) {
        DETBOX_TRACE_BEGIN(DETBOX_K_ROI_BWD, st);
        static const bool no_sep = getenv("DETBOX_ROI_BWD_GATHER") != nullptr;     // the exact-term gather kernel, for A/B timing and its own test
        if (cfg->mode == DETBOX_ROI_GRID_POINT) roi_bwd_nchw_gather_kernel<true><<<grid, RG_THREADS, RG_GFLOATS * sizeof(float), st>>>(a);
        else if (cfg->pooled_w <= RS2_PW && !no_sep) roi_bwd_nchw_sep_kernel<<<grid, RG_THREADS, RG_GFLOATS * sizeof(float), st>>>(a);
        else roi_bwd_nchw_gather_kernel<false><<<grid, RG_THREADS, RG_GFLOATS * sizeof(float), st>>>(a);
        DETBOX_TRACE_END(DETBOX_K_ROI_BWD, st);
        DETBOX_LAUNCH_CHECK();
        return DETBOX_OK;
    }
    static const bool bwd_staged = getenv("DETBOX_ROI_BWD_STAGED") != nullptr;
    if (cfg->layout == DETBOX_LAYOUT_NCHW && cfg->out_layout == DETBOX_LAYOUT_NCHW && cfg->pool == DETBOX_POOL_NONE && bwd_staged) {
        DETBOX_TRACE_BEGIN(DETBOX_K_ROI_BWD, st);
        if (cfg->mode == DETBOX_ROI_GRID_POINT) roi_bwd_nchw_staged_kernel<true><<<grid, RS_THREADS, RS_FLOATS * sizeof(float), st>>>(a);
        else roi_bwd_nchw_staged_kernel<false><<<grid, RS_THREADS, RS_FLOATS * sizeof(float), st>>>(a);
        DETBOX_TRACE_END(DETBOX_K_ROI_BWD, st);
        DETBOX_LAUNCH_CHECK();
        return DETBOX_OK;
    }
    return dispatch(*cfg, [&](auto K) -> int {
        constexpr int k = decltype(K)::value;
        DETBOX_TRACE_BEGIN(DETBOX_K_ROI_BWD, st);
        roi_bwd_kernel<(k & 4) != 0, (k & 2) != 0, (k & 1) != 0><<<grid, RA_THREADS, 0, st>>>(a);
        DETBOX_TRACE_END(DETBOX_K_ROI_BWD, st);
        DETBOX_LAUNCH_CHECK();
        return DETBOX_OK;
    });
}

static bool tiled_supported(const detbox_roi_cfg *c)
{
    if (c->layout != DETBOX_LAYOUT_NCHW || c->out_layout != DETBOX_LAYOUT_NCHW || c->pool != DETBOX_POOL_NONE || c->all_levels) return false;
    if (c->mode != DETBOX_ROI_BIN_CENTRE && c->mode != DETBOX_ROI_CROP_NORM) return false;
    if (c->pooled_h > RT_PMAX || c->pooled_w > RT_PMAX) return false;
    for (int l = 0; l < c->n_levels; ++l) if (c->height[l] > 32767 || c->width[l] > 32767) return false;
    return c->batch < (1 << 24);
}

extern "C" size_t detbox_roi_align_backward_tiled_workspace_bytes(const detbox_roi_cfg *cfg, long R)
{
    if (!cfg || R < 0 || cfg->pooled_h <= 0 || cfg->pooled_w <= 0) return 0;
    const size_t n = (size_t)(R > 0 ? R : 1);
    return align_up(n * sizeof(int4)) + align_up(n * (size_t)(cfg->pooled_h + cfg->pooled_w) * sizeof(RtTap));
}

extern "C" int detbox_roi_align_backward_tiled(const detbox_roi_cfg *cfg, const float *rois, long R, const int32_t *levels,
                                               const float *grad_out, void *ws, size_t ws_bytes, void *stream)
{
    int rc = check_cfg(cfg, true);
    if (rc != DETBOX_OK) return rc;
    if (R < 0) return DETBOX_ERR_INVALID;
    if (!tiled_supported(cfg)) return DETBOX_ERR_UNSUPPORTED;
    if (R > 0 && (!rois || !grad_out)) return DETBOX_ERR_INVALID;
    if (!ws || ws_bytes < detbox_roi_align_backward_tiled_workspace_bytes(cfg, R)) return DETBOX_ERR_WORKSPACE;
    cudaStream_t st = as_stream(stream);
    Carver cv(ws);
    int4 *hdr = cv.take<int4>((size_t)(R > 0 ? R : 1));
    RtTap *taps = cv.take<RtTap>((size_t)(R > 0 ? R : 1) * (size_t)(cfg->pooled_h + cfg->pooled_w));
    RoiArgs a; a.cfg = *cfg; a.rois = rois; a.R = R; a.levels = levels; a.out = nullptr; a.grad_out = grad_out;
    TileGeom geo;
    long off = 0;
    for (int l = 0; l < DETBOX_MAX_LEVELS; ++l) {
        geo.off[l] = (int)off; geo.nty[l] = geo.ntx[l] = 1;
        if (l >= cfg->n_levels) continue;
        geo.nty[l] = (cfg->height[l] + RT_TH - 1) / RT_TH; geo.ntx[l] = (cfg->width[l] + RT_TW - 1) / RT_TW;
        off += (long)cfg->batch * geo.nty[l] * geo.ntx[l];
        if (off >= (1l << 31)) return DETBOX_ERR_UNSUPPORTED;
    }
    geo.off[DETBOX_MAX_LEVELS] = (int)off;
    for (int l = cfg->n_levels; l < DETBOX_MAX_LEVELS; ++l) geo.off[l] = (int)off;
    if (off == 0) return DETBOX_OK;
    DETBOX_TRACE_BEGIN(DETBOX_K_ROI_BWD, st);
    if (R > 0) {
        roi_tile_prep_kernel<<<(unsigned)std::min<long>((R + 255) / 256, kNumSMs * 4), 256, 0, st>>>(a, hdr, taps);
        DETBOX_LAUNCH_CHECK();
    }
    const int bins = cfg->pooled_h * cfg->pooled_w;
    const int nb = RT_NBMAX;
    const size_t smem = ((size_t)RT_THREADS * RT_APITCH + (size_t)nb * RT_CH * bins) * sizeof(float);     // 33 + 25 KB at 7 x 7
    {
        static bool ready[64] = {};
        int dev = 0;
        DETBOX_CUDA_TRY(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !ready[dev]) {
            DETBOX_CUDA_TRY(cudaFuncSetAttribute(roi_bwd_nchw_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)(((size_t)RT_THREADS * RT_APITCH + (size_t)RT_NBMAX * RT_CH * RT_PMAX * RT_PMAX) * sizeof(float))));
            if (dev >= 0 && dev < 64) ready[dev] = true;
        }
    }
    roi_bwd_nchw_tile_kernel<<<(unsigned)off, RT_THREADS, smem, st>>>(a, hdr, taps, geo, nb);
    DETBOX_TRACE_END(DETBOX_K_ROI_BWD, st);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

// ---- L2-friendly RoI processing order ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
roi_order_key_kernel(const float *__restrict__ rois, long R, const int32_t *__restrict__ levels, detbox_roi_cfg c,
                     float im_h, float im_w, float *__restrict__ keys)
{
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < R; i += (long)gridDim.x * blockDim.x) {
        const float *roi = rois + i * 5;
        int l = 0;
        if (levels) l = levels[i];
        else if (c.n_levels > 1) l = pick_level(c, roi);
        l = min(max(l, 0), DETBOX_MAX_LEVELS - 1);
        const int b = min(max((int)roi[0], 0), 255);
        const float cy = 0.5f * (roi[2] + roi[4]), cx = 0.5f * (roi[1] + roi[3]);
        const int yb = min(max((int)(cy / im_h * 64.0f), 0), 63), xb = min(max((int)(cx / im_w * 32.0f), 0), 31);
        // 3 + 8 + 6 + 5 = 22 bits: an exact float32 integer.  detbox_sort_desc sorts descending and stable: any monotone key does.
        keys[i] = (float)((((l << 8) | b) << 11) | (yb << 5) | xb);
    }
}

extern "C" size_t detbox_roi_locality_order_workspace_bytes(long R)
{
    if (R < 0) return 0;
    return align_up((size_t)(R > 0 ? R : 1) * sizeof(float)) + detbox_sort_desc_workspace_bytes(R);
}

extern "C" int detbox_roi_locality_order(const float *rois, long R, const int32_t *levels, const float *ref_areas, int n_levels,
                                         float im_h, float im_w, int32_t *order, void *ws, size_t ws_bytes, void *stream)
{
    if (R < 0 || n_levels < 1 || n_levels > DETBOX_MAX_LEVELS || !(im_h > 0) || !(im_w > 0)) return DETBOX_ERR_INVALID;
    if (R == 0) return DETBOX_OK;
    if (!rois || !order || (!levels && n_levels > 1 && !ref_areas)) return DETBOX_ERR_INVALID;
    if (!ws || ws_bytes < detbox_roi_locality_order_workspace_bytes(R)) return DETBOX_ERR_WORKSPACE;
    detbox_roi_cfg c{};
    c.n_levels = n_levels;
    for (int l = 0; l < n_levels; ++l) c.ref_area[l] = ref_areas ? ref_areas[l] : 1.0f;
    float *keys = reinterpret_cast<float *>(ws);
    const size_t koff = align_up((size_t)R * sizeof(float));
    long gb = (R + 255) / 256; if (gb > kNumSMs * 8) gb = kNumSMs * 8;
    roi_order_key_kernel<<<(unsigned)gb, 256, 0, as_stream(stream)>>>(rois, R, levels, c, im_h, im_w, keys);
    DETBOX_LAUNCH_CHECK();
    return detbox_sort_desc(keys, R, order, reinterpret_cast<char *>(ws) + koff, ws_bytes - koff, stream);
}

extern "C" int detbox_assign_levels(const float *boxes, long R, const float *ref_areas, int n_levels,
                                    int32_t *levels, void *stream)
{
    if (R < 0 || n_levels < 1 || n_levels > DETBOX_MAX_LEVELS || !ref_areas) return DETBOX_ERR_INVALID;
    if (R == 0) return DETBOX_OK;
    if (!boxes || !levels) return DETBOX_ERR_INVALID;
    detbox_roi_cfg c{};
    c.n_levels = n_levels;
    for (int l = 0; l < n_levels; ++l) c.ref_area[l] = ref_areas[l];
    long gb = (R + 255) / 256; if (gb > kNumSMs * 8) gb = kNumSMs * 8;
    assign_levels_kernel<<<(unsigned)gb, 256, 0, as_stream(stream)>>>(boxes, R, c, levels);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}
