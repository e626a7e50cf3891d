// anchor_target.cu — fused anchor-target assignment (reference libs/layers/anchor.py:17-145,
// 'simple' sampling; batch wrapper libs/layers/data_layer.py:228-239).
//
// The reference materialises the [A,G] float64 overlap matrix (25-207 MB per image) and makes
// three NumPy passes over it.  Here nothing of size A*G ever exists:
//
//   pass 0  anchor_bbox_kernel (once per call, A anchors, not B*A): bounding box of every 2048-anchor tile
//           and of every 32-anchor warp slice.  They depend on the anchors only, so all images share them.
//   pass 1  colmax_kernel: per-GT max IoU over all anchors as order-preserving u64 bit patterns.  GT-centric:
//           slice boxes are tested against the staged GT, slices some GT may touch go on a work list, one
//           warp per listed slice computes the exact float64 IoUs; warp max, then one atomicMax.
//   pass 2  anchor_target_kernel.  Decide phase, only for slices a GT may touch: the anchor's row (first-max
//           argmax, thresholds, per-GT promotion ov == colmax incl. all-zero columns, clobber, ignore
//           regions) -> (label, GT row, flag) in shared memory.  Emit phase, every anchor: label, float32
//           regression targets against the chosen GT (GT 0 for an all-zero row, like np.argmax), inside
//           weights: 36 B per anchor; fg/bg counts by block reduction.
//   pass 3  label_weight_kernel: label_weights from the per-image counts (needs the global fg/bg totals).
//
// HBM bytes per anchor: 36 written in pass 2 (labels 4 + targets 16 + inside weights 16),
// 4 written + 4 read in pass 3; the 16-byte anchor rows are shared by the whole batch and stay
// L2-resident.  No A*G traffic.
//
// Culling keeps the float64 work off the critical path: exact float64 IoU (reference bbox.pyx:33-54, same
// operation order, --fmad=false) is evaluated only for (slice, GT) pairs whose boxes may intersect, so labels
// are bit-exact while most pairs cost one float32 predicate.
//
// Monotonicity makes the culling exact: min/max are exact, and a rounded (x - y + 1) is a
// monotone function of its operands, so iw(anchor, gt) <= iw(bbox, gt) also after rounding.
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <atomic>
#include <thread>
#include <vector>
#if defined(__x86_64__)
#include <immintrin.h>
#endif
#include "box_math.cuh"

using namespace detbox;

namespace {

constexpr int AT_THREADS = 256;
constexpr int AT_ITEMS = 8;                         // anchors per thread
constexpr int AT_TILE = AT_THREADS * AT_ITEMS;      // anchors per block
constexpr int AT_WARPS = AT_THREADS / 32;
constexpr int GCH = 128;                            // GT boxes staged per chunk

struct TileBoxes { float4 tile; float4 warp[AT_ITEMS][AT_WARPS]; };   // (min x1, min y1, max x2, max y2)

struct GtChunk {
    double x1[GCH], y1[GCH], x2[GCH], y2[GCH], area[GCH];
    unsigned long long colmax[GCH];
    float4 fbox[GCH];                                // the same boxes as float32, for the conservative slice test
    float4 genc[GCH];                                // box_math.cuh::gt_pre — the GT half of the encode, once per GT
    float cls[GCH];
    float area32[GCH];                               // float32 screen of the decide phase: GT area ...
    float cmlo[GCH];                                 // ... and its column maximum lowered by the screen's margin (-1: no screen for this GT)
};

// ---- float32 screen of the decide phase ---------------------------------------------------------------------------
// The exact decide (float64 IoU in the reference's operation order, one IEEE division per overlapping pair) is needed
// only where a decision can depend on the last bits.  Everything else is decided by float32 values with a PROVEN bound:
// inputs are float32, so fl(m - M) is one rounding of the exact difference and s = fl(fl(m - M) + 1) satisfies
// |s - exact| <= (2|s| + 1) * 2^-23: for |s| >= SCR_T = 2^-10 its sign is certain and, if positive, its relative error
// is <= (2 + 1/s) * 2^-24 <= 6.2e-5.  With both intersection sides and all four box sides >= SCR_T the float32 IoU
// (inter / (a + g - inter), fast division) is within 5e-4 relative of the exact one, and so is every term inter / a of
// the ignore sum.  A decision is taken from float32 values only if it holds with a margin of SCR_M = 2e-3 relative:
//   argmax    second best < best * (1 - SCR_M)                      (a tie or near-tie -> exact: first maximum wins)
//   labels    |best - t| > SCR_M * t for t in {bg, fg, 0.5};  |ignore sum - ign| > SCR_M * ign
//   promotion every overlap < colmax(GT) * (1 - SCR_M)              (otherwise ov == colmax must be tested exactly)
//   zero      a side in (-SCR_T, SCR_T), a non-positive union or any NaN -> exact
// If ANY lane of a slice is undecided the whole slice takes the exact path (warp-uniform, no divergence), whose results
// are bit-identical by construction; decided lanes carry the same argmax, the same threshold outcomes and the same
// ignore outcome as the exact path, which is all label_of() and the emit phase consume.
constexpr float SCR_T = 9.765625e-4f;                // 2^-10
constexpr float SCR_M = 2e-3f;
__device__ __forceinline__ bool scr_near(const float v, const float t) { return !(fabsf(v - t) > SCR_M * fabsf(t) + 1e-7f); }
__device__ __forceinline__ void scr_stage(GtChunk &sm, const int i, const float4 f, const unsigned long long cm) {
    const float gw = f.z - f.x + 1.0f, gh = f.w - f.y + 1.0f;
    sm.area32[i] = gw * gh;
    sm.cmlo[i] = (gw >= SCR_T && gh >= SCR_T) ? __double2float_rd(__longlong_as_double((long long)cm)) * (1.0f - SCR_M) : -1.0f;
}

__device__ __forceinline__ unsigned long long iou_bits(double v) {
    return (unsigned long long)__double_as_longlong(v);
}

// float64 test "gt may intersect bbox" with the exact +1 arithmetic of bbox.pyx:39-47
__device__ __forceinline__ bool may_touch(const float4 bb, double x1, double y1, double x2, double y2) {
    const double iw = fmin((double)bb.z, x2) - fmax((double)bb.x, x1) + 1;
    const double ih = fmin((double)bb.w, y2) - fmax((double)bb.y, y1) + 1;
    return (iw > 0) && (ih > 0);
}

// Conservative float32 version for culling only: with m = min(x2), M = max(x1), iw > 0 in exact
// arithmetic means m - M > -1, and rounding is monotone, so fl32(m - M) >= -1 can only let extra pairs
// through (they then get the exact float64 treatment and contribute an overlap of 0).
__device__ __forceinline__ bool may_touch32(const float4 bb, const float4 g) {
    return (fminf(bb.z, g.z) - fmaxf(bb.x, g.x) >= -1.0f) && (fminf(bb.w, g.w) - fmaxf(bb.y, g.y) >= -1.0f);
}

// Pyramid geometry on the device (detbox_pyramid_desc + derived constants).  Anchor row i of level l is
// (y, x, cell) with i - off[l] = (y * W + x) * ncell + cell, so rel = row * rowlen + xin with rowlen = W * ncell.
struct PyrDev {
    int n_levels, E;                                  // E: table entries per image
    long long off[DETBOX_MAX_LEVELS + 1];             // first anchor row of every level
    unsigned rowlen[DETBOX_MAX_LEVELS], m_rowlen[DETBOX_MAX_LEVELS];   // divisor and floor(2^32 / divisor)
    unsigned ncell[DETBOX_MAX_LEVELS], m_ncell[DETBOX_MAX_LEVELS];
    int eoff[DETBOX_MAX_LEVELS];                      // first table entry of the level: rowlen X entries, then H * ncell Y entries
};

struct TargetParams {
    const float4 *anchors; long A;
    const float *gt; const int32_t *gt_off;
    const TileBoxes *tiles;
    unsigned long long *colmax; int32_t *counts;
    double bg, fg, ign;
    int clobber, enc;
    float4 bbw;
    float *labels; float4 *targets; float4 *inside;
    uint8_t *packed;                                  // DETBOX_TARGETS_COMPACT: (label + 1) | flag << 7, one byte per anchor
    const float2 *tab;                                // [batch, E] separable GT-0 targets (pass 0), or NULL
    const int32_t *sep_bad;                           // != 0: the anchors are not separable over the pyramid -> tables unused
    PyrDev pyr;
};

// n / d and n % d for a divisor known on the host: m = floor(2^32 / d) under-estimates the quotient by at most one
__device__ __forceinline__ void divmod(const unsigned n, const unsigned d, const unsigned m, unsigned &q, unsigned &r) {
    q = __umulhi(n, m); r = n - q * d;
    if (r >= d) { ++q; r -= d; }
}
__device__ __forceinline__ int level_of(const PyrDev &y, const long long a) {
    int l = 0;
#pragma unroll
    for (int k = 1; k < DETBOX_MAX_LEVELS; ++k)
        if (k < y.n_levels && a >= y.off[k]) l = k;
    return l;
}
constexpr int TAB_SENT = 0x7fc0dead;                 // "take the full encode" marker in a table entry's second half

// ---- pass 0 ------------------------------------------------------------------------------------
// Blocks [0, n_tiles): tile / warp-slice bounding boxes and, given a pyramid descriptor, the separability check:
// anchor (l, y, x, c) must repeat the x-interval of (l, 0, x, c) and the y-interval of (l, y, 0, c) bit for bit (what
// anchor.py:47-70 produces: x1 = base_x1 + x * stride does not depend on y).  Any mismatch raises *sep_bad and the emit
// pass ignores the tables.  Blocks [n_tiles, ...): the separable GT-0 tables, see table_entry().
//
// The fastrcnn encode is separable per axis: (dx, dw) depend on the anchor's x-interval and the GT's (cx, w) only,
// (dy, dh) on the y-interval and (cy, h).  92 % of the CP-R anchors overlap no GT and are encoded against GT 0 of their
// image (np.argmax of an all-zero row), so per image 2 * E half-encodes (E = sum_l ncell * (W_l + H_l) = 4,464 for
// CP-R) replace 261,888 full ones; the emit pass then needs two 8-byte loads per anchor.  One axis of
// encode_fastrcnn_pre, same operations in the same order; *ok = this axis passes the range test of the fast path.
__device__ __forceinline__ float2 axis_encode(const float lo, const float hi, const float gc, const float gs, const float gmark, bool &ok)
{
    const float e = (hi - lo) + 1.0f;
    const float ec = __fmaf_rn(e, 0.5f, lo);
    const float n = gc - ec;
    const float mn = fminf(fabsf(n), e), mx = fmaxf(fmaxf(fabsf(n), e), gmark);
    ok = mn >= 9.5367431640625e-07f && mx <= 1048576.0f;
    if (!ok) return make_float2(0.0f, 0.0f);
    const float y = rcp_rn_inrange(e);
    float qc = __fmul_rn(n, y), qs = __fmul_rn(gs, y);
    qc = __fmaf_rn(__fmaf_rn(-e, qc, n), y, qc);
    qs = log_inrange(__fmaf_rn(__fmaf_rn(-e, qs, gs), y, qs));
    float tc = __fmul_rn(qc, 10.0f), ts = __fmul_rn(qs, 5.0f);
    tc = __fmaf_rn(__fmaf_rn(-0.1f, tc, qc), 10.0f, tc);
    ts = __fmaf_rn(__fmaf_rn(-0.2f, ts, qs), 5.0f, ts);
    return make_float2(tc, ts);
}

__global__ void __launch_bounds__(AT_THREADS)
anchor_bbox_kernel(const float4 *__restrict__ anchors, long A, TileBoxes *__restrict__ tiles, int n_tiles,
                   const PyrDev pyr, const float *__restrict__ gt, const int32_t *__restrict__ gt_off,
                   float2 *__restrict__ tab, int32_t *__restrict__ sep_bad)
{
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if ((int)blockIdx.x >= n_tiles) {                 // ---- table blocks: (image, 256 entries) --------------------
        const int per_img = (pyr.E + AT_THREADS - 1) / AT_THREADS;
        const int k = (int)blockIdx.x - n_tiles;
        const int b = k / per_img, e = (k - b * per_img) * AT_THREADS + tid;
        const int g0 = gt_off[b];
        if (e >= pyr.E || gt_off[b + 1] == g0) return;                         // no GT: the emit pass writes zeros
        int l = 0;
#pragma unroll
        for (int q = 1; q < DETBOX_MAX_LEVELS; ++q)
            if (q < pyr.n_levels && e >= pyr.eoff[q]) l = q;
        const float *q = gt + (size_t)g0 * 5;
        const float4 ge = gt_pre(make_float4(q[0], q[1], q[2], q[3]));         // (gcx, gcy, gw | inf, gh)
        const unsigned k_in = (unsigned)(e - pyr.eoff[l]);
        bool ok; float2 t;
        if (k_in < pyr.rowlen[l]) {                                             // X entry: row y = 0
            const float4 a = anchors[pyr.off[l] + k_in];
            t = axis_encode(a.x, a.z, ge.x, ge.z, ge.z, ok);
        } else {                                                                // Y entry: column x = 0
            unsigned row, cell;
            divmod(k_in - pyr.rowlen[l], pyr.ncell[l], pyr.m_ncell[l], row, cell);
            const float4 a = anchors[pyr.off[l] + (long long)row * pyr.rowlen[l] + cell];
            t = axis_encode(a.y, a.w, ge.y, ge.w, 0.0f, ok);
        }
        if (!ok) t.y = __int_as_float(TAB_SENT);
        tab[(size_t)b * pyr.E + e] = t;
        return;
    }
    __shared__ unsigned bb[4];
    const long tile0 = (long)blockIdx.x * AT_TILE;
    const float INF = __int_as_float(0x7f800000);
    if (tid < 4) bb[tid] = (tid < 2) ? 0xffffffffu : 0u;
    __syncthreads();
    unsigned m0 = 0xffffffffu, m1 = 0xffffffffu, m2 = 0u, m3 = 0u;
    bool bad = false;
#pragma unroll
    for (int i = 0; i < AT_ITEMS; ++i) {
        const long a = tile0 + i * AT_THREADS + tid;
        const float4 v = a < A ? anchors[a] : make_float4(INF, INF, -INF, -INF);
        if (tab && a < A) {                           // separability over the pyramid
            const int l = level_of(pyr, a);
            unsigned row, xin, pos, cell;
            divmod((unsigned)(a - pyr.off[l]), pyr.rowlen[l], pyr.m_rowlen[l], row, xin);
            divmod(xin, pyr.ncell[l], pyr.m_ncell[l], pos, cell);
            const float4 rx = anchors[pyr.off[l] + xin];
            const float4 ry = anchors[pyr.off[l] + (long long)row * pyr.rowlen[l] + cell];
            bad |= !(v.x == rx.x && v.z == rx.z && v.y == ry.y && v.w == ry.w);
        }
        const unsigned r0 = __reduce_min_sync(0xffffffffu, float_to_ordered(v.x));
        const unsigned r1 = __reduce_min_sync(0xffffffffu, float_to_ordered(v.y));
        const unsigned r2 = __reduce_max_sync(0xffffffffu, float_to_ordered(v.z));
        const unsigned r3 = __reduce_max_sync(0xffffffffu, float_to_ordered(v.w));
        if (lane == 0)
            tiles[blockIdx.x].warp[i][w] = make_float4(ordered_to_float(r0), ordered_to_float(r1),
                                                       ordered_to_float(r2), ordered_to_float(r3));
        m0 = min(m0, r0); m1 = min(m1, r1); m2 = max(m2, r2); m3 = max(m3, r3);
    }
    if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(sep_bad, 1);
    if (lane == 0) {
        atomicMin(&bb[0], m0); atomicMin(&bb[1], m1);
        atomicMax(&bb[2], m2); atomicMax(&bb[3], m3);
    }
    __syncthreads();
    if (tid == 0)
        tiles[blockIdx.x].tile = make_float4(ordered_to_float(bb[0]), ordered_to_float(bb[1]),
                                             ordered_to_float(bb[2]), ordered_to_float(bb[3]));
}

// ---- pass 1: per-GT column maxima ---------------------------------------------------------------
// The column maximum of a GT row is taken over ALL anchors and the anchors are shared by the whole batch, so this
// pass has no per-image structure at all: a block owns CM_SLICES warp slices (32 anchors each; dealt round-robin) and
// CM_GT consecutive rows of the concatenated GT array, whichever images they belong to.  At CP-R (8,192 slices) with
// 64 images x 12 GT and CM_SLICES = 512 / CM_GT = 32 that is 16 x 24 blocks — ONE wave — where a (slice group, image)
// grid was 4096 blocks of a few hundred instructions each, i.e. seven waves of pure load latency.  Inside a block: the slice boxes and the GT rows are
// fetched together (one latency), every (slice, 32-GT word) pair is culled with the conservative float32 test,
// pairs with a non-empty mask go on a shared work list, and the warps take pairs off the list — 32 anchors, exact
// float64 IoU (bbox.pyx:33-54) against the GT of the mask, warp maximum, one shared-memory atomicMax on the
// order-preserving bit pattern; the block's maxima leave through fire-and-forget global atomics.
constexpr int CM_THREADS = 256;
constexpr int CM_SLICES = 512;
constexpr int CM_GT = 32;                           // GT rows per block: few GT x many slices, so that the running maximum screens
constexpr int CM_WORDS = CM_GT / 32;

__global__ void __launch_bounds__(CM_THREADS)
colmax_kernel(const TargetParams p, int n_slices, int total_gt)
{
    __shared__ double gx1[CM_GT], gy1[CM_GT], gx2[CM_GT], gy2[CM_GT], garea[CM_GT];
    __shared__ float4 gbox[CM_GT];
    __shared__ unsigned long long cmax[CM_GT];
    __shared__ float4 s_bb[CM_SLICES];
    __shared__ unsigned s_mask[CM_SLICES * CM_WORDS];
    __shared__ unsigned short s_list[CM_SLICES * CM_WORDS];
    __shared__ int s_n;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int r0 = blockIdx.y * CM_GT;                        // first GT row of this block
    const int gn = min(CM_GT, total_gt - r0);
    // slices are dealt round-robin (local slice s = global slice s * gridDim.x + blockIdx.x): the slices of the coarse
    // levels, which nearly every GT touches, are contiguous in memory and would otherwise all land in one block
    const int nb = gridDim.x;
    const int ns = (n_slices - (int)blockIdx.x + nb - 1) / nb;         // <= CM_SLICES by the launch geometry
    const float INF = __int_as_float(0x7f800000);
    for (int t = tid; t < ns; t += CM_THREADS) {              // flat slice -> (tile, item, warp) of pass 0's layout
        const int gs = t * nb + blockIdx.x;
        s_bb[t] = (&p.tiles[gs / (AT_ITEMS * AT_WARPS)].warp[0][0])[gs % (AT_ITEMS * AT_WARPS)];
    }
    if (tid < gn) {
        const float *q = p.gt + (size_t)(r0 + tid) * 5;
        const float4 f = make_float4(q[0], q[1], q[2], q[3]);
        const double x1 = f.x, y1 = f.y, x2 = f.z, y2 = f.w;
        gx1[tid] = x1; gy1[tid] = y1; gx2[tid] = x2; gy2[tid] = y2;
        garea[tid] = (x2 - x1 + 1) * (y2 - y1 + 1);           // bbox.pyx:34-37
        gbox[tid] = f;
        cmax[tid] = 0ull;
    }
    if (tid == 0) s_n = 0;
    __syncthreads();
    // coordinates beyond 2^15 void the float32 screen's error bound (slice boxes bound their anchors): decided once per block
    bool mybig = false;
    for (int t = tid; t < ns; t += CM_THREADS) {              // (an all-padding slice has an infinite box and never passes the cull)
        const float4 q = s_bb[t];
        const float v = fmaxf(fmaxf(fabsf(q.x), fabsf(q.z)), fmaxf(fabsf(q.y), fabsf(q.w)));
        mybig |= !(v <= 32768.0f) && v != INF;
    }
    if (tid < gn) { const float4 q = gbox[tid]; mybig |= !(fmaxf(fmaxf(fabsf(q.x), fabsf(q.z)), fmaxf(fabsf(q.y), fabsf(q.w))) <= 32768.0f); }
    const bool big = __syncthreads_or(mybig) != 0;
    const int nw = (gn + 31) >> 5;                            // mask words in use
    for (int t = tid; t < ns * nw; t += CM_THREADS) {         // cull: which GT of word `wd` may touch slice `sl`
        const int sl = t / nw, wd = t - sl * nw;
        const float4 bb = s_bb[sl];
        const int jn = min(32, gn - wd * 32);
        unsigned m = 0;
        for (int j = 0; j < jn; ++j) m |= may_touch32(bb, gbox[wd * 32 + j]) ? (1u << j) : 0u;
        if (m) { const int e = atomicAdd(&s_n, 1); s_mask[e] = m; s_list[e] = (unsigned short)(sl * CM_WORDS + wd); }
    }
    __syncthreads();
    const int n = s_n;
    // one warp per listed (slice, word), two per trip so that two anchor loads are in flight
    auto eval = [&](const float4 anc, unsigned m, const int jb) {
        const double ax1 = anc.x, ay1 = anc.y, ax2 = anc.z, ay2 = anc.w;
        const double aarea = (ax2 - ax1 + 1) * (ay2 - ay1 + 1);
        const float aarea32 = (anc.z - anc.x + 1.0f) * (anc.w - anc.y + 1.0f);
        while (m) {
            const int j = jb + __ffs(m) - 1;
            m &= m - 1;
            // Float32 screen: only a slice whose best overlap comes within 2 % of the GT's current maximum can raise
            // it.  The float32 quotient is within ~1e-3 relative of the exact one when both intersection sides are
            // at least 2^-10 and coordinates stay below 2^15 (`big`, per block) (inputs are exact, every rounding is relative to its
            // result, the worst term is span / side * 2^-24); anything outside that, or a non-positive union, is
            // "uncertain" and takes the exact path.  A stale (smaller) maximum only screens less.
            {
                const float4 g = gbox[j];
                const float iw = fminf(anc.z, g.z) - fmaxf(anc.x, g.x) + 1.0f, ih = fminf(anc.w, g.w) - fmaxf(anc.y, g.y) + 1.0f;
                float ov32 = 0.0f; bool unsure = false;
                if (iw > 0.0f && ih > 0.0f) {
                    const float inter = iw * ih;
                    const float uni = aarea32 + (g.z - g.x + 1.0f) * (g.w - g.y + 1.0f) - inter;
                    ov32 = __fdividef(inter, uni);
                    unsure = !(iw >= 9.765625e-4f && ih >= 9.765625e-4f && uni > 0.0f && ov32 < 4.0f);
                }
                const unsigned a32 = __reduce_max_sync(0xffffffffu, __float_as_uint(ov32));
                // the running maximum is read by ONE lane and broadcast: other warps atomicMax the same word meanwhile, and
                // the `continue` below must be taken by all 32 lanes or none (full-mask sync operations follow)
                unsigned long long cm = 0ull;
                if (lane == 0) cm = cmax[j];
                cm = __shfl_sync(0xffffffffu, cm, 0);
                if (!big && !__any_sync(0xffffffffu, unsure) &&
                    (double)__uint_as_float(a32) * 1.02 < __longlong_as_double((long long)cm)) continue;
            }
            unsigned long long bits = 0ull;
            const double iw = fmin(ax2, gx2[j]) - fmax(ax1, gx1[j]) + 1;          // bbox.pyx:39-42
            if (iw > 0) {
                const double ih = fmin(ay2, gy2[j]) - fmax(ay1, gy1[j]) + 1;      // bbox.pyx:44-47
                if (ih > 0) {
                    const double inter = iw * ih;
                    const double ov = inter / (aarea + garea[j] - inter);         // bbox.pyx:49-54
                    if (ov > 0.0) bits = iou_bits(ov);
                }
            }
            // warp maximum first (positive doubles order like their bit patterns: high word, then low word),
            // then ONE shared-memory atomic — a 64-bit shared atomicMax is a compare-and-swap loop
            const unsigned hi = __reduce_max_sync(0xffffffffu, (unsigned)(bits >> 32));
            const unsigned lo = __reduce_max_sync(0xffffffffu, (unsigned)(bits >> 32) == hi ? (unsigned)bits : 0u);
            if (lane == 0 && (hi | lo)) atomicMax(&cmax[j], ((unsigned long long)hi << 32) | lo);
        }
    };
    const float4 none = make_float4(INF, INF, -INF, -INF);
    for (int e = w; e < n; e += 2 * (CM_THREADS / 32)) {
        const int e1 = e + CM_THREADS / 32;
        const int ia = s_list[e], ib = e1 < n ? s_list[e1] : -1;
        const int sa = ia / CM_WORDS, sb = max(ib, 0) / CM_WORDS;
        const long a0 = ((long)(sa * nb + (int)blockIdx.x) << 5) + lane, a1 = ((long)(sb * nb + (int)blockIdx.x) << 5) + lane;
        const float4 anc0 = a0 < p.A ? p.anchors[a0] : none;
        const float4 anc1 = (ib >= 0 && a1 < p.A) ? p.anchors[a1] : none;
        eval(anc0, s_mask[e], (ia % CM_WORDS) * 32);
        if (ib >= 0) eval(anc1, s_mask[e1], (ib % CM_WORDS) * 32);
    }
    __syncthreads();
    if (tid < gn && cmax[tid] != 0ull) atomicMax(&p.colmax[r0 + tid], cmax[tid]);
}

// label and regression flag of one anchor from its row summary — anchor.py:49-62,86-87
template <class P>                                  // P: anything with bg / fg / ign / clobber (TargetParams, LabelRule)
__device__ __forceinline__ void label_of(const P &p, float gcls, double mo, bool promoted, double isum,
                                         float &l_out, bool &flag_out)
{
    float l = gcls;                                                         // anchor.py:49
    if (mo < p.bg) l = 0.0f;                                                // :50
    if (mo < p.fg && mo >= p.bg) l = -1.0f;                                 // :51-52
    bool flag = mo >= 0.5;                                                  // :53 (hard-coded 0.5)
    if (promoted) l = gcls;                                                 // :56 own-argmax class
    if (p.clobber && mo < p.bg) l = 0.0f;                                   // :60-61
    if (l >= 1.0f) flag = true;                                             // :62
    if (isum > p.ign) { l = -1.0f; flag = false; }                          // :86-87
    l_out = l; flag_out = flag;
}

// One anchor's outputs.  Reference format: label, targets, inside weights (36 B).  Compact format: the targets and one
// byte (label + 1) | flag << 7 staged in shared memory (flushed 16 bytes per thread by pk_flush) — the inside weights
// are flag ? bbweights : 0 and the label weights a function of the label and the image's (nfg, nbg), so the byte and
// the two counts carry everything (detbox_expand_targets_host rebuilds the reference's arrays bit for bit).
__device__ __forceinline__ unsigned pack_label(const float l, const bool flag) {
    const int li = (int)l;
    const bool ok = (l == (float)li) && li >= -1 && li <= 125;
    return (ok ? (unsigned)(li + 1) : 127u) | (flag ? 0x80u : 0u);
}
template <bool COMPACT>
__device__ __forceinline__ void store_anchor(const TargetParams &p, const long o, const int off, uint8_t *s_pk,
                                             const float4 t, const float l, const bool flag)
{
    // targets and inside weights (32 of the 36 bytes) are streaming stores: they are not read again on the device, and
    // evicting them first lets the 4-byte labels stay in L2 for the label-weight pass that follows
    __stcs(&p.targets[o], t);                                                   // anchor.py:133-135 every anchor
    if (COMPACT) s_pk[off] = (uint8_t)pack_label(l, flag);
    else {
        p.labels[o] = l;
        __stcs(&p.inside[o], flag ? p.bbw : make_float4(0, 0, 0, 0));           // :136-138
    }
}
// tile's packed bytes -> global: 16 bytes per thread when the tile is whole and its first byte 16-byte aligned
__device__ __forceinline__ void pk_flush(const TargetParams &p, const uint8_t *s_pk, const long o0, const int rem)
{
    __syncthreads();
    uint8_t *dst = p.packed + o0;
    if (rem == AT_TILE && ((size_t)dst & 15) == 0) {
        if (threadIdx.x < AT_TILE / 16)
            reinterpret_cast<uint4 *>(dst)[threadIdx.x] = reinterpret_cast<const uint4 *>(s_pk)[threadIdx.x];
    } else {
        for (int i = threadIdx.x; i < rem; i += AT_THREADS) dst[i] = s_pk[i];
    }
}

// GT-0 targets of anchor row ai from the separable tables of its image (tb = tab + b * E); false: an entry carries the
// "full encode" marker.  lvl >= 0: the level is known (tile inside one level).
__device__ __forceinline__ bool tab_targets(const PyrDev &y, const float2 *__restrict__ tb, const int lvl, const long long ai, float4 &t)
{
    const int l = lvl >= 0 ? lvl : level_of(y, ai);
    unsigned row, xin, pos, cell;
    divmod((unsigned)(ai - y.off[l]), y.rowlen[l], y.m_rowlen[l], row, xin);
    divmod(xin, y.ncell[l], y.m_ncell[l], pos, cell);
    const float2 X = __ldg(tb + y.eoff[l] + xin);
    const float2 Y = __ldg(tb + y.eoff[l] + y.rowlen[l] + row * y.ncell[l] + cell);
    t = make_float4(X.x, Y.x, X.y, Y.y);
    return (__float_as_int(X.y) != TAB_SENT) & (__float_as_int(Y.y) != TAB_SENT);
}

// ---- decide phase for 33 .. GCH GT per image (one staged chunk, several 32-GT mask words per slice) -------------------
// anchor_target_kernel's common path (at most 32 GT: one mask word per slice, everything in registers) sits exactly at 64
// registers and ~2,900 instructions, and both are load-bearing (four blocks per SM; the kernel is sensitive to its code
// size): folding a loop over mask words into it cost 10-13 % at G = 12 (profiles/r02_anchor_target_experiments.txt).
// Images with more GT therefore take this FUNCTION — not inlined, so it has a register allocation of its own and the
// common path is unchanged to the instruction: the same cull-up-front + float32 screen + exact fallback, over
// nw = ceil(G / 32) words.  Before, such images went through the chunked loop: no up-front cull, no screen, exact
// float64 for every (slice, GT) hit.  Writes s_res for the touched slices and the OR of a slice's words into s_any
// (the emit phase's "skipped" test); returns the thread's (fg, bg) counts.
struct LabelRule { double bg, fg, ign; int clobber; };
constexpr int AT_MW = GCH / 32;
template <bool SCREEN>
__device__ __noinline__ int2 decide_multi(const LabelRule rule, const float4 *__restrict__ anchors, const long A, const long tile0,
                                          const float *__restrict__ gt_img /* the image's first GT row */, const int G,
                                          const GtChunk *__restrict__ smp, const float4 *__restrict__ s_wbb,
                                          int2 (*__restrict__ s_res)[AT_THREADS], unsigned (*__restrict__ s_any)[AT_WARPS],
                                          const int zero_any, const float l_zero, const int flag_zero)
{
    __shared__ unsigned s_mw[AT_MW][AT_ITEMS][AT_WARPS];
    const GtChunk &sm = *smp;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int nw = (G + 31) >> 5;
    const float INF = __int_as_float(0x7f800000);
#pragma unroll 1
    for (int i = 0; i < AT_ITEMS; ++i) {
        const float4 wbb = s_wbb[i * AT_WARPS + w];
        unsigned any = 0;
#pragma unroll 1
        for (int wd = 0; wd < nw; ++wd) {
            const int j = wd * 32 + lane;
            const unsigned m = __ballot_sync(0xffffffffu, j < G && may_touch32(wbb, sm.fbox[min(j, GCH - 1)]));
            if (lane == 0) s_mw[wd][i][w] = m;
            any |= m;
        }
        if (lane == 0) s_any[i][w] = any;
    }
    __syncwarp();
    int nfg = 0, nbg = 0;
#pragma unroll 1
    for (int i = 0; i < AT_ITEMS; ++i) {
        if (s_any[i][w] == 0) continue;
        const long a = tile0 + i * AT_THREADS + tid;
        const bool valid = a < A;
        const float4 anc = valid ? anchors[a] : make_float4(INF, INF, -INF, -INF);
        double best = 0.0, isum = 0.0;
        int best_g = 0;
        bool promoted = zero_any != 0;
        bool amb = !SCREEN;
        float best32 = 0.0f, second32 = 0.0f, isum32 = 0.0f; int bj = 0;
        if (SCREEN) {                                            // the float32 screen of the kernel, word after word
            const float aw = anc.z - anc.x + 1.0f, ah = anc.w - anc.y + 1.0f, aarea32 = aw * ah;
            if (valid && !(aw >= SCR_T && ah >= SCR_T)) amb = true;
#pragma unroll 1
            for (int wd = 0; wd < nw; ++wd) {
                unsigned mm = s_mw[wd][i][w];
                while (mm) {                                     // ascending GT index, like the exact loop
                    const int j = wd * 32 + __ffs(mm) - 1;
                    mm &= mm - 1;
                    const float4 g = sm.fbox[j];
                    const float iw = fminf(anc.z, g.z) - fmaxf(anc.x, g.x) + 1.0f, ih = fminf(anc.w, g.w) - fmaxf(anc.y, g.y) + 1.0f;
                    if (iw <= -SCR_T || ih <= -SCR_T) continue;
                    if (!(iw >= SCR_T && ih >= SCR_T)) { amb = true; continue; }
                    const float inter = iw * ih, ua = aarea32 + sm.area32[j] - inter;
                    const float ov = __fdividef(inter, ua);
                    if (!(ua > 0.0f && ov < 2.0f && ov < sm.cmlo[j])) amb = true;
                    if (ov > best32) { second32 = best32; best32 = ov; bj = j; } else second32 = fmaxf(second32, ov);
                    if (sm.cls[j] < 0.0f) isum32 += __fdividef(inter, aarea32);
                }
            }
            if (best32 > 0.0f) {
                amb |= !(second32 < best32 * (1.0f - SCR_M));
                amb |= scr_near(best32, (float)rule.bg) | scr_near(best32, (float)rule.fg) | scr_near(best32, 0.5f);
                amb |= isum32 > 0.0f ? scr_near(isum32, (float)rule.ign) : !(isum32 == 0.0f);
            }
        }
        if (__any_sync(0xffffffffu, amb)) {                      // exact, the whole slice (bbox.pyx:33-54, anchor.py:42-46,81-85)
            const double ax1 = anc.x, ay1 = anc.y, ax2 = anc.z, ay2 = anc.w;
            const double aarea = (ax2 - ax1 + 1) * (ay2 - ay1 + 1);
#pragma unroll 1
            for (int wd = 0; wd < nw; ++wd) {
                unsigned m = s_mw[wd][i][w];
                while (m) {
                    const int j = wd * 32 + __ffs(m) - 1;
                    m &= m - 1;
                    double ov = 0.0;
                    const double iw = fmin(ax2, sm.x2[j]) - fmax(ax1, sm.x1[j]) + 1;
                    if (iw > 0) {
                        const double ih = fmin(ay2, sm.y2[j]) - fmax(ay1, sm.y1[j]) + 1;
                        if (ih > 0) {
                            const double inter = iw * ih;
                            const double ua = aarea + sm.area[j] - inter;
                            ov = inter / ua;
                            if (sm.cls[j] < 0.0f) isum = isum + inter / aarea;
                        }
                    }
                    if (ov > 0.0) {
                        if (ov > best) { best = ov; best_g = j; }
                        if (valid && iou_bits(ov) == sm.colmax[j]) promoted = true;
                    }
                }
            }
        } else { best = (double)best32; best_g = bj; isum = (double)isum32; }
        float l = l_zero; bool flag = flag_zero != 0;
        int g = 0;
        if (best > 0.0) {
            g = best_g;
            label_of(rule, gt_img[(size_t)g * 5 + 4], best, promoted, isum, l, flag);
        }
        s_res[i][tid] = make_int2(__float_as_int(l), g | (flag ? (int)0x80000000 : 0));
        if (valid) { nfg += (l >= 1.0f); nbg += (l == 0.0f); }
    }
    return make_int2(nfg, nbg);
}

// EMIT_ILP: anchors encoded side by side in the emit loops; MINB: blocks per SM; TABLE: GT-0 targets come from the
// separable tables when the pyramid check passed; COMPACT: output format; SCREEN: float32 screen of the decide phase
template <int EMIT_ILP, bool FRCNN, int MINB, bool TABLE, bool COMPACT, bool SCREEN>
__global__ void __launch_bounds__(AT_THREADS, MINB)
anchor_target_kernel(const TargetParams p)
{
    __shared__ GtChunk sm;
    __shared__ float4 s_wbb[AT_ITEMS * AT_WARPS];    // bounding boxes of the tile's warp slices
    __shared__ __align__(16) uint8_t s_pk[COMPACT ? AT_TILE : 16];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int b = blockIdx.y;
    // loads that do not depend on the image are issued first, so their latency overlaps the GT's
    float4 my_wbb = make_float4(0, 0, 0, 0);
    if (tid < AT_ITEMS * AT_WARPS) my_wbb = (&p.tiles[blockIdx.x].warp[0][0])[tid];
    const float4 tbox = p.tiles[blockIdx.x].tile;
    const int g0 = p.gt_off[b];
    const int G = p.gt_off[b + 1] - g0;
    const long tile0 = (long)blockIdx.x * AT_TILE;
    const int rem = (int)min((long)AT_TILE, p.A - tile0);      // anchors of this tile (the last one may be partial)
    const float INF = __int_as_float(0x7f800000);
    const long o0 = (long)b * p.A + tile0;

    if (G == 0) {                                    // anchor.py:89-91,133: all background, zero targets
        {
            int nbg = 0;
#pragma unroll
            for (int i = 0; i < AT_ITEMS; ++i) {
                const int off = i * AT_THREADS + tid;
                if (off >= rem) continue;
                const long o = o0 + off;
                p.targets[o] = make_float4(0, 0, 0, 0);
                if (COMPACT) s_pk[off] = 1;
                else {
                    p.labels[o] = 0.0f;
                    p.inside[o] = make_float4(0, 0, 0, 0);
                }
                ++nbg;
            }
            nbg = __reduce_add_sync(0xffffffffu, nbg);
            if (lane == 0 && nbg) atomicAdd(&p.counts[b * 2 + 1], nbg);
        }
        if (COMPACT) pk_flush(p, s_pk, o0, rem);
        return;
    }
    // tables: usable when pass 0 found the anchors separable; the tile's level when it lies inside one
    const bool use_tab = TABLE && p.tab != nullptr && *p.sep_bad == 0;
    const float2 *tb = nullptr; int lvl = -1;
    if (TABLE && use_tab) {
        tb = p.tab + (size_t)b * p.pyr.E;
        const int l0 = level_of(p.pyr, tile0), l1 = level_of(p.pyr, tile0 + rem - 1);
        lvl = l0 == l1 ? l0 : -1;
    }

    // ---- can any GT of this image touch this tile?  is any column all-zero? ---------------------
    // Up to GCH GT (always, at CityPersons / COCO counts): the GT rows are read ONCE — staged in shared memory
    // (float64 copies, areas, column maxima) and tested against the tile's box in the same step.
    const bool single = G <= GCH;
    if (tid < AT_ITEMS * AT_WARPS) s_wbb[tid] = my_wbb;
    int flags = 0;                                   // bit 0: a GT may touch the tile; bit 1: some column is all-zero
    if (single) {
        if (tid < G) {
            const float *q = p.gt + (size_t)(g0 + tid) * 5;
            const float4 f = make_float4(q[0], q[1], q[2], q[3]);
            const double x1 = f.x, y1 = f.y, x2 = f.z, y2 = f.w;
            sm.x1[tid] = x1; sm.y1[tid] = y1; sm.x2[tid] = x2; sm.y2[tid] = y2;
            sm.area[tid] = (x2 - x1 + 1) * (y2 - y1 + 1);                 // bbox.pyx:34-37
            sm.cls[tid] = q[4];
            sm.fbox[tid] = f;
            sm.genc[tid] = gt_pre(f);
            const unsigned long long cm = p.colmax[g0 + tid];
            sm.colmax[tid] = cm;
            scr_stage(sm, tid, f, cm);
            flags = (may_touch(tbox, x1, y1, x2, y2) ? 1 : 0) | (cm == 0ull ? 2 : 0);   // all-zero column: every anchor ties (anchor.py:46)
        }
    } else {
        for (int g = tid; g < G; g += AT_THREADS) {
            const float *q = p.gt + (size_t)(g0 + g) * 5;
            flags |= may_touch(tbox, q[0], q[1], q[2], q[3]) ? 1 : 0;
            if (p.colmax[g0 + g] == 0ull) flags |= 2;
        }
    }
    // (__syncthreads_or is a logical OR of predicates, not a bitwise one: one call per flag; the first also
    //  publishes the staged GT and slice boxes)
    const int cand_any = __syncthreads_or(flags & 1);
    const int zero_any = __syncthreads_or(flags & 2);
    const float4 *anc_t = p.anchors + tile0;

    float4 gbox0; float gcls0;
    if (single) { gbox0 = sm.fbox[0]; gcls0 = sm.cls[0]; }
    else { const float *q0 = p.gt + (size_t)g0 * 5; gbox0 = make_float4(q0[0], q0[1], q0[2], q0[3]); gcls0 = q0[4]; }
    float l_zero; bool flag_zero;                    // outcome of an anchor that overlaps nothing: argmax = GT 0, max overlap 0
    label_of(p, gcls0, 0.0, zero_any != 0, 0.0, l_zero, flag_zero);

    if (!cand_any) {
        // streaming path: every row is all-zero -> (l_zero, GT 0) for every anchor of the tile
        const float4 ge = gt_pre(gbox0);
        if (TABLE && use_tab) {
#pragma unroll 1
            for (int h = 0; h < AT_ITEMS; h += 2) {              // two lookups (four 8-byte loads) in flight
                float4 t[2]; bool ok[2];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int off = min((h + i) * AT_THREADS + tid, rem - 1);
                    ok[i] = tab_targets(p.pyr, tb, lvl, tile0 + off, t[i]);
                }
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int off = (h + i) * AT_THREADS + tid;
                    if (off >= rem) continue;
                    if (!ok[i]) t[i] = FRCNN ? encode_fastrcnn_pre(anc_t[off], ge, [&]() { return gbox0; }) : t[i];
                    store_anchor<COMPACT>(p, o0 + off, off, s_pk, t[i], l_zero, flag_zero);
                }
            }
        } else {
#pragma unroll 1
            for (int h = 0; h < AT_ITEMS; h += EMIT_ILP) {      // EMIT_ILP independent load -> encode -> store chains
                float4 anc[EMIT_ILP];
#pragma unroll
                for (int i = 0; i < EMIT_ILP; ++i) {
                    const int off = (h + i) * AT_THREADS + tid;
                    anc[i] = off < rem ? anc_t[off] : make_float4(0, 0, 0, 0);
                }
#pragma unroll
                for (int i = 0; i < EMIT_ILP; ++i) {
                    const int off = (h + i) * AT_THREADS + tid;
                    if (off >= rem) continue;
                    const float4 t = FRCNN ? encode_fastrcnn_pre(anc[i], ge, [&]() { return gbox0; })
                                           : encode_linear(anc[i], gbox0, p.enc == DETBOX_ENC_LINEAR);
                    store_anchor<COMPACT>(p, o0 + off, off, s_pk, t, l_zero, flag_zero);
                }
            }
        }
        if (tid == 0) {                                          // every anchor of the tile has the label l_zero
            if (l_zero >= 1.0f) atomicAdd(&p.counts[b * 2 + 0], rem);
            if (l_zero == 0.0f) atomicAdd(&p.counts[b * 2 + 1], rem);
        }
        if (COMPACT) pk_flush(p, s_pk, o0, rem);
        return;
    }

    // ---- tiles some GT may touch: decide again per 32-anchor warp slice ---------------------------
    // (a tile is a long horizontal stripe of the level grid and meets some GT more often than not;
    //  a warp slice spans ~5 grid positions and rarely does)
    auto stage = [&](int c0) {                       // GT chunk -> shared memory (float64 copies, areas, column maxima)
        const int gn = min(GCH, G - c0);
        __syncthreads();                             // previous chunk fully consumed
        if (tid < gn) {
            const float *q = p.gt + (size_t)(g0 + c0 + tid) * 5;
            const double x1 = q[0], y1 = q[1], x2 = q[2], y2 = q[3];
            sm.x1[tid] = x1; sm.y1[tid] = y1; sm.x2[tid] = x2; sm.y2[tid] = y2;
            sm.area[tid] = (x2 - x1 + 1) * (y2 - y1 + 1);                 // bbox.pyx:34-37
            sm.cls[tid] = q[4];
            sm.fbox[tid] = make_float4(q[0], q[1], q[2], q[3]);
            sm.genc[tid] = gt_pre(sm.fbox[tid]);
            sm.colmax[tid] = p.colmax[g0 + c0 + tid];
            scr_stage(sm, tid, sm.fbox[tid], sm.colmax[tid]);
        }
        __syncthreads();
    };

    const float4 *wbb_base = &s_wbb[w];
    int nfg = 0, nbg = 0;
    // pass 2, decide phase -> emit phase: (label, GT row | regression flag << 31) of every anchor of the tile
    __shared__ int2 s_res[AT_ITEMS][AT_THREADS];
    // With at most 32 GT (one lane per GT) the cull of all the warp's slices is done up front — AT_ITEMS
    // independent bounding-box loads and ballots — instead of one dependent load per slice inside the loop.
    __shared__ unsigned smask[AT_ITEMS][AT_WARPS];
    const bool use_masks = G <= 32;
    const bool multi = single && !use_masks;         // 33 .. GCH GT: decide_multi() (its own register allocation)
    const bool have_masks = single;                  // smask[i][w] == 0 <=> no GT near the slice (both routes)
    if (multi) {
        LabelRule rule; rule.bg = p.bg; rule.fg = p.fg; rule.ign = p.ign; rule.clobber = p.clobber;
        const int2 c = decide_multi<SCREEN>(rule, p.anchors, p.A, tile0, p.gt + (size_t)g0 * 5, G, &sm, s_wbb, s_res, smask,
                                            zero_any, l_zero, flag_zero ? 1 : 0);
        nfg = c.x; nbg = c.y;
    }
    if (use_masks) {
        // four threads per slice (64 slices = 256 threads), each testing every fourth GT against the slice's box, the
        // partial masks combined by two shuffles: G / 4 short trips per thread instead of eight 32-lane ballots per
        // warp with at most G useful lanes each (6.5 M of the kernel's 105 M warp instructions at G = 12)
        static_assert(AT_THREADS == 4 * AT_ITEMS * AT_WARPS, "one thread per (slice, GT quarter)");
        const float4 wbb = s_wbb[tid >> 2];
        unsigned m = 0;
        for (int j = tid & 3; j < G; j += 4) m |= may_touch32(wbb, sm.fbox[j]) ? 1u << j : 0u;
        m |= __shfl_xor_sync(0xffffffffu, m, 1);
        m |= __shfl_xor_sync(0xffffffffu, m, 2);
        if ((tid & 3) == 0) (&smask[0][0])[tid >> 2] = m;
        __syncthreads();
    }
    // Decide phase.  One slice at a time keeps the float64 row state (argmax, ignore sum, area) of a single
    // anchor in registers; only slices some GT may touch load their anchors here.  With more than GCH GT the
    // chunks are re-staged per slice (all warps together).
    // (Measured and rejected, B200, G = 12: fusing this loop into the emit loop — decision in registers from the anchor
    //  the emit is about to encode, next anchors prefetched meanwhile — ran 0.164 vs 0.153 ms: the decide chain then sits in
    //  front of the stores and the extra live state spills.  The float32 screen below removes the float64 division from
    //  the common path (17 % of the kernel's instructions) but only 1 % of its time at G = 12 (7 % at G = 32): the
    //  arithmetic of this loop is hidden, its per-slice latency chain is not.  profiles/r02_anchor_target_experiments.txt)
#pragma unroll 1
    for (int i = 0; i < (multi ? 0 : AT_ITEMS); ++i) {
        if (use_masks && smask[i][w] == 0) continue;             // no GT near this slice: the emit phase uses (l_zero, GT 0)
        const long a = tile0 + i * AT_THREADS + tid;
        const bool valid = a < p.A;
        float4 anc = make_float4(INF, INF, -INF, -INF);
        bool have_anc = false;
        double best = 0.0, isum = 0.0;
        int best_g = 0;
        bool promoted = zero_any != 0;
        // exact float64 IoU of this lane's anchor against the GT whose bits are set in m (GT index cb + bit)
        auto eval_hits = [&](unsigned m, int cb, int c0) {
            if (!have_anc) { if (valid) anc = p.anchors[a]; have_anc = true; }
            const double ax1 = anc.x, ay1 = anc.y, ax2 = anc.z, ay2 = anc.w;
            const double aarea = (ax2 - ax1 + 1) * (ay2 - ay1 + 1);
            while (m) {                          // ascending GT index: first max wins (anchor.py:42)
                const int j = cb + __ffs(m) - 1;
                m &= m - 1;
                const double gx1 = sm.x1[j], gy1 = sm.y1[j], gx2 = sm.x2[j], gy2 = sm.y2[j];
                double ov = 0.0;
                const double iw = fmin(ax2, gx2) - fmax(ax1, gx1) + 1;          // bbox.pyx:39-42
                if (iw > 0) {
                    const double ih = fmin(ay2, gy2) - fmax(ay1, gy1) + 1;      // bbox.pyx:44-47
                    if (ih > 0) {
                        const double inter = iw * ih;
                        const double ua = aarea + sm.area[j] - inter;           // bbox.pyx:49-53
                        ov = inter / ua;                                        // bbox.pyx:54
                        if (sm.cls[j] < 0.0f)
                            isum = isum + inter / aarea;                        // anchor.py:81-85 (bbox.pyx:93)
                    }
                }
                if (ov > 0.0) {
                    if (ov > best) { best = ov; best_g = c0 + j; }
                    if (valid && iou_bits(ov) == sm.colmax[j]) promoted = true; // anchor.py:46
                }
            }
        };
        if (use_masks) {
            const unsigned m = smask[i][w];                      // != 0 here
            bool amb = !SCREEN;
            float best32 = 0.0f, second32 = 0.0f, isum32 = 0.0f; int bj = 0;
            if (SCREEN) {
                if (valid) anc = p.anchors[a];
                have_anc = true;
                const float aw = anc.z - anc.x + 1.0f, ah = anc.w - anc.y + 1.0f, aarea32 = aw * ah;
                if (valid && !(aw >= SCR_T && ah >= SCR_T)) amb = true;
                unsigned mm = m;
                while (mm) {                                     // ascending GT index, like the exact loop
                    const int j = __ffs(mm) - 1;
                    mm &= mm - 1;
                    const float4 g = sm.fbox[j];
                    const float iw = fminf(anc.z, g.z) - fmaxf(anc.x, g.x) + 1.0f, ih = fminf(anc.w, g.w) - fmaxf(anc.y, g.y) + 1.0f;
                    // no per-lane early exits (they diverge and cost convergence barriers): one warp-uniform skip, then
                    // predicates.  dis: certainly disjoint (padding lanes: -inf); ok: certainly overlapping (false for NaN);
                    // neither: within rounding of touching -> exact path.  A lane that is not `ok` contributes ov = 0,
                    // which changes neither the maximum nor the runner-up.
                    const bool dis = iw <= -SCR_T || ih <= -SCR_T;
                    if (__all_sync(0xffffffffu, dis)) continue;
                    const bool ok = iw >= SCR_T && ih >= SCR_T;
                    const float inter = iw * ih, ua = aarea32 + sm.area32[j] - inter;
                    const float ovr = __fdividef(inter, ua);
                    amb |= !dis & !(ok && ua > 0.0f && ovr < 2.0f && ovr < sm.cmlo[j]);   // touching / odd union / may equal the column maximum
                    const float ov = ok ? ovr : 0.0f;
                    if (ov > best32) { second32 = best32; best32 = ov; bj = j; } else second32 = fmaxf(second32, ov);
                    if (sm.cls[j] < 0.0f) isum32 += ok ? __fdividef(inter, aarea32) : 0.0f;
                }
                if (best32 > 0.0f) {
                    amb |= !(second32 < best32 * (1.0f - SCR_M));
                    amb |= scr_near(best32, (float)p.bg) | scr_near(best32, (float)p.fg) | scr_near(best32, 0.5f);
                    amb |= isum32 > 0.0f ? scr_near(isum32, (float)p.ign) : !(isum32 == 0.0f);
                }
            }
            if (__any_sync(0xffffffffu, amb)) eval_hits(m, 0, 0);                     // exact, the whole slice
            else { best = (double)best32; best_g = bj; isum = (double)isum32; }      // promoted stays zero_any: no overlap reaches a column maximum
        } else {
            const float4 wbb = wbb_base[i * AT_WARPS];
            for (int c0 = 0; c0 < G; c0 += GCH) {
                if (!single) stage(c0);
                const int gn = min(GCH, G - c0);
                for (int cb = 0; cb < gn; cb += 32) {    // 32 GT at a time: one lane tests one GT against the slice's box
                    const int jl = cb + lane;
                    const unsigned m = __ballot_sync(0xffffffffu, jl < gn && may_touch32(wbb, sm.fbox[jl]));
                    if (m) eval_hits(m, cb, c0);         // warp-uniform
                }
            }
        }
        {
            float l = l_zero; bool flag = flag_zero;
            int g = 0;
            if (best > 0.0) {                        // (promotion and ignore overlap both imply best > 0)
                g = best_g;
                label_of(p, p.gt[(size_t)(g0 + g) * 5 + 4], best, promoted, isum, l, flag);
            }
            s_res[i][tid] = make_int2(__float_as_int(l), g | (flag ? (int)0x80000000 : 0));
            if (valid) { nfg += (l >= 1.0f); nbg += (l == 0.0f); }
        }
    }
    // Emit phase: a plain streaming loop over the tile — anchor in, 36 bytes out.  Every thread reads back only
    // what it wrote itself, so no barrier is needed in between.
    const int2 r_zero = make_int2(__float_as_int(l_zero), flag_zero ? (int)0x80000000 : 0);
    int nskip = 0;                                   // anchors of skipped slices: all carry l_zero
    auto raw_gt_of = [&](const int g) {
        if (single) return sm.fbox[g];
        const float *q = p.gt + (size_t)(g0 + g) * 5; return make_float4(q[0], q[1], q[2], q[3]);
    };
    if (TABLE && use_tab) {
        // GT 0 (every anchor that overlaps nothing, 92 % at CP-R) comes from the tables; only the lanes matched to
        // another GT, or whose entry carries the marker, load their anchor and run the encode
#pragma unroll 1
        for (int h = 0; h < AT_ITEMS; ++h) {
            const int off = h * AT_THREADS + tid;
            const bool skipped = have_masks && smask[h][w] == 0;
            const int2 r = skipped ? r_zero : s_res[h][tid];
            nskip += (skipped && off < rem);
            const int g = r.y & 0x7fffffff;
            float4 t;
            const bool ok = tab_targets(p.pyr, tb, lvl, tile0 + min(off, rem - 1), t);
            if (off >= rem) continue;
            if (g != 0 || !ok) {
                const float4 ge = single ? sm.genc[g] : gt_pre(raw_gt_of(g));
                t = FRCNN ? encode_fastrcnn_pre(anc_t[off], ge, [&]() { return raw_gt_of(g); }) : t;
            }
            store_anchor<COMPACT>(p, o0 + off, off, s_pk, t, __int_as_float(r.x), r.y < 0);
        }
    } else {
        // EMIT_ILP independent encode chains per thread (the encode is a long dependent sequence: reciprocal, log, divisions)
        float4 nxt[EMIT_ILP];                            // the next trip's anchors are requested before this trip's encode
#pragma unroll
        for (int i = 0; i < EMIT_ILP; ++i) { const int off = i * AT_THREADS + tid; nxt[i] = off < rem ? anc_t[off] : make_float4(0, 0, 0, 0); }
#pragma unroll 1
        for (int h = 0; h < AT_ITEMS; h += EMIT_ILP) {
            float4 anc[EMIT_ILP], ge[EMIT_ILP]; int2 r[EMIT_ILP];
#pragma unroll
            for (int i = 0; i < EMIT_ILP; ++i) {
                const int off = (h + i) * AT_THREADS + tid;
                anc[i] = nxt[i];
                const int noff = off + EMIT_ILP * AT_THREADS;
                if (MINB < 4) nxt[i] = (h + EMIT_ILP < AT_ITEMS && noff < rem) ? anc_t[noff] : make_float4(0, 0, 0, 0);
                const bool skipped = have_masks && smask[h + i][w] == 0;
                r[i] = skipped ? r_zero : s_res[h + i][tid];
                nskip += (skipped && off < rem);
                const int g = r[i].y & 0x7fffffff;
                if (single) ge[i] = sm.genc[g];
                else ge[i] = gt_pre(raw_gt_of(g));
            }
#pragma unroll
            for (int i = 0; i < EMIT_ILP; ++i) {
                const int off = (h + i) * AT_THREADS + tid;
                const int g = r[i].y & 0x7fffffff;
                if (off >= rem) continue;
                const float4 t = FRCNN ? encode_fastrcnn_pre(anc[i], ge[i], [&]() { return raw_gt_of(g); })
                                       : encode_linear(anc[i], raw_gt_of(g), p.enc == DETBOX_ENC_LINEAR);
                store_anchor<COMPACT>(p, o0 + off, off, s_pk, t, __int_as_float(r[i].x), r[i].y < 0);
            }
            if (MINB >= 4) {                             // 64 registers: no room to hold the next anchors across the encode
#pragma unroll
                for (int i = 0; i < EMIT_ILP; ++i) {
                    const int noff = (h + EMIT_ILP + i) * AT_THREADS + tid;
                    nxt[i] = (h + EMIT_ILP < AT_ITEMS && noff < rem) ? anc_t[noff] : make_float4(0, 0, 0, 0);
                }
            }
        }
    }
    if (l_zero >= 1.0f) nfg += nskip;
    if (l_zero == 0.0f) nbg += nskip;
    nfg = __reduce_add_sync(0xffffffffu, nfg);
    nbg = __reduce_add_sync(0xffffffffu, nbg);
    if (lane == 0) {
        if (nfg) atomicAdd(&p.counts[b * 2 + 0], nfg);
        if (nbg) atomicAdd(&p.counts[b * 2 + 1], nbg);
    }
    if (COMPACT) pk_flush(p, s_pk, o0, rem);
}

// pass 3 — anchor.py:111-118: fg weight 1, bg weight 3*max(nfg,1)/max(nbg,nfg,1), others 0.
__global__ void __launch_bounds__(256)
label_weight_kernel(const float *__restrict__ labels, long A, const int32_t *__restrict__ counts,
                    float *__restrict__ weights, int32_t *__restrict__ counts_out)
{
    const int b = blockIdx.y;
    const int nfg = counts[b * 2], nbg = counts[b * 2 + 1];
    const double num = 3 * fmax((double)nfg, 1.0);
    const double den = fmax((double)max(nbg, nfg), 1.0);
    const float wbg = (float)(num / den);
    if (counts_out && blockIdx.x == 0 && threadIdx.x < 2) counts_out[b * 2 + threadIdx.x] = counts[b * 2 + threadIdx.x];
    const float *l = labels + (size_t)b * A;
    float *w = weights + (size_t)b * A;
    const long A4 = ((A & 3) == 0 && ((size_t)b * A & 3) == 0) ? A / 4 : 0;
    auto wt = [&](float v) { return v >= 1.0f ? 1.0f : (v == 0.0f ? wbg : 0.0f); };
    const long stride = (long)gridDim.x * blockDim.x;
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < A4; i += 4 * stride) {              // four independent 16-byte loads in flight per thread
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = __ldg(reinterpret_cast<const float4 *>(l) + i + k * stride);
#pragma unroll
        for (int k = 0; k < 4; ++k)
            __stcs(reinterpret_cast<float4 *>(w) + i + k * stride, make_float4(wt(v[k].x), wt(v[k].y), wt(v[k].z), wt(v[k].w)));
    }
    for (; i < A4; i += stride) {
        const float4 v = reinterpret_cast<const float4 *>(l)[i];
        reinterpret_cast<float4 *>(w)[i] = make_float4(wt(v.x), wt(v.y), wt(v.z), wt(v.w));
    }
    for (long i = A4 * 4 + (long)blockIdx.x * blockDim.x + threadIdx.x; i < A; i += (long)gridDim.x * blockDim.x) {
        const float v = l[i];
        w[i] = v >= 1.0f ? 1.0f : (v == 0.0f ? wbg : 0.0f);
    }
}

}  // namespace

// detbox_pyramid_desc -> device constants; false when the descriptor does not describe exactly A rows
static bool make_pyr(const detbox_pyramid_desc *d, long A, PyrDev &y)
{
    memset(&y, 0, sizeof(y));
    if (!d || d->n_levels < 1 || d->n_levels > DETBOX_MAX_LEVELS) return false;
    long long off = 0; long long e = 0;
    y.n_levels = d->n_levels;
    for (int l = 0; l < DETBOX_MAX_LEVELS; ++l) {
        y.off[l] = off;
        y.rowlen[l] = y.ncell[l] = 1; y.m_rowlen[l] = y.m_ncell[l] = 0xffffffffu; y.eoff[l] = (int)e;
        if (l >= d->n_levels) continue;
        const long long nc = d->n_cell[l], H = d->height[l], W = d->width[l];
        if (nc < 1 || H < 0 || W < 0) return false;
        if (H == 0 || W == 0) continue;                                   // empty level: no rows, no entries
        const long long rowlen = W * nc, rows = H * rowlen;
        if (rowlen >= (1ll << 31) || rows >= (1ll << 31)) return false;
        y.rowlen[l] = (unsigned)rowlen; y.m_rowlen[l] = (unsigned)(0x100000000ull / (unsigned long long)rowlen > 0xffffffffull ? 0xffffffffull : 0x100000000ull / (unsigned long long)rowlen);
        y.ncell[l] = (unsigned)nc; y.m_ncell[l] = (unsigned)(0x100000000ull / (unsigned long long)nc > 0xffffffffull ? 0xffffffffull : 0x100000000ull / (unsigned long long)nc);
        off += rows;
        e += rowlen + H * nc;
        if (e >= (1ll << 30)) return false;
    }
    y.off[DETBOX_MAX_LEVELS] = off;
    for (int l = d->n_levels; l <= DETBOX_MAX_LEVELS; ++l) y.off[l] = off;
    y.E = (int)e;
    return off == A && e > 0;
}

static size_t target_ws_bytes(long A, int batch, long total_gt, long E)
{
    return align_up((size_t)(total_gt > 0 ? total_gt : 1) * sizeof(unsigned long long)) +
           align_up((size_t)(batch > 0 ? batch : 1) * 2 * sizeof(int32_t)) + align_up(4 * sizeof(int32_t)) +
           align_up((size_t)((A + AT_TILE - 1) / AT_TILE + 1) * sizeof(TileBoxes)) +
           align_up((size_t)(batch > 0 ? batch : 1) * (size_t)E * sizeof(float2));
}

extern "C" size_t detbox_anchor_target_workspace_bytes(long A, int batch, long total_gt)
{
    if (A < 0 || batch < 0 || total_gt < 0) return 0;
    return target_ws_bytes(A, batch, total_gt, 0);
}

extern "C" size_t detbox_anchor_target_ex_workspace_bytes(long A, int batch, long total_gt, const detbox_pyramid_desc *pyr)
{
    if (A < 0 || batch < 0 || total_gt < 0) return 0;
    PyrDev y;
    return target_ws_bytes(A, batch, total_gt, make_pyr(pyr, A, y) ? y.E : 0);
}

extern "C" int detbox_anchor_target_ex(const float *anchors, long A, const detbox_pyramid_desc *pyr,
                                       const float *gt, const int32_t *gt_off, int batch, long total_gt,
                                       const detbox_target_cfg *cfg, int format,
                                       float *labels, float *label_weights, float *bbox_targets,
                                       float *bbox_inside_weights, uint8_t *packed, int32_t *counts_out,
                                       void *ws, size_t ws_bytes, void *stream)
{
    if (A < 0 || batch < 0 || total_gt < 0 || !cfg) return DETBOX_ERR_INVALID;
    if (cfg->box_encoding != DETBOX_ENC_FASTRCNN && cfg->box_encoding != DETBOX_ENC_LINEAR) return DETBOX_ERR_INVALID;
    if (format != DETBOX_TARGETS_REFERENCE && format != DETBOX_TARGETS_COMPACT) return DETBOX_ERR_INVALID;
    const bool compact = format == DETBOX_TARGETS_COMPACT;
    if (A == 0 || batch == 0) return DETBOX_OK;
    if (!anchors || !gt_off || !bbox_targets) return DETBOX_ERR_INVALID;
    if (compact ? (!packed || !counts_out) : (!labels || !label_weights || !bbox_inside_weights)) return DETBOX_ERR_INVALID;
    if (total_gt > 0 && !gt) return DETBOX_ERR_INVALID;
    TargetParams p;
    memset(&p, 0, sizeof(p));
    const bool frcnn = cfg->box_encoding == DETBOX_ENC_FASTRCNN;
    bool tables = false;
    if (pyr) {
        if (!make_pyr(pyr, A, p.pyr)) return DETBOX_ERR_INVALID;        // the descriptor must describe exactly A rows
        static const bool off = getenv("DETBOX_AT_NOTABLE") != nullptr;
        tables = frcnn && total_gt > 0 && !off;
    }
    if (!ws || ws_bytes < target_ws_bytes(A, batch, total_gt, tables ? p.pyr.E : 0)) return DETBOX_ERR_WORKSPACE;
    if (batch > 65535) return DETBOX_ERR_UNSUPPORTED;
    cudaStream_t st = as_stream(stream);
    Carver cv(ws);
    unsigned long long *colmax = cv.take<unsigned long long>(total_gt > 0 ? total_gt : 1);
    int32_t *counts = cv.take<int32_t>((size_t)batch * 2);
    int32_t *flags = cv.take<int32_t>(4);
    DETBOX_CUDA_TRY(cudaMemsetAsync(ws, 0, cv.off, st));
    const unsigned n_tiles = (unsigned)((A + AT_TILE - 1) / AT_TILE);
    TileBoxes *tiles = cv.take<TileBoxes>(n_tiles + 1);
    float2 *tab = tables ? cv.take<float2>((size_t)batch * p.pyr.E) : nullptr;
    if (compact) {                                   // the counts are a result: accumulate straight into the caller's array
        DETBOX_CUDA_TRY(cudaMemsetAsync(counts_out, 0, (size_t)batch * 2 * sizeof(int32_t), st));
        counts = counts_out;
    }

    p.anchors = (const float4 *)anchors; p.A = A; p.gt = gt; p.gt_off = gt_off;
    p.colmax = colmax; p.counts = counts; p.tiles = tiles;
    p.bg = cfg->bg_threshold; p.fg = cfg->fg_threshold; p.ign = cfg->ignore_fraction;
    p.clobber = cfg->clobber_positives; p.enc = cfg->box_encoding;
    p.bbw = make_float4(cfg->bbweights[0], cfg->bbweights[1], cfg->bbweights[2], cfg->bbweights[3]);
    p.labels = labels; p.targets = (float4 *)bbox_targets; p.inside = (float4 *)bbox_inside_weights;
    p.packed = packed; p.tab = tab; p.sep_bad = flags;

    dim3 grid(n_tiles, (unsigned)batch);
    if (total_gt > 0) {
        const unsigned tab_blocks = tables ? (unsigned)batch * (unsigned)((p.pyr.E + AT_THREADS - 1) / AT_THREADS) : 0u;
        anchor_bbox_kernel<<<n_tiles + tab_blocks, AT_THREADS, 0, st>>>((const float4 *)anchors, A, tiles, (int)n_tiles, p.pyr, gt, gt_off, tab, flags);
        DETBOX_LAUNCH_CHECK();
        DETBOX_TRACE_BEGIN(DETBOX_K_TARGET_COLMAX, st);
        const int n_slices = (int)n_tiles * AT_ITEMS * AT_WARPS;
        // slices per block: at most CM_SLICES, fewer when three blocks per SM allow it (CP-R, 64 x 12 GT: 18 x 24 = 432 blocks
        // instead of 16 x 24 = 384 = 2.6 per SM, an uneven single wave).  Measured 0.051 -> 0.047 ms; 4 / 6 / 8 blocks per SM
        // 0.055 / 0.049 / 0.050 ms: more blocks also means more copies of the cull and a weaker running-maximum screen.
        const long cm_gy = (total_gt + CM_GT - 1) / CM_GT;
        long cm_gx = (n_slices + CM_SLICES - 1) / CM_SLICES;
        static const int cm_wave = getenv("DETBOX_CM_WAVE") ? atoi(getenv("DETBOX_CM_WAVE")) : 3;     // blocks per SM aimed at (0: off)
        if (cm_wave > 0 && (long)kNumSMs * cm_wave / cm_gy > cm_gx) cm_gx = (long)kNumSMs * cm_wave / cm_gy;
        if (cm_gx > n_slices) cm_gx = n_slices;
        colmax_kernel<<<dim3((unsigned)cm_gx, (unsigned)cm_gy), CM_THREADS, 0, st>>>(p, n_slices, (int)total_gt);
        DETBOX_TRACE_END(DETBOX_K_TARGET_COLMAX, st);
        DETBOX_LAUNCH_CHECK();
    }
    DETBOX_TRACE_BEGIN(DETBOX_K_TARGET_ASSIGN, st);
    static const bool screen = getenv("DETBOX_AT_NOSCREEN") == nullptr;
#define AT_LAUNCH(ILP, FR, TB, CP) do { if (screen) anchor_target_kernel<ILP, FR, 4, TB, CP, true><<<grid, AT_THREADS, 0, st>>>(p); \
                                        else anchor_target_kernel<ILP, FR, 4, TB, CP, false><<<grid, AT_THREADS, 0, st>>>(p); } while (0)
    if (!frcnn) { if (compact) AT_LAUNCH(1, false, false, true); else AT_LAUNCH(1, false, false, false); }
    else if (tables) { if (compact) AT_LAUNCH(2, true, true, true); else AT_LAUNCH(2, true, true, false); }
    else { if (compact) AT_LAUNCH(2, true, false, true); else AT_LAUNCH(2, true, false, false); }
#undef AT_LAUNCH
    DETBOX_TRACE_END(DETBOX_K_TARGET_ASSIGN, st);
    DETBOX_LAUNCH_CHECK();
    if (compact) return DETBOX_OK;                   // label weights = f(label, counts): the expander's job
    long wb = (A / 4 + 255) / 256;
    if (wb < 1) wb = 1;
    wb = (wb + 3) / 4;                       // 4 float4 per thread per trip
    if (wb < 1) wb = 1;
    DETBOX_TRACE_BEGIN(DETBOX_K_TARGET_WEIGHTS, st);
    label_weight_kernel<<<dim3((unsigned)wb, (unsigned)batch), 256, 0, st>>>(labels, A, counts, label_weights, counts_out);
    DETBOX_TRACE_END(DETBOX_K_TARGET_WEIGHTS, st);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

extern "C" int detbox_anchor_target(const float *anchors, long A, const float *gt, const int32_t *gt_off,
                                    int batch, long total_gt, const detbox_target_cfg *cfg,
                                    float *labels, float *label_weights, float *bbox_targets,
                                    float *bbox_inside_weights, int32_t *counts_out,
                                    void *ws, size_t ws_bytes, void *stream)
{
    if (A > 0 && batch > 0 && (!labels || !label_weights || !bbox_targets || !bbox_inside_weights)) return DETBOX_ERR_INVALID;
    return detbox_anchor_target_ex(anchors, A, nullptr, gt, gt_off, batch, total_gt, cfg, DETBOX_TARGETS_REFERENCE,
                                   labels, label_weights, bbox_targets, bbox_inside_weights, nullptr, counts_out,
                                   ws, ws_bytes, stream);
}

// ---- padded GT -> concatenated rows + offsets (the collate contract of SURVEY 8f-4) -------------------------------
__global__ void __launch_bounds__(1024)
gt_offsets_kernel(const int32_t *__restrict__ counts, int B, int gmax, int32_t *__restrict__ off)
{
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    if (threadIdx.x == 0) { s_carry = 0; off[0] = 0; }
    __syncthreads();
    for (int base = 0; base < B; base += 1024) {
        const int i = base + threadIdx.x;
        int v = i < B ? min(max(counts[i], 0), gmax) : 0;
        int x = v;                                     // inclusive warp scan
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, d); if ((threadIdx.x & 31) >= d) x += y; }
        if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            int w = s_warp[threadIdx.x];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, w, d); if (threadIdx.x >= d) w += y; }
            s_warp[threadIdx.x] = w;
        }
        __syncthreads();
        const int incl = x + ((threadIdx.x >> 5) ? s_warp[(threadIdx.x >> 5) - 1] : 0) + s_carry;
        if (i < B) off[i + 1] = incl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = incl;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256)
gt_pack_kernel(const float *__restrict__ padded, const int32_t *__restrict__ off, int B, int gmax, float *__restrict__ out)
{
    const long total = (long)B * gmax;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int b = (int)(i / gmax), j = (int)(i - (long)b * gmax);
        const int o = off[b];
        if (j >= off[b + 1] - o) continue;
        const float *src = padded + i * 5;
        float *dst = out + (size_t)(o + j) * 5;
#pragma unroll
        for (int k = 0; k < 5; ++k) dst[k] = src[k];
    }
}

extern "C" int detbox_pack_padded_gt(const float *gt_padded, const int32_t *gt_counts, int batch, int gmax,
                                     float *gt_out, int32_t *gt_offsets_out, void *stream)
{
    if (batch < 0 || gmax < 0) return DETBOX_ERR_INVALID;
    if (batch == 0) return DETBOX_OK;
    if (!gt_counts || !gt_offsets_out || (gmax > 0 && (!gt_padded || !gt_out))) return DETBOX_ERR_INVALID;
    cudaStream_t st = as_stream(stream);
    gt_offsets_kernel<<<1, 1024, 0, st>>>(gt_counts, batch, gmax, gt_offsets_out);
    DETBOX_LAUNCH_CHECK();
    if (gmax > 0) {
        DETBOX_CUDA_TRY(cudaMemsetAsync(gt_out, 0, (size_t)batch * gmax * 5 * sizeof(float), st));   // rows past offsets[batch] are unused but finite
        long gb = ((long)batch * gmax + 255) / 256;
        if (gb > kNumSMs * 8) gb = kNumSMs * 8;
        gt_pack_kernel<<<(unsigned)gb, 256, 0, st>>>(gt_padded, gt_offsets_out, batch, gmax, gt_out);
        DETBOX_LAUNCH_CHECK();
    }
    return DETBOX_OK;
}

#if defined(__x86_64__) && defined(__GNUC__)
// Eight anchors per iteration, no table look-ups: label = (byte & 0x7f) - 1 as float, weight by two blends, inside weights by
// lane permutes of the flag mask; 32-byte non-temporal stores.  Requires all three outputs and 32-byte alignment at index i0
// (i0 a multiple of 8 on 32-byte-aligned arrays).  Returns the index it stopped at (a multiple of 8 past i0); *bad is set if a
// label code 127 was met.  Bit-identical to the scalar loop (the same float values are produced).
__attribute__((target("avx2")))
static long expand_avx2(const uint8_t *packed, long i0, long i1, const float wbg, const float *bbweights,
                        float *labels, float *label_weights, float *inside, int *bad)
{
    const __m256i c7f = _mm256_set1_epi32(0x7f), c80 = _mm256_set1_epi32(0x80), one_i = _mm256_set1_epi32(1), zero_i = _mm256_setzero_si256();
    const __m256 wbgv = _mm256_set1_ps(wbg), onev = _mm256_set1_ps(1.0f), zerov = _mm256_setzero_ps();
    const __m256 w2 = _mm256_set_ps(bbweights[3], bbweights[2], bbweights[1], bbweights[0], bbweights[3], bbweights[2], bbweights[1], bbweights[0]);
    const __m256i p01 = _mm256_set_epi32(1, 1, 1, 1, 0, 0, 0, 0), p23 = _mm256_set_epi32(3, 3, 3, 3, 2, 2, 2, 2);
    const __m256i p45 = _mm256_set_epi32(5, 5, 5, 5, 4, 4, 4, 4), p67 = _mm256_set_epi32(7, 7, 7, 7, 6, 6, 6, 6);
    __m256i badv = zero_i;
    long i = i0;
    for (; i + 8 <= i1; i += 8) {
        const __m256i v = _mm256_cvtepu8_epi32(_mm_loadl_epi64(reinterpret_cast<const __m128i *>(packed + i)));
        const __m256i code = _mm256_and_si256(v, c7f);
        badv = _mm256_or_si256(badv, _mm256_cmpeq_epi32(code, c7f));
        const __m256i lab = _mm256_sub_epi32(code, one_i);
        _mm256_stream_ps(labels + i, _mm256_cvtepi32_ps(lab));
        const __m256 m_bg = _mm256_castsi256_ps(_mm256_cmpeq_epi32(lab, zero_i)), m_fg = _mm256_castsi256_ps(_mm256_cmpgt_epi32(lab, zero_i));
        _mm256_stream_ps(label_weights + i, _mm256_blendv_ps(_mm256_blendv_ps(zerov, wbgv, m_bg), onev, m_fg));
        const __m256 fm = _mm256_castsi256_ps(_mm256_cmpeq_epi32(_mm256_and_si256(v, c80), c80));
        float *o = inside + i * 4;
        _mm256_stream_ps(o, _mm256_and_ps(_mm256_permutevar8x32_ps(fm, p01), w2));
        _mm256_stream_ps(o + 8, _mm256_and_ps(_mm256_permutevar8x32_ps(fm, p23), w2));
        _mm256_stream_ps(o + 16, _mm256_and_ps(_mm256_permutevar8x32_ps(fm, p45), w2));
        _mm256_stream_ps(o + 24, _mm256_and_ps(_mm256_permutevar8x32_ps(fm, p67), w2));
    }
    if (!_mm256_testz_si256(badv, badv)) *bad = 1;
    return i;
}
#endif

// HOST: compact format -> the reference's arrays (libs/layers/anchor.py:111-118 label weights, :136-138 inside weights)
extern "C" int detbox_expand_targets_host(const uint8_t *packed, const int32_t *counts, long A, int batch,
                                          const float *bbweights, float *labels, float *label_weights,
                                          float *inside, int n_threads)
{
    if (A < 0 || batch < 0 || !bbweights) return DETBOX_ERR_INVALID;
    if (A == 0 || batch == 0) return DETBOX_OK;
    if (!packed || !counts) return DETBOX_ERR_INVALID;
    if (n_threads < 1) n_threads = 1;
    std::atomic<int> bad(0);
    auto work = [&](long i0, long i1) {               // rows [i0, i1) of the flattened [batch * A]
        long i = i0;
        while (i < i1) {
            const int b = (int)(i / A);
            const long e = std::min(i1, (long)(b + 1) * A);
            const int nfg = counts[b * 2], nbg = counts[b * 2 + 1];
            const float wbg = (float)(3 * std::max((double)nfg, 1.0) / std::max((double)std::max(nbg, nfg), 1.0));
            float lut_l[256], lut_w[256];
            for (int v = 0; v < 256; ++v) {
                const int l = (v & 0x7f) - 1;
                lut_l[v] = (float)l;
                lut_w[v] = l >= 1 ? 1.0f : (l == 0 ? wbg : 0.0f);
            }
            int seen_bad = 0;
#if defined(__x86_64__) && defined(__GNUC__)
            if (labels && label_weights && inside && __builtin_cpu_supports("avx2") &&
                (((size_t)labels | (size_t)label_weights | (size_t)inside) & 31) == 0) {
                for (; i < e && (i & 7); ++i) {                  // scalar head up to a multiple of 8 anchors
                    const unsigned v = packed[i];
                    seen_bad |= ((v & 0x7f) == 127);
                    labels[i] = lut_l[v]; label_weights[i] = lut_w[v];
                    float *o = inside + i * 4;
                    if (v & 0x80) { o[0] = bbweights[0]; o[1] = bbweights[1]; o[2] = bbweights[2]; o[3] = bbweights[3]; }
                    else { o[0] = o[1] = o[2] = o[3] = 0.0f; }
                }
                i = expand_avx2(packed, i, e, wbg, bbweights, labels, label_weights, inside, &seen_bad);
            }
#endif
#if defined(__x86_64__) && defined(__SSE2__)
            // The three arrays are written once and not read here: non-temporal 16-byte stores skip the read-for-ownership
            // of ordinary stores (half the memory traffic of the expansion, which competes with the device->host DMA into the
            // same host memory when run_host(wire="compact") overlaps the two).  Groups of 4 anchors on 16-byte boundaries.
            const bool al = (!labels || ((size_t)labels & 15) == 0) && (!label_weights || ((size_t)label_weights & 15) == 0) &&
                            (!inside || ((size_t)inside & 15) == 0);
            if (al) {
                const __m128 wv = _mm_loadu_ps(bbweights), zv = _mm_setzero_ps();
                for (; i < e && (i & 3); ++i) {                  // head up to a multiple of 4 anchors
                    const unsigned v = packed[i];
                    seen_bad |= ((v & 0x7f) == 127);
                    if (labels) labels[i] = lut_l[v];
                    if (label_weights) label_weights[i] = lut_w[v];
                    if (inside) _mm_stream_ps(inside + i * 4, (v & 0x80) ? wv : zv);
                }
                for (; i + 4 <= e; i += 4) {
                    const unsigned v0 = packed[i], v1 = packed[i + 1], v2 = packed[i + 2], v3 = packed[i + 3];
                    seen_bad |= ((v0 & 0x7f) == 127) | ((v1 & 0x7f) == 127) | ((v2 & 0x7f) == 127) | ((v3 & 0x7f) == 127);
                    if (labels) _mm_stream_ps(labels + i, _mm_set_ps(lut_l[v3], lut_l[v2], lut_l[v1], lut_l[v0]));
                    if (label_weights) _mm_stream_ps(label_weights + i, _mm_set_ps(lut_w[v3], lut_w[v2], lut_w[v1], lut_w[v0]));
                    if (inside) {
                        float *o = inside + i * 4;
                        _mm_stream_ps(o, (v0 & 0x80) ? wv : zv); _mm_stream_ps(o + 4, (v1 & 0x80) ? wv : zv);
                        _mm_stream_ps(o + 8, (v2 & 0x80) ? wv : zv); _mm_stream_ps(o + 12, (v3 & 0x80) ? wv : zv);
                    }
                }
            }
#endif
            for (; i < e; ++i) {
                const unsigned v = packed[i];
                seen_bad |= ((v & 0x7f) == 127);
                if (labels) labels[i] = lut_l[v];
                if (label_weights) label_weights[i] = lut_w[v];
                if (inside) {
                    float *o = inside + i * 4;
                    if (v & 0x80) { o[0] = bbweights[0]; o[1] = bbweights[1]; o[2] = bbweights[2]; o[3] = bbweights[3]; }
                    else { o[0] = o[1] = o[2] = o[3] = 0.0f; }
                }
            }
#if defined(__x86_64__) && defined(__SSE2__)
            _mm_sfence();                                        // non-temporal stores are weakly ordered: publish before the thread ends
#endif
            if (seen_bad) bad.store(1);
        }
    };
    const long total = (long)batch * A;
    if (n_threads == 1 || total < (1 << 16)) work(0, total);
    else {
        std::vector<std::thread> th;
        const long per = (total + n_threads - 1) / n_threads;
        for (int t = 0; t < n_threads; ++t) {
            const long i0 = std::min(total, t * per), i1 = std::min(total, i0 + per);
            if (i0 < i1) th.emplace_back(work, i0, i1);
        }
        for (auto &t : th) t.join();
    }
    return bad.load() ? DETBOX_ERR_UNSUPPORTED : DETBOX_OK;       // a GT class outside [-1, 125] cannot be packed
}
