// nms.cu — greedy non-maximum suppression entirely on the device.
// Replaces: libs/nms/gpu_nms.pyx + nms_kernel.cu:46-155 (bitmask kernel + HOST scan + per-call
// cudaMalloc/Memcpy), libs/nms/cpu_nms.pyx:17-68, libs/layers/box.py:148-173 + layers/nms/src/*,
// and — through segments — the per-image Python loop of libs/nets/model.py:343-355.
//
// Pipeline (one stream, no host round trip, no allocation):
//   1. keys = ~ordered(score), stable sort (sort.cu: one-block bitonic for n <= 4096, else LSD radix)
//      -> score-descending order, ties by index; with segments one more stable pass on the segment
//      id -> segment-major order.  DETBOX_NMS_PRESORTED skips 1 and 2 (rows already in that order).
//   2. gather boxes into sorted float4 rows (+ segment id).
//   3. suppressor bitmask, LOWER triangle only (the reference computes the full square,
//      nms_kernel.cu:51): row i gets one 64-bit word per earlier column tile of its own segment.
//      Broad phase: four sign-bit tests per pair against conservatively rounded column boxes reject
//      the disjoint pairs; narrow phase: the reference's float32 IoU sequence (cpu_nms.pyx:57-64 ==
//      nms_kernel.cu:24-32): (area_i + area_j) - inter, IEEE division, --fmad=false.  The division is
//      only executed when an approximate quotient lands within 2^-20 of the threshold.
//   4. resolve: greedy keep set as a monotone fixed point (kept iff no KEPT earlier suppressor),
//      all boxes in parallel, 4-8 rounds.  Segmented input: one block per segment with the segment's
//      mask rows and bitsets in shared memory; one segment: a cooperative grid with grid.sync
//      between rounds.  The fixed point is unique, so the result is the exact greedy set.
//   5. ordered compaction of the kept bits into original indices + per-segment counts.
#include <mutex>
#include <cooperative_groups.h>
#include "sort.cuh"

namespace cg = cooperative_groups;

using namespace detbox;
typedef unsigned long long u64;

namespace {

struct NmsCmp { int mode; float thr_f; double thr_d; bool nonpos; bool fast; float q_hi, q_lo; int variant; };

__device__ __forceinline__ bool suppresses(const NmsCmp &c, float ovr) {
    if (c.mode == DETBOX_NMS_GE_F64) return (double)ovr >= c.thr_d;      // cpu_nms.pyx:65
    if (c.mode == DETBOX_NMS_GE_F32) return ovr >= c.thr_f;             // nms_cpu.c:76
    return ovr > c.thr_f;                                               // nms_kernel.cu:83
}

// Does the earlier box b suppress box a?  The reference's arithmetic (cpu_nms.pyx:57-64); b plays its
// `i`, a its `j`: (area_i + area_j) - inter.  The IEEE division only runs when the approximate quotient
// (MUFU.RCP * inter, <= 2 ulp off) is within 2^-20 of the threshold: outside that band the correctly
// rounded quotient is on the same side (make_cmp() widens the band by one more ulp).
template <int VARIANT>
__device__ __forceinline__ bool pair_suppresses(const NmsCmp &c, const float4 &a, float aarea, const float4 &b, float barea)
{
    const float xx1 = fmaxf(a.x, b.x), yy1 = fmaxf(a.y, b.y);
    const float xx2 = fminf(a.z, b.z), yy2 = fminf(a.w, b.w);
    const float w = fmaxf(0.0f, xx2 - xx1 + 1), h = fmaxf(0.0f, yy2 - yy1 + 1);
    const float inter = w * h;
    if (VARIANT == DETBOX_NMS_VARIANT_MULTIANCHOR) {
        // libs/boxes/nms.pyx:125-195.  `+ 0.1` promotes to double, the quotient is stored in a float32;
        // then the centre rule: sqrt(d2) <= 0.2 * min(w, h) with Python-float (double) centres.  b is `i`, a is `j`.
        const double od = ((double)inter + 0.1) / ((double)(barea + aarea - inter) + 0.1);
        if (suppresses(c, (float)od)) return true;
        const double icx = 0.5 * (double)(b.x + b.z), icy = 0.5 * (double)(b.y + b.w);
        const double jcx = 0.5 * (double)(a.x + a.z), jcy = 0.5 * (double)(a.y + a.w);
        const double dist = (jcx - icx) * (jcx - icx) + (jcy - icy) * (jcy - icy);
        return sqrt(dist) <= 0.2 * (double)fminf(w, h);
    }
    if (!(inter > 0.0f) && !c.nonpos) return false;        // quotient 0 or NaN: cannot suppress when thr > 0
    const float uni = barea + aarea - inter;
    if (VARIANT == DETBOX_NMS_VARIANT_NEW) {
        // libs/boxes/nms.pyx:70-123: also containment, inter/area_i > 0.95 or inter/area_j > 0.95 (compared as doubles)
        return suppresses(c, inter / uni) || (double)(inter / barea) > 0.95 || (double)(inter / aarea) > 0.95;
    }
    if (c.fast && uni > 1e-30f && uni < 1e30f) {
        // (uni is a normal number here: the bare MUFU.RCP, without __fdividef's rescaling of huge divisors — six
        //  instructions of the ~50 of a candidate pair; the same <= 2 ulp the band below allows for)
        float y;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(uni));
        const float q = inter * y;
        if (q > c.q_hi) return true;
        if (q < c.q_lo) return false;
    }
    return suppresses(c, inter / uni);
}

__global__ void __launch_bounds__(256)
nms_gather_kernel(const float *__restrict__ boxes, int stride, const uint32_t *__restrict__ order, long n,
              const int32_t *__restrict__ seg_ids, float4 *__restrict__ sorted, int32_t *__restrict__ seg_sorted)
{
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const uint32_t o = order ? order[i] : (uint32_t)i;
        const float *b = boxes + (size_t)o * stride;
        sorted[i] = make_float4(b[0], b[1], b[2], b[3]);
        if (seg_sorted) seg_sorted[i] = seg_ids[o];
    }
}

// segment sizes -> offsets in the (segment-major) processing order.  One atomic per distinct id per warp
// (__match_any): ids arrive sorted or in long runs, so a warp usually issues one or two.
// ID = int64_t (DETBOX_NMS_SEG_I64): the ids are narrowed into `narrow` on the way, so every later pass reads int32.
template <typename ID>
__global__ void __launch_bounds__(256)
nms_seg_count_kernel(const ID *__restrict__ seg_ids, long n, int n_seg, int32_t *__restrict__ seg_off, int32_t *__restrict__ narrow)
{
    const int lane = threadIdx.x & 31;
    for (long base = (long)blockIdx.x * blockDim.x; base < n; base += (long)gridDim.x * blockDim.x) {   // block-uniform trip count
        const long i = base + threadIdx.x;
        const ID raw = i < n ? seg_ids[i] : (ID)-1;
        // An id outside [0, n_seg) is a caller error.  It must never be used as an index (seg_off[id] in the later
        // passes): the row is counted in, and handed on as, the nearest valid segment — memory-safe, its keep decision
        // unspecified — and every later pass reads this narrowed array, whatever the width of the caller's ids.
        int s = i < n ? (raw < 0 ? 0 : (raw >= (ID)n_seg ? n_seg - 1 : (int)raw)) : -1;
        if (i < n) narrow[i] = s;
        const unsigned peers = __match_any_sync(0xffffffffu, s);
        if (s >= 0 && lane == __ffs(peers) - 1) atomicAdd(&seg_off[s + 1], __popc(peers));
    }
}
__global__ void nms_seg_scan_kernel(int32_t *seg_off, int n_seg)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        int run = 0;
        for (int s = 0; s <= n_seg; ++s) { run += seg_off[s]; seg_off[s] = run; }
    }
}

// ---- 3. suppressor mask (lower triangle) --------------------------------------------------------
// Row i holds, for every EARLIER box j < i (processing order) of the same segment, bit j%64 of word
// j/64 iff j would suppress i.  A block owns MASK_RT row tiles and visits groups of `cg` column tiles:
// blockIdx.x, + gridDim.x, ... counted from the first tile of the rows' segment; it stops at the diagonal.
// Segments are contiguous in processing order, so "same segment" for an earlier column is just
// "column index >= start of the row's segment".
constexpr int MASK_CT = 4;     // column tiles staged per iteration (one per 64-thread group)
constexpr int MASK_RT = 2;     // row tiles per block: every staged column box is tested against 2 rows per
                               // thread, so the shared-memory read (the limiter at 1 row) is amortised

template <int VARIANT>
__global__ void __launch_bounds__(64 * MASK_CT)
nms_mask_kernel(const float4 *__restrict__ boxes, const int32_t *__restrict__ seg, const int32_t *__restrict__ seg_off,
                int n, int ncol, int cg, NmsCmp cmp, u64 *__restrict__ mask)
{
    const int rt0 = blockIdx.y * MASK_RT;
    const int rt_last = min(rt0 + MASK_RT, ncol) - 1;
    const int row0 = rt0 * 64, row_last = min(n, (rt_last + 1) * 64) - 1;
    int first_ct = 0;                                          // first column tile any row of this block can see
    if (seg) first_ct = seg_off[seg[row0]] >> 6;
    const int ct_base = first_ct & ~(MASK_CT - 1);
    if (ct_base + (int)blockIdx.x * cg > rt_last) return;      // strictly right of the diagonal
    __shared__ float4 cb[64 * MASK_CT];
    __shared__ float4 ccull[64 * MASK_CT];      // (-(x1 - 1 rounded down), -(y1 - 1 down), x2 + 1 up, y2 + 1 up)
    __shared__ float carea[64 * MASK_CT];
    const int tid = threadIdx.x, sub = tid >> 6, rl = tid & 63;
    float4 a[MASK_RT]; float aarea[MASK_RT]; int astart[MASK_RT];
#pragma unroll
    for (int k = 0; k < MASK_RT; ++k) {
        const int r = row0 + k * 64 + rl;
        const float NaN = __int_as_float(0x7fc00000);
        a[k] = make_float4(NaN, NaN, NaN, NaN); aarea[k] = 0.0f; astart[k] = 0;
        if (r < n) {
            a[k] = boxes[r]; aarea[k] = (a[k].z - a[k].x + 1) * (a[k].w - a[k].y + 1);
            if (seg) astart[k] = seg_off[seg[r]];
        }
    }
    for (int ct_lo = ct_base + (int)blockIdx.x * cg; ct_lo <= rt_last; ct_lo += (int)gridDim.x * cg) {
    const int ct_hi = min(rt_last, ct_lo + cg - 1);
    for (int c0 = ct_lo; c0 <= ct_hi; c0 += MASK_CT) {
        __syncthreads();
        {
            const int c = c0 * 64 + tid;
            if (c <= row_last) {
                const float4 b = boxes[c];
                cb[tid] = b;
                ccull[tid] = make_float4(-__fadd_rd(b.x, -1.0f), -__fadd_rd(b.y, -1.0f), __fadd_ru(b.z, 1.0f), __fadd_ru(b.w, 1.0f));
                carea[tid] = (b.z - b.x + 1) * (b.w - b.y + 1);    // cpu_nms.pyx:24
            }
        }
        __syncthreads();
        const int ct = c0 + sub;
        if (ct > rt_last) continue;
        const int cbase = sub * 64;
        // Broad phase (branch-free, fully unrolled): bit i of a row's candidate mask survives unless box i
        // lies strictly beyond one side of the row's box by more than the +1 of the reference's pixel
        // convention.  Exact by monotone rounding: a.x > up(b.x2 + 1) means min(x2) - max(x1) < -1 in real
        // arithmetic, so the reference's fl(fl(min - max) + 1) is <= 0 and the intersection is empty.
        unsigned lo[MASK_RT], hi[MASK_RT];
        float2 naxy[MASK_RT], azw[MASK_RT];
#pragma unroll
        for (int k = 0; k < MASK_RT; ++k) { lo[k] = 0; hi[k] = 0; naxy[k] = make_float2(-a[k].x, -a[k].y); azw[k] = make_float2(a[k].z, a[k].w); }
        // The four strict comparisons are evaluated as sign bits of float differences (FADD runs on the FMA
        // pipe; FSETP would saturate the half-rate ALU pipe): a.x > q.z  <=>  sign(q.z - a.x).  A difference is
        // -0 only for (-0) - (+0), which the directed roundings above exclude (q.z is never -0, q.x never +0; the
        // staged -q.x is therefore never -0);
        // NaN / Inf operands may set a sign bit, but such pairs have an empty intersection (fmaxf(0, NaN) = 0).
#pragma unroll 1
        for (int ib = 0; ib < 32; ib += 8) {         // 8 + 8 column boxes per trip keeps the live registers bounded
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 q0 = ccull[cbase + ib + j], q1 = ccull[cbase + 32 + ib + j];
#pragma unroll
                for (int k = 0; k < MASK_RT; ++k) {
                    // packed float32 adds (FADD2): (q.z, q.w) + (-a.x, -a.y) and (a.z, a.w) + (-q.x, -q.y) — the four
                    // differences of a pair in two instructions, each lane the same IEEE subtraction as before
                    const float2 d0 = __fadd2_rn(make_float2(q0.z, q0.w), naxy[k]), e0 = __fadd2_rn(azw[k], make_float2(q0.x, q0.y));
                    const float2 d1 = __fadd2_rn(make_float2(q1.z, q1.w), naxy[k]), e1 = __fadd2_rn(azw[k], make_float2(q1.x, q1.y));
                    const unsigned o0 = __float_as_uint(d0.x) | __float_as_uint(e0.x) | __float_as_uint(d0.y) | __float_as_uint(e0.y);
                    const unsigned o1 = __float_as_uint(d1.x) | __float_as_uint(e1.x) | __float_as_uint(d1.y) | __float_as_uint(e1.y);
                    lo[k] = (lo[k] >> 1) | (~o0 & 0x80000000u);      // bit 31 <- "not outside"; 32 shifts put box i at bit i
                    hi[k] = (hi[k] >> 1) | (~o1 & 0x80000000u);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < MASK_RT; ++k) {
            const int rt = rt0 + k, r = rt * 64 + rl;
            const int lb = astart[k] - ct * 64;                    // columns below bit lb belong to earlier segments
            if (r >= n || ct > rt || lb >= 64) continue;
            u64 cand = ((u64)hi[k] << 32) | lo[k];
            if (cmp.nonpos || VARIANT == DETBOX_NMS_VARIANT_MULTIANCHOR) cand = ~0ull;   // empty intersections matter too
            if (ct == rt) cand &= (1ull << rl) - 1ull;             // diagonal tile: only j < i
            if (lb > 0) cand &= ~0ull << lb;
            // narrow phase, one 32-bit half at a time (32-bit find-first-set / clear-lowest are 1/3 of the 64-bit cost).
            // Plain variant: a branch-free first pass decides every candidate whose approximate quotient is outside the
            // band around the threshold (pair_suppresses' own rule, same arithmetic) and collects the others in `und`;
            // the full routine then runs for those only.  On clustered detections a third of all pairs reach this loop
            // (profiles/r02_nms_mask_experiments.txt), so its length is the kernel's: boxes by value, no early returns.
            const bool fast = VARIANT == DETBOX_NMS_VARIANT_PLAIN && cmp.fast && !cmp.nonpos;
            unsigned wlo = 0, whi = 0;
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
                const unsigned cw = hf ? (unsigned)(cand >> 32) : (unsigned)cand;
                const int col0 = cbase + hf * 32;
                unsigned res = 0, und = cw;
                if (fast) {
                    und = 0;
                    const float4 ar = a[k]; const float aa = aarea[k];
                    const float qhi = cmp.q_hi, qlo = cmp.q_lo;
                    for (unsigned c32 = cw; c32; c32 &= c32 - 1) {
                        const int i = __ffs((int)c32) - 1;
                        const float4 b = cb[col0 + i];
                        const float ba = carea[col0 + i];
                        const float w = fmaxf(0.0f, fminf(ar.z, b.z) - fmaxf(ar.x, b.x) + 1), h = fmaxf(0.0f, fminf(ar.w, b.w) - fmaxf(ar.y, b.y) + 1);
                        const float inter = w * h;
                        const float uni = ba + aa - inter;                  // (area_i + area_j) - inter, b is `i`
                        float y;
                        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(uni));
                        const float q = inter * y;
                        const bool rng = uni > 1e-30f && uni < 1e30f;
                        const bool yes = q > qhi, no = q < qlo;      // (yes implies inter > 0: q_hi > 0)
                        const unsigned bit = 1u << i;
                        if (rng & yes) res |= bit;
                        if (!(rng & (yes | no)) & (inter > 0.0f)) und |= bit;
                    }
                }
                for (; und; und &= und - 1) {
                    const int i = __ffs((int)und) - 1;
                    if (pair_suppresses<VARIANT>(cmp, a[k], aarea[k], cb[col0 + i], carea[col0 + i])) res |= 1u << i;
                }
                if (hf) whi = res; else wlo = res;
            }
            mask[(size_t)r * ncol + ct] = ((u64)whi << 32) | wlo;
        }
    }
    }
}

// ---- 4. resolve: the greedy keep set as a fixed point, all boxes in parallel ----------------------
// Greedy NMS keeps box i iff no KEPT earlier box suppresses it.  Iterate over the still undecided
// boxes: suppressed as soon as one suppressor is known kept; kept as soon as every suppressor is
// known suppressed.  Decisions are final and the lowest undecided box is always decidable, so the
// iteration ends in the exact greedy set; on detection-shaped inputs it takes 4-6 rounds and reads
// the mask ~1.2-2x (the reference scans it once, sequentially, on the host: nms_kernel.cu:136-151).
// One warp per box scans the row's words (coalesced); kept / suppressed are global bitsets updated
// with atomicOr and read through L2 (monotone, so a stale read only delays a decision).
__global__ void __launch_bounds__(256)
nms_resolve_kernel(const u64 *__restrict__ mask, int n, int ncol, const int32_t *__restrict__ seg,
                   const int32_t *__restrict__ seg_off, u64 *kept, u64 *supp, unsigned *cnt /*[3]*/)
{
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (int round = 0;; ++round) {
        if (blockIdx.x == 0 && threadIdx.x == 0) cnt[(round + 1) % 3] = 0;
        unsigned undecided = 0;
        for (int i = gw; i < n; i += nw) {
            const int wi = i >> 6;
            const u64 bit = 1ull << (i & 63);
            if ((__ldcg(&kept[wi]) | __ldcg(&supp[wi])) & bit) continue;
            const int c_lo = seg ? (seg_off[seg[i]] >> 6) : 0;
            bool hit = false, open = false;
            const u64 *row = mask + (size_t)i * ncol;
            for (int c = c_lo + lane; c <= wi; c += 32) {
                const u64 sw = __ldg(&row[c]);
                if (sw) {
                    const u64 k = __ldcg(&kept[c]), d = __ldcg(&supp[c]);
                    hit |= (sw & k) != 0ull;
                    open |= (sw & ~(k | d)) != 0ull;
                }
            }
            hit = __any_sync(0xffffffffu, hit);
            open = __any_sync(0xffffffffu, open);
            if (lane == 0) {
                if (hit) atomicOr(&supp[wi], bit);
                else if (!open) atomicOr(&kept[wi], bit);
                else ++undecided;
            }
        }
        if (lane == 0 && undecided) atomicAdd(&cnt[round % 3], undecided);
        grid.sync();
        if (__ldcg(&cnt[round % 3]) == 0u) break;
    }
}

// ---- 4b. resolve for segmented input: one block per segment, state in shared memory --------------
// Same fixed point as above, but a segment (an image's top-k) is small: its mask rows (count x words
// spanned, <= ~200 KB up to ~1200 boxes) and both bitsets live in shared memory, a round ends in a
// __syncthreads instead of a grid.sync, and no atomics are needed — warps own aligned 32-box groups, so
// a ballot is the group's 32-bit half-word of the bitset.  Segments too large for shared memory read
// their rows from global memory (L2) instead.
constexpr int RSEG_THREADS = 1024;

__global__ void __launch_bounds__(RSEG_THREADS)
nms_resolve_seg_kernel(const u64 *__restrict__ mask, int n, int ncol, const int32_t *__restrict__ seg_off,
                       u64 *__restrict__ kept_g, int smem_words)
{
    extern __shared__ u64 rsm[];
    const int b0 = seg_off ? seg_off[blockIdx.x] : 0, b1 = seg_off ? seg_off[blockIdx.x + 1] : n;   // no offsets: one segment = everything
    if (b1 <= b0) return;
    const int w0 = b0 >> 6, w = ((b1 - 1) >> 6) - w0 + 1, cnt = b1 - b0;
    const int ws = w | 1;                                  // odd row stride: 64-bit column reads hit distinct banks
    u64 *kept = rsm, *supp = rsm + w, *rows = rsm + 2 * w;
    const bool cached = (long)cnt * ws + 2L * w <= (long)smem_words;
    const int tid = threadIdx.x, lane = tid & 31;
    for (int c = tid; c < 2 * w; c += RSEG_THREADS) rsm[c] = 0ull;
    if (cached) {
        for (int e = tid; e < cnt * w; e += RSEG_THREADS) {
            const int t = e / w, c = e - t * w;
            const int i = b0 + t;
            rows[t * ws + c] = (w0 + c <= (i >> 6)) ? __ldg(&mask[(size_t)i * ncol + w0 + c]) : 0ull;
        }
    }
    const u64 *rbase = cached ? rows : mask + (size_t)b0 * ncol + w0;
    const size_t rstride = cached ? (size_t)ws : (size_t)ncol;
    volatile unsigned *k32 = reinterpret_cast<volatile unsigned *>(kept);
    volatile unsigned *d32 = reinterpret_cast<volatile unsigned *>(supp);
    volatile u64 *kv = kept, *dv = supp;
    __syncthreads();
    const int i_lo = b0 & ~31, i_hi = (b1 + 31) & ~31;
    for (;;) {
        bool undecided = false;
        for (int i = i_lo + tid; i < i_hi; i += RSEG_THREADS) {       // warp-uniform trip count
            const int hw = (i >> 5) - 2 * w0;                         // the warp's half-word of the bitsets
            const bool live = i >= b0 && i < b1 && !(((k32[hw] | d32[hw]) >> (i & 31)) & 1u);
            bool hit = false, open = false;
            if (live) {
                const u64 *row = rbase + (size_t)(i - b0) * rstride;
                const int wi = (i >> 6) - w0;
                for (int c = 0; c <= wi; ++c) {
                    const u64 sw = row[c];
                    if (sw) {
                        const u64 k = kv[c], d = dv[c];
                        hit |= (sw & k) != 0ull;
                        open |= (sw & ~(k | d)) != 0ull;
                    }
                }
            }
            const unsigned mk = __ballot_sync(0xffffffffu, live && !hit && !open);
            const unsigned ms = __ballot_sync(0xffffffffu, live && hit);
            if (lane == 0) { if (mk) k32[hw] |= mk; if (ms) d32[hw] |= ms; }
            undecided |= live && !hit && open;
        }
        if (!__syncthreads_or(undecided)) break;
    }
    for (int c = tid; c < w; c += RSEG_THREADS) if (kept[c]) atomicOr(&kept_g[w0 + c], kept[c]);
}

// nms_multianchor's `clusters` (nms.pyx:150, 177, 184): a kept box is its own cluster, a suppressed box
// belongs to the kept box that suppressed it in the greedy order = its FIRST kept suppressor.
__global__ void __launch_bounds__(256)
nms_cluster_kernel(const u64 *__restrict__ mask, const u64 *__restrict__ kept, int n, int ncol,
                   const uint32_t *__restrict__ order, const int32_t *__restrict__ seg, const int32_t *__restrict__ seg_off,
                   int64_t *__restrict__ clusters)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t self = order ? order[i] : (uint32_t)i;
    int64_t cl = self;
    const int wi = i >> 6;
    if (!((kept[wi] >> (i & 63)) & 1ull)) {
        const u64 *row = mask + (size_t)i * ncol;
        for (int c = seg ? (seg_off[seg[i]] >> 6) : 0; c <= wi; ++c) {
            const u64 hit = row[c] & kept[c];
            if (hit) { const int p = c * 64 + __ffsll((long long)hit) - 1; cl = order ? order[p] : (uint32_t)p; break; }
        }
    }
    clusters[self] = cl;
}

// Ordered compaction of the kept bits -> original indices, and the per-segment counts.  A block owns CP_WORDS words
// (2048 boxes): it counts the kept boxes in front of its words (every block re-reads at most ncol words — cheaper than
// a second launch), ranks its own words with one warp scan, and then every thread places ONE box, so the dependent
// order[] loads of a word's 64 boxes overlap instead of queueing behind each other in one thread.
// Per-segment count = rank(end) - rank(begin): the block that owns a boundary adds / subtracts its rank (integer
// atomics on the zeroed n_keep — order-independent).
constexpr int CP_WORDS = 32;

__global__ void __launch_bounds__(1024)
nms_compact_kernel(const u64 *__restrict__ kept, int n, int ncol, const uint32_t *__restrict__ order,
                   const int32_t *__restrict__ seg_off, int n_seg, int64_t *__restrict__ keep, int32_t *__restrict__ n_keep)
{
    __shared__ int warp_sum[32];
    __shared__ int s_rank[CP_WORDS + 1];
    __shared__ u64 s_word[CP_WORDS];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int w_lo = blockIdx.x * CP_WORDS;
    const bool last = blockIdx.x == gridDim.x - 1;
    int c = 0;
    for (int i = tid; i < w_lo; i += 1024) c += __popcll(kept[i]);
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0) warp_sum[w] = c;
    __syncthreads();
    if (w == 0) {
        const int base = __reduce_add_sync(0xffffffffu, warp_sum[lane]);
        const int wi = w_lo + lane;
        const u64 word = wi < ncol ? kept[wi] : 0ull;
        const int pc = __popcll(word);
        int incl = pc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        s_word[lane] = word;
        s_rank[lane] = base + incl - pc;                      // kept boxes before this word
        if (lane == 31) s_rank[CP_WORDS] = base + incl;
    }
    __syncthreads();
    for (int t = tid; t < CP_WORDS * 64; t += 1024) {
        const int wl = t >> 6, bit = t & 63;
        const u64 word = s_word[wl];
        if ((word >> bit) & 1ull) {
            const int sp = (w_lo + wl) * 64 + bit;
            keep[s_rank[wl] + __popcll(word & ((1ull << bit) - 1ull))] = (int64_t)(order ? order[sp] : (uint32_t)sp);
        }
    }
    if (last && tid == 0) { n_keep[n_seg] = s_rank[CP_WORDS]; if (n_seg == 1) n_keep[0] = s_rank[CP_WORDS]; }
    if (n_seg > 1) {
        for (int s = tid; s <= n_seg; s += 1024) {
            const int p = seg_off[s];
            const int wl = (p >> 6) - w_lo;
            if (wl < 0 || (wl >= CP_WORDS && !last)) continue;    // another block's boundary
            const int r = wl >= CP_WORDS ? s_rank[CP_WORDS] : s_rank[wl] + __popcll(s_word[wl] & ((1ull << (p & 63)) - 1ull));
            if (s < n_seg) atomicSub(&n_keep[s], r);
            if (s > 0) atomicAdd(&n_keep[s - 1], r);
        }
    }
}

struct NmsBuffers {
    uint32_t *ka, *kb, *va, *vb, *counts;
    float4 *sorted; int32_t *seg_sorted, *seg_narrow, *seg_off; u64 *mask, *kept, *supp; unsigned *cnt;
    size_t total;
};

NmsBuffers carve_nms(void *ws, long n, int n_seg) {
    Carver cv(ws);
    NmsBuffers b;
    const long ncol = (n + 63) / 64;
    b.ka = cv.take<uint32_t>(n); b.kb = cv.take<uint32_t>(n);
    b.va = cv.take<uint32_t>(n); b.vb = cv.take<uint32_t>(n);
    b.counts = cv.take<uint32_t>(sort_counts_bytes(n) / 4);
    b.sorted = cv.take<float4>(n);
    b.seg_sorted = cv.take<int32_t>(n_seg > 1 ? n : 1);
    b.seg_narrow = cv.take<int32_t>(n_seg > 1 ? n : 1);
    b.seg_off = cv.take<int32_t>(n_seg + 2);
    b.kept = cv.take<u64>(2 * ncol + 2);           // kept | supp | round counters, cleared together
    b.supp = b.kept + ncol;
    b.cnt = reinterpret_cast<unsigned *>(b.supp + ncol);
    b.mask = cv.take<u64>((size_t)n * ncol);
    b.total = cv.off;
    return b;
}

NmsCmp make_cmp(double thresh, int mode, int variant = 0) {
    NmsCmp c; c.mode = mode; c.variant = variant; c.thr_d = thresh; c.thr_f = (float)thresh;
    c.nonpos = (mode == DETBOX_NMS_GE_F64) ? !(thresh > 0.0) : !(c.thr_f > 0.0f);
    // band around the threshold outside which the approximate quotient decides (pair_suppresses)
    const double t = (mode == DETBOX_NMS_GE_F64) ? thresh : (double)c.thr_f;
    c.fast = variant == 0 && t > 1e-20 && t < 1e20;
    c.q_hi = c.q_lo = 0.0f;
    if (c.fast) {
        const double eps = 1.0 / (1 << 20);
        c.q_hi = nextafterf((float)(t * (1.0 + eps)), INFINITY);
        c.q_lo = nextafterf((float)(t * (1.0 - eps)), 0.0f);
    }
    return c;
}

// mask + scan + compact over boxes that are already gathered in processing order
int nms_sorted(const NmsBuffers &b, const uint32_t *order, long n, int n_seg,
               NmsCmp cmp, int64_t *keep, int32_t *n_keep, cudaStream_t st, int64_t *clusters = nullptr)
{
    const int32_t *seg_off = b.seg_off;
    const int ncol = (int)((n + 63) / 64);
    const int32_t *seg_sorted = n_seg > 1 ? b.seg_sorted : nullptr;
    DETBOX_CUDA_TRY(cudaMemsetAsync(b.kept, 0, (size_t)(2 * ncol + 2) * sizeof(u64), st));
    DETBOX_CUDA_TRY(cudaMemsetAsync(n_keep, 0, (size_t)(n_seg + 1) * sizeof(int32_t), st));
    // column tiles per visit / visits in flight: whole 32-tile groups for one big triangle; single 4-tile
    // chunks for segmented input, whose rows see only the handful of tiles of their own segment
    const bool small_segs = seg_sorted && n / n_seg <= 4096;
    const int cg = small_segs ? MASK_CT : 32;
    unsigned gx = (unsigned)((ncol + cg - 1) / cg);
    // (any gx is correct — a block strides over its row tiles' column chunks — but every block right of the diagonal
    //  costs two dependent loads before it can exit: ceil(L / 256) chunks + 2 for the alignment of the first chunk and
    //  a row-tile pair that straddles two segments; the grid used to be twice that, 3,400 of 4,500 blocks idle at 64 x 900)
    if (small_segs) { const unsigned want = (unsigned)((n / n_seg + 255) / 256) + 2u; if (gx > want) gx = want; }
    dim3 grid(gx, (unsigned)((ncol + MASK_RT - 1) / MASK_RT));
    DETBOX_TRACE_BEGIN(DETBOX_K_NMS_MASK, st);
    if (cmp.variant == DETBOX_NMS_VARIANT_PLAIN)
        nms_mask_kernel<0><<<grid, 64 * MASK_CT, 0, st>>>(b.sorted, seg_sorted, seg_sorted ? seg_off : nullptr, (int)n, ncol, cg, cmp, b.mask);
    else if (cmp.variant == DETBOX_NMS_VARIANT_NEW)
        nms_mask_kernel<1><<<grid, 64 * MASK_CT, 0, st>>>(b.sorted, seg_sorted, seg_sorted ? seg_off : nullptr, (int)n, ncol, cg, cmp, b.mask);
    else
        nms_mask_kernel<2><<<grid, 64 * MASK_CT, 0, st>>>(b.sorted, seg_sorted, seg_sorted ? seg_off : nullptr, (int)n, ncol, cg, cmp, b.mask);
    DETBOX_TRACE_END(DETBOX_K_NMS_MASK, st);
    DETBOX_LAUNCH_CHECK();
    int dev = 0;
    DETBOX_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return DETBOX_ERR_UNSUPPORTED;
    const long wmax = (n >> 6) + 2, smem_cap = 200 * 1024 / 8;      // 64-bit words
    const bool one_small = !seg_sorted && n * (wmax | 1) + 2 * wmax <= smem_cap;   // a single segment whose rows fit one block's smem
    if ((small_segs && 2 * wmax <= smem_cap) || one_small) {
        // one block per segment; shared memory sized for the largest segment n allows, capped at 200 KB
        // (the two bitsets always fit; rows that do not are read from global memory)
        static thread_local bool smem_set[64] = {false};
        long words = n * (wmax | 1) + 2 * wmax;
        if (words > smem_cap) words = smem_cap;
        const int bytes = (int)(words * 8);
        if (!smem_set[dev]) {
            DETBOX_CUDA_TRY(cudaFuncSetAttribute(nms_resolve_seg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(smem_cap * 8)));
            smem_set[dev] = true;
        }
        DETBOX_TRACE_BEGIN(DETBOX_K_NMS_SCAN, st);
        nms_resolve_seg_kernel<<<(unsigned)n_seg, RSEG_THREADS, (size_t)bytes, st>>>(b.mask, (int)n, ncol, seg_sorted ? seg_off : nullptr, b.kept, (int)words);
        DETBOX_TRACE_END(DETBOX_K_NMS_SCAN, st);
        DETBOX_LAUNCH_CHECK();
    } else {
    // cooperative launch: every block must be resident (grid.sync between rounds)
    static thread_local int coop_blocks[64] = {0};
    if (coop_blocks[dev] == 0) {
        int per_sm = 0, sms = 0, coop = 0;
        DETBOX_CUDA_TRY(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
        if (!coop) return DETBOX_ERR_UNSUPPORTED;
        DETBOX_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, nms_resolve_kernel, 256, 0));
        DETBOX_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        coop_blocks[dev] = per_sm * sms > 0 ? per_sm * sms : 1;
    }
    {
        long want = (n + 7) / 8;                           // 8 warps per block, one box per warp per sweep
        int blocks = (int)(want < coop_blocks[dev] ? want : coop_blocks[dev]);
        const u64 *mask_c = b.mask; int n_i = (int)n, ncol_i = ncol;
        const int32_t *seg_c = seg_sorted, *off_c = seg_sorted ? seg_off : nullptr;
        u64 *kept = b.kept, *supp = b.supp; unsigned *cnt = b.cnt;
        void *args[] = {&mask_c, &n_i, &ncol_i, &seg_c, &off_c, &kept, &supp, &cnt};
        DETBOX_TRACE_BEGIN(DETBOX_K_NMS_SCAN, st);
        DETBOX_CUDA_TRY(cudaLaunchCooperativeKernel((const void *)nms_resolve_kernel, dim3(blocks), dim3(256), args, 0, st));
        DETBOX_TRACE_END(DETBOX_K_NMS_SCAN, st);
    }
    }
    nms_compact_kernel<<<(unsigned)((ncol + CP_WORDS - 1) / CP_WORDS), 1024, 0, st>>>(b.kept, (int)n, ncol, order, seg_sorted ? seg_off : nullptr, n_seg, keep, n_keep);
    DETBOX_LAUNCH_CHECK();
    if (clusters) {
        nms_cluster_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(b.mask, b.kept, (int)n, ncol, order, seg_sorted,
                                                                        seg_sorted ? seg_off : nullptr, clusters);
        DETBOX_LAUNCH_CHECK();
    }
    return DETBOX_OK;
}

}  // namespace

extern "C" size_t detbox_nms_workspace_bytes(long n, int n_seg)
{
    if (n <= 0) return 256;
    return carve_nms(nullptr, n, n_seg > 1 ? n_seg : 1).total;
}

static int nms_impl(const float *boxes, int box_stride, const float *scores, long n,
                    const int32_t *seg_ids, int n_seg, double thresh, int cmp_mode, int variant,
                    int64_t *keep, int32_t *n_keep, int64_t *clusters, void *ws, size_t ws_bytes, void *stream)
{
    if (variant < 0 || variant > 2) return DETBOX_ERR_INVALID;
    const bool presorted = (cmp_mode & DETBOX_NMS_PRESORTED) != 0;
    const bool seg_i64 = (cmp_mode & DETBOX_NMS_SEG_I64) != 0;
    cmp_mode &= ~(DETBOX_NMS_PRESORTED | DETBOX_NMS_SEG_I64);
    if (n < 0 || box_stride < 4 || cmp_mode < 0 || cmp_mode > 2 || !n_keep) return DETBOX_ERR_INVALID;
    if (presorted && (box_stride != 4 || ((uintptr_t)boxes & 15))) return DETBOX_ERR_INVALID;
    if (!seg_ids || n_seg <= 1) { n_seg = 1; seg_ids = nullptr; }
    cudaStream_t st = as_stream(stream);
    if (n == 0) { DETBOX_CUDA_TRY(cudaMemsetAsync(n_keep, 0, (size_t)(n_seg + 1) * sizeof(int32_t), st)); return DETBOX_OK; }
    if (n >= (1L << 31) - 64) return DETBOX_ERR_UNSUPPORTED;
    if (!boxes || (!scores && !presorted) || !keep) return DETBOX_ERR_INVALID;
    if (!ws || ws_bytes < detbox_nms_workspace_bytes(n, n_seg)) return DETBOX_ERR_WORKSPACE;
    NmsBuffers b = carve_nms(ws, n, n_seg);
    long gb0 = (n + 255) / 256; if (gb0 > kNumSMs * 8) gb0 = kNumSMs * 8;
    auto count_segments = [&]() -> int {          // segment offsets; int64 ids are narrowed into the workspace
        DETBOX_CUDA_TRY(cudaMemsetAsync(b.seg_off, 0, (size_t)(n_seg + 2) * sizeof(int32_t), st));
        if (seg_i64) {
            nms_seg_count_kernel<int64_t><<<(unsigned)gb0, 256, 0, st>>>(reinterpret_cast<const int64_t *>(seg_ids), n, n_seg, b.seg_off, b.seg_narrow);
        } else {
            nms_seg_count_kernel<int32_t><<<(unsigned)gb0, 256, 0, st>>>(seg_ids, n, n_seg, b.seg_off, b.seg_narrow);
        }
        seg_ids = b.seg_narrow;                       // narrowed AND clamped: the only id array the later passes index with
        DETBOX_LAUNCH_CHECK();
        nms_seg_scan_kernel<<<1, 32, 0, st>>>(b.seg_off, n_seg);
        DETBOX_LAUNCH_CHECK();
        return DETBOX_OK;
    };
    if (presorted) {                              // rows are in processing order: no sort, no gather
        if (n_seg > 1) { const int rcs = count_segments(); if (rcs != DETBOX_OK) return rcs; }
        b.sorted = const_cast<float4 *>(reinterpret_cast<const float4 *>(boxes));
        b.seg_sorted = const_cast<int32_t *>(seg_ids);
        return nms_sorted(b, nullptr, n, n_seg, make_cmp(thresh, cmp_mode, variant), keep, n_keep, st, clusters);
    }
    int rc;
    if (n <= SMALL_SORT_MAX && n_seg == 1) {      // one launch instead of 13
        rc = small_sort_scores_desc(scores, n, b.va, st);
        if (rc != DETBOX_OK) return rc;
    } else {
        rc = make_score_keys(scores, n, b.ka, b.va, st);
        if (rc != DETBOX_OK) return rc;
        rc = radix_sort_pairs(b.ka, b.va, b.kb, b.vb, n, b.counts, st);
        if (rc != DETBOX_OK) return rc;
    }
    uint32_t *order = b.va;
    long gb = (n + 255) / 256; if (gb > kNumSMs * 8) gb = kNumSMs * 8;
    if (n_seg > 1) {                              // stable grouping by segment: 1 pass per 8 bits of id
        rc = count_segments();
        if (rc != DETBOX_OK) return rc;
        uint32_t *ki = b.ka, *vi = b.va, *ko = b.kb, *vo = b.vb;
        for (int shift = 0; (n_seg - 1) >> shift; shift += 8) {
            DigitSrc d{2, shift, 1u, seg_ids};
            rc = radix_pass(ki, vi, ko, vo, n, d, b.counts, st);
            if (rc != DETBOX_OK) return rc;
            uint32_t *t = ki; ki = ko; ko = t; t = vi; vi = vo; vo = t;
        }
        order = vi;
    }
    nms_gather_kernel<<<(unsigned)gb, 256, 0, st>>>(boxes, box_stride, order, n, seg_ids, b.sorted,
                                               n_seg > 1 ? b.seg_sorted : nullptr);
    DETBOX_LAUNCH_CHECK();
    return nms_sorted(b, order, n, n_seg, make_cmp(thresh, cmp_mode, variant), keep, n_keep, st, clusters);
}

extern "C" int detbox_nms(const float *boxes, int box_stride, const float *scores, long n,
                          const int32_t *seg_ids, int n_seg, double thresh, int cmp_mode,
                          int64_t *keep, int32_t *n_keep, void *ws, size_t ws_bytes, void *stream)
{
    return nms_impl(boxes, box_stride, scores, n, seg_ids, n_seg, thresh, cmp_mode, DETBOX_NMS_VARIANT_PLAIN,
                    keep, n_keep, nullptr, ws, ws_bytes, stream);
}

extern "C" int detbox_nms_variant(const float *boxes, int box_stride, const float *scores, long n,
                                  const int32_t *seg_ids, int n_seg, double thresh, int cmp_mode, int variant,
                                  int64_t *keep, int32_t *n_keep, int64_t *clusters,
                                  void *ws, size_t ws_bytes, void *stream)
{
    return nms_impl(boxes, box_stride, scores, n, seg_ids, n_seg, thresh, cmp_mode, variant,
                    keep, n_keep, clusters, ws, ws_bytes, stream);
}

// ------------------------------------------------------------------------------------------------
// host-buffer entry points
// ------------------------------------------------------------------------------------------------
namespace {
struct Scratch { void *ptr = nullptr; size_t size = 0; };
std::mutex g_mu;
Scratch g_scratch[64];

int get_scratch(int dev, size_t bytes, void **out) {
    if (dev < 0 || dev >= 64) return DETBOX_ERR_INVALID;
    Scratch &s = g_scratch[dev];
    if (s.size < bytes) {
        if (s.ptr) cudaFree(s.ptr);
        s.ptr = nullptr; s.size = 0;
        const size_t want = bytes + bytes / 4;
        DETBOX_CUDA_TRY(cudaMalloc(&s.ptr, want));
        s.size = want;
    }
    *out = s.ptr;
    return DETBOX_OK;
}
}  // namespace

namespace detbox {
int host_scratch(int dev, size_t bytes, void **out) { return get_scratch(dev, bytes, out); }
std::mutex &host_mutex() { return g_mu; }
}

static int nms_host_impl(const float *dets, int n, int dim, bool presorted, double thresh, int cmp_mode,
                         int *keep_out, int *num_out, int device_id)
{
    if (n < 0 || dim < 4 + (presorted ? 0 : 1) || !num_out) return DETBOX_ERR_INVALID;
    *num_out = 0;
    if (n == 0) return DETBOX_OK;
    if (!dets || !keep_out) return DETBOX_ERR_INVALID;
    std::lock_guard<std::mutex> lock(g_mu);
    int cur = 0;
    DETBOX_CUDA_TRY(cudaGetDevice(&cur));
    if (cur != device_id) DETBOX_CUDA_TRY(cudaSetDevice(device_id));
    const size_t in_bytes = align_up((size_t)n * dim * sizeof(float));
    const size_t sc_bytes = align_up((size_t)n * sizeof(float));
    const size_t keep_bytes = align_up((size_t)n * sizeof(int64_t));
    const size_t ws_bytes = detbox_nms_workspace_bytes(n, 1);
    void *base = nullptr;
    int rc = get_scratch(device_id, in_bytes + sc_bytes + keep_bytes + 256 + ws_bytes, &base);
    if (rc == DETBOX_OK) {
        char *p = (char *)base;
        float *d_in = (float *)p; p += in_bytes;
        float *d_sc = (float *)p; p += sc_bytes;
        int64_t *d_keep = (int64_t *)p; p += keep_bytes;
        int32_t *d_nk = (int32_t *)p; p += 256;
        void *ws = p;
        cudaStream_t st = 0;
        auto body = [&]() -> int {
            DETBOX_CUDA_TRY(cudaMemcpyAsync(d_in, dets, (size_t)n * dim * sizeof(float), cudaMemcpyHostToDevice, st));
            if (presorted) {
                NmsBuffers b = carve_nms(ws, n, 1);
                long gb = (n + 255) / 256; if (gb > kNumSMs * 8) gb = kNumSMs * 8;
                nms_gather_kernel<<<(unsigned)gb, 256, 0, st>>>(d_in, dim, nullptr, n, nullptr, b.sorted, nullptr);
                DETBOX_LAUNCH_CHECK();
                int r = nms_sorted(b, nullptr, n, 1, make_cmp(thresh, cmp_mode), d_keep, d_nk, st);
                if (r != DETBOX_OK) return r;
            } else {
                DETBOX_CUDA_TRY(cudaMemcpy2DAsync(d_sc, sizeof(float), d_in + 4, (size_t)dim * sizeof(float),
                                                  sizeof(float), n, cudaMemcpyDeviceToDevice, st));
                int r = detbox_nms(d_in, dim, d_sc, n, nullptr, 1, thresh, cmp_mode, d_keep, d_nk, ws, ws_bytes, st);
                if (r != DETBOX_OK) return r;
            }
            int32_t nk[2] = {0, 0};
            DETBOX_CUDA_TRY(cudaMemcpyAsync(nk, d_nk, sizeof(nk), cudaMemcpyDeviceToHost, st));
            DETBOX_CUDA_TRY(cudaStreamSynchronize(st));
            if (nk[0] > 0) {
                int64_t *tmp = (int64_t *)malloc((size_t)nk[0] * sizeof(int64_t));
                if (!tmp) return DETBOX_ERR_INVALID;
                cudaError_t e = cudaMemcpy(tmp, d_keep, (size_t)nk[0] * sizeof(int64_t), cudaMemcpyDeviceToHost);
                if (e == cudaSuccess) for (int i = 0; i < nk[0]; ++i) keep_out[i] = (int)tmp[i];
                free(tmp);
                DETBOX_CUDA_TRY(e);
            }
            *num_out = nk[0];
            return DETBOX_OK;
        };
        rc = body();
    }
    if (cur != device_id) cudaSetDevice(cur);
    return rc;
}

// The reference's FFI symbol (libs/nms/gpu_nms.hpp:1-2).  Errors cannot be returned through this
// signature; like the reference (CUDA_CHECK only prints, nms_kernel.cu:12-19) we report zero keeps,
// and detbox_last_cuda_error() tells why.
extern "C" void _nms(int *keep_out, int *num_out, const float *boxes_host, int boxes_num, int boxes_dim,
                     float nms_overlap_thresh, int device_id)
{
    int n = 0;
    const int rc = nms_host_impl(boxes_host, boxes_num, boxes_dim, true, (double)nms_overlap_thresh,
                                 DETBOX_NMS_GT_F32, keep_out, &n, device_id);
    if (num_out) *num_out = (rc == DETBOX_OK) ? n : 0;
}

extern "C" int detbox_nms_host(const float *dets, int n, double thresh, int cmp_mode,
                               int *keep_out, int *num_out, int device_id)
{
    return nms_host_impl(dets, n, 5, false, thresh, cmp_mode, keep_out, num_out, device_id);
}
