// dets.cu — the device-side epilogue of stage one.
// Replaces: the NumPy `to_Dets*` conversions of libs/nets/model.py:357-469 (D2H of boxes / probs / ids, a Python
//           loop over images, np.where + np.concatenate per image) and the score thresholding of
//           _decoding_and_thresholding_stage1 (model.py:323-340: nonzero() + a host-side fallback).
// Both are ordered expansions of a few thousand rows, so each is one single-block kernel: every step is a block
// scan, nothing is read back, and the NMS in between keeps its own sync-free pipeline.
#include "common.cuh"

using namespace detbox;

namespace {

constexpr int DT_THREADS = 1024;
constexpr int DT_MAX_IMAGES = 2048;

__device__ __forceinline__ int block_excl_scan_i(int v, int *warp_sum, int &total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    __syncthreads();
    if (lane == 31) warp_sum[w] = incl;
    __syncthreads();
    if (w == 0) {
        const int s = warp_sum[lane]; int is = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, is, o); if (lane >= o) is += t; }
        warp_sum[lane] = is - s;
        if (lane == 31) warp_sum[32] = is;
    }
    __syncthreads();
    total = warp_sum[32];
    return warp_sum[w] + incl - v;
}

// ---- model.py:326-333: objectness >= threshold, or the single best row when nothing passes ---------------
// Rows failing the test are not removed (that would need their count on the host): they are moved to a
// dummy segment `n_images`, so the segmented NMS keeps them apart and the epilogue drops them.
__global__ void __launch_bounds__(DT_THREADS)
threshold_segments_kernel(const float *__restrict__ scores, const int64_t *__restrict__ img_ids, long n, int n_images,
                          float thr, int32_t *__restrict__ seg32, int64_t *__restrict__ seg64)
{
    __shared__ unsigned long long best;      // (ordered score << 32) | ~index : max = best score, lowest index
    __shared__ int any_pass;
    const int tid = threadIdx.x;
    if (tid == 0) { best = 0ull; any_pass = 0; }
    __syncthreads();
    unsigned long long mine = 0ull; int pass = 0;
    for (long i = tid; i < n; i += DT_THREADS) {
        const float s = scores[i];
        pass |= s >= thr;
        const unsigned long long c = ((unsigned long long)float_to_ordered(s) << 32) | (unsigned)(~(unsigned)i);
        mine = c > mine ? c : mine;
    }
    if (__any_sync(0xffffffffu, pass) && (tid & 31) == 0) any_pass = 1;
    if (n > 0) atomicMax(&best, mine);
    __syncthreads();
    const bool none = any_pass == 0;
    const long best_i = (long)(unsigned)(~(unsigned)best);
    for (long i = tid; i < n; i += DT_THREADS) {
        const bool valid = none ? (i == best_i) : (scores[i] >= thr);
        const int64_t im = img_ids[i];
        const int64_t s = (valid && im >= 0 && im < n_images) ? im : (int64_t)n_images;
        seg32[i] = (int32_t)s; seg64[i] = s;
    }
}

// ---- to_Dets / to_Dets_sigmoid / to_Dets2 / to_Dets2_sigmoid ------------------------------------------------
// Row r of the input is row keep[r] (or r) of boxes / probs; rows must be image-major.  first = index of the
// first foreground column (1 softmax, 0 sigmoid), class id = column - first + 1.
struct DetsArgs {
    const float4 *boxes; const float *probs; int C; const int64_t *img_ids; const int64_t *keep;
    const int32_t *m_dev; long m_max; int n_images; int first; int multi; float thr;
    double *dets; long cap; int32_t *offsets;
};

__global__ void __launch_bounds__(DT_THREADS)
to_dets_kernel(const DetsArgs a)
{
    __shared__ int tot[DT_MAX_IMAGES + 1];     // (row, class) pairs above the threshold, then rows emitted, per image
    __shared__ int fin[DT_MAX_IMAGES + 1];
    __shared__ int warp_sum[33];
    const int tid = threadIdx.x;
    long m = a.m_dev ? (long)*a.m_dev : a.m_max;
    if (m > a.m_max) m = a.m_max;
    const int nfg = a.C - a.first;              // foreground columns
    const bool single = nfg <= 1;               // two-class softmax / one-logit sigmoid: one detection per row, no threshold
    for (int i = tid; i <= a.n_images; i += DT_THREADS) { tot[i] = 0; fin[i] = 0; }
    __syncthreads();
    auto row_of = [&](long r) { return a.keep ? (long)a.keep[r] : r; };
    auto count_above = [&](const float *p) { int c = 0; for (int k = a.first; k < a.C; ++k) c += p[k] > a.thr; return c; };
    if (a.multi && !single) {
        for (long r = tid; r < m; r += DT_THREADS) {
            const long src = row_of(r);
            const int64_t im = a.img_ids[src];
            if (im < 0 || im >= a.n_images) continue;
            const int c = count_above(a.probs + (size_t)src * a.C);
            if (c) atomicAdd(&tot[(int)im], c);
        }
        __syncthreads();
    }
    int running = 0;
    for (long r0 = 0; r0 < m; r0 += DT_THREADS) {
        const long r = r0 + tid;
        int c = 0; long src = 0; int im = -1; bool fallback = false;
        if (r < m) {
            src = row_of(r);
            const int64_t i64 = a.img_ids[src];
            if (i64 >= 0 && i64 < a.n_images) {
                im = (int)i64;
                if (single || !a.multi) c = 1;
                else if (tot[im] == 0) { c = 1; fallback = true; }               // model.py:428-434: argmax for every row
                else c = count_above(a.probs + (size_t)src * a.C);
            }
        }
        int total;
        long pos = running + block_excl_scan_i(c, warp_sum, total);
        running += total;
        if (c) {
            atomicAdd(&fin[im], c);
            const float4 b = a.boxes[src];
            const float *p = a.probs + (size_t)src * a.C;
            auto put = [&](float prob, int cls) {
                if (pos < a.cap) {
                    double *d = a.dets + (size_t)pos * 6;
                    d[0] = b.x; d[1] = b.y; d[2] = b.z; d[3] = b.w; d[4] = prob; d[5] = cls;
                }
                ++pos;
            };
            if (single) {
                put(p[a.C - 1], 1);                                               // model.py:365-367, 387-389
            } else if (!a.multi || fallback) {
                int arg = a.first; float best = p[a.first];                       // first maximum (np.argmax)
                for (int k = a.first + 1; k < a.C; ++k) if (p[k] > best) { best = p[k]; arg = k; }
                put(best, arg - a.first + 1);
            } else {
                for (int k = a.first; k < a.C; ++k) if (p[k] > a.thr) put(p[k], k - a.first + 1);   // np.where order: row, then class
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int i = 0; i < a.n_images; ++i) { a.offsets[i] = run; run += fin[i]; }
        a.offsets[a.n_images] = run;             // true total: > cap means the output was truncated
    }
}


// ---- single_shot.py:401-415 dets_cluster: score-weighted box voting over nms_multianchor's clusters -----------
// One thread per kept box.  clusters[] is read by every thread of a warp at the same index (broadcast), members are
// met in ascending index order — np.where's order — so the float32 sums below reproduce NumPy's: the score sum is
// np.sum's pairwise scheme (8 running sums over blocks of 8, halves above 128 elements), the row sum over axis 0 is
// sequential.  Three passes over clusters[] (count, score sum, vote); n and n_keep are a few thousand here.
struct ClusterCursor {
    const int64_t *cl; const float *dets; long i; int64_t mine;
    __device__ __forceinline__ long next() { while (cl[i] != mine) ++i; return i++; }
    __device__ __forceinline__ float next_score() { return dets[next() * 5 + 4]; }
};

__device__ float pairwise_sum(ClusterCursor &c, int n)       // numpy/_core/src/umath/loops_utils.h.src pairwise sum
{
    if (n < 8) {
        float r = 0.0f;
        for (int i = 0; i < n; ++i) r += c.next_score();
        return r;
    }
    if (n <= 128) {
        float r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = c.next_score();
        int i = 8;
        for (; i < n - (n % 8); i += 8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) r[j] += c.next_score();
        }
        float res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += c.next_score();
        return res;
    }
    int n2 = n / 2; n2 -= n2 % 8;
    const float a = pairwise_sum(c, n2);
    return a + pairwise_sum(c, n - n2);
}

__global__ void __launch_bounds__(128)
dets_cluster_kernel(const float *__restrict__ dets, long n, const int64_t *__restrict__ keep, const int32_t *__restrict__ n_keep_dev,
                    long n_keep_max, const int64_t *__restrict__ clusters, float *__restrict__ out)
{
    const long k = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long nk = n_keep_dev ? min((long)*n_keep_dev, n_keep_max) : n_keep_max;
    if (k >= nk) return;
    const int64_t mine = keep[k];
    int m = 0; long first = -1;
    for (long i = 0; i < n; ++i) if (clusters[i] == mine) { if (m == 0) first = i; ++m; }
    float *o = out + k * 5;
    if (m <= 1) {                                  // a cluster of one keeps its row (m == 0: the keeper's own row)
        const long src = m == 1 ? first : (long)mine;
        const bool ok = src >= 0 && src < n;
        for (int c = 0; c < 5; ++c) o[c] = ok ? dets[src * 5 + c] : 0.0f;
        return;
    }
    ClusterCursor cur{clusters, dets, first, mine};
    const float S = pairwise_sum(cur, m);
    cur.i = first;
    float acc[4], best = 0.0f;
    for (int j = 0; j < m; ++j) {
        const long i = cur.next();
        const float s = dets[i * 5 + 4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float v = dets[i * 5 + c] * s / S;
            acc[c] = j == 0 ? v : acc[c] + v;
        }
        best = (j == 0 || s > best) ? s : best;    // np.max
    }
    o[0] = acc[0]; o[1] = acc[1]; o[2] = acc[2]; o[3] = acc[3]; o[4] = best;
}

}  // namespace

extern "C" int detbox_threshold_segments(const float *scores, const int64_t *img_ids, long n, int n_images,
                                         float score_threshold, int32_t *seg32, int64_t *seg64, void *stream)
{
    if (n < 0 || n_images < 1) return DETBOX_ERR_INVALID;
    if (n == 0) return DETBOX_OK;
    if (!scores || !img_ids || !seg32 || !seg64) return DETBOX_ERR_INVALID;
    if (n >= (1L << 31)) return DETBOX_ERR_UNSUPPORTED;
    threshold_segments_kernel<<<1, DT_THREADS, 0, as_stream(stream)>>>(scores, img_ids, n, n_images, score_threshold, seg32, seg64);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

extern "C" int detbox_to_dets(const float *boxes, const float *probs, int C, const int64_t *img_ids,
                              const int64_t *keep, const int32_t *m_dev, long m_max, int n_images,
                              int activation, int multi_label, float score_threshold,
                              double *dets, long cap, int32_t *det_offsets, void *stream)
{
    if (C < 1 || m_max < 0 || n_images < 1 || cap < 0 || !det_offsets) return DETBOX_ERR_INVALID;
    if (activation != DETBOX_ACT_SOFTMAX && activation != DETBOX_ACT_SIGMOID) return DETBOX_ERR_INVALID;
    if (n_images > DT_MAX_IMAGES) return DETBOX_ERR_UNSUPPORTED;
    if (m_max > 0 && (!boxes || !probs || !img_ids || (cap > 0 && !dets))) return DETBOX_ERR_INVALID;
    if ((uintptr_t)boxes & 15) return DETBOX_ERR_INVALID;
    DetsArgs a;
    a.boxes = (const float4 *)boxes; a.probs = probs; a.C = C; a.img_ids = img_ids; a.keep = keep;
    a.m_dev = m_dev; a.m_max = m_max; a.n_images = n_images;
    a.first = activation == DETBOX_ACT_SOFTMAX ? 1 : 0;
    a.multi = multi_label != 0; a.thr = score_threshold;
    a.dets = dets; a.cap = cap; a.offsets = det_offsets;
    to_dets_kernel<<<1, DT_THREADS, 0, as_stream(stream)>>>(a);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

extern "C" int detbox_dets_cluster(const float *dets, long n, const int64_t *keep, const int32_t *n_keep_dev, long n_keep_max,
                                   const int64_t *clusters, float *out, void *stream)
{
    if (n < 0 || n_keep_max < 0) return DETBOX_ERR_INVALID;
    if (n == 0 || n_keep_max == 0) return DETBOX_OK;
    if (!dets || !keep || !clusters || !out) return DETBOX_ERR_INVALID;
    const long blocks = (n_keep_max + 127) / 128;
    if (blocks >= (1L << 31)) return DETBOX_ERR_UNSUPPORTED;
    dets_cluster_kernel<<<(unsigned)blocks, 128, 0, as_stream(stream)>>>(dets, n, keep, n_keep_dev, n_keep_max, clusters, out);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

// ---- fixed-size detection buffers for the optional all-gather epilogue -------------------------------------
// keep[] holds the kept ORIGINAL row indices segment-major and score-descending inside a segment (detbox_nms), so the
// first max_dets entries of a segment are its best detections.  One block per image: start = sum of the counts of the
// earlier segments, then rows -> (x1, y1, x2, y2, score, class) with zero padding.  No host round trip.
namespace {
__global__ void __launch_bounds__(128)
pack_dets_kernel(const float4 *__restrict__ boxes, const float *__restrict__ scores, const float *__restrict__ classes,
                 const int64_t *__restrict__ keep, const int32_t *__restrict__ n_keep, int n_seg, int max_dets,
                 float *__restrict__ out, int32_t *__restrict__ counts)
{
    __shared__ int s_part[4];
    const int b = blockIdx.x, tid = threadIdx.x;
    int part = 0;
    for (int i = tid; i < b; i += 128) part += n_keep[i];
    part = __reduce_add_sync(0xffffffffu, part);
    if ((tid & 31) == 0) s_part[tid >> 5] = part;
    __syncthreads();
    const int start = s_part[0] + s_part[1] + s_part[2] + s_part[3];
    const int m = min(max(n_keep[b], 0), max_dets);
    if (tid == 0) counts[b] = m;
    float *o = out + (size_t)b * max_dets * 6;
    for (int j = tid; j < max_dets; j += 128) {
        float v[6] = {0, 0, 0, 0, 0, 0};
        if (j < m) {
            const int64_t r = keep[start + j];
            const float4 bx = boxes[r];
            v[0] = bx.x; v[1] = bx.y; v[2] = bx.z; v[3] = bx.w; v[4] = scores[r]; v[5] = classes ? classes[r] : 1.0f;
        }
#pragma unroll
        for (int k = 0; k < 6; ++k) o[j * 6 + k] = v[k];
    }
}
}  // namespace

extern "C" int detbox_pack_detections(const float *boxes, const float *scores, const float *classes,
                                      const int64_t *keep, const int32_t *n_keep, int n_seg, int max_dets,
                                      float *out, int32_t *counts, void *stream)
{
    if (n_seg < 0 || max_dets < 0) return DETBOX_ERR_INVALID;
    if (n_seg == 0) return DETBOX_OK;
    if (!boxes || !scores || !keep || !n_keep || !counts || (max_dets > 0 && !out)) return DETBOX_ERR_INVALID;
    if (((uintptr_t)boxes & 15) != 0) return DETBOX_ERR_INVALID;
    pack_dets_kernel<<<(unsigned)n_seg, 128, 0, as_stream(stream)>>>((const float4 *)boxes, scores, classes, keep, n_keep,
                                                                    n_seg, max_dets, out, counts);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}
