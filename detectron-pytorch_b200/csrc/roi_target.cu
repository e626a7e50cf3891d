// roi_target.cu — RoI -> ground-truth matching, RoI targets and the deterministic half of RoI sampling
// (the step right after NMS in MaskRCNN training).
// Replaces: libs/layers/box.py:12-69 (matching_box: D2H of the RoIs, Cython IoU per image, H2D),
//           libs/layers/box.py:175-292 (sample_rois: same round trip + filter / ignore / append GT),
//           libs/layers/roi_target/src/cuda/roi_target_kernel.cu:49-90 (ROITargetForward).
// One thread per RoI; the RoIs of a warp belong to one image almost always (NMS output is image-major),
// so the GT reads are warp-uniform L1 hits.  R is thousands, G tens: latency-bound by design, the gain is
// that nothing leaves the device.  --fmad=false as everywhere.
#include "common.cuh"

using namespace detbox;

namespace {

// ---- matching_box / sample_rois labelling --------------------------------------------------------
// float64 IoU with the +1 convention in bbox.pyx:15-55's operation order; argmax = first maximum
// (np.argmax).  cfg.min_size > 0: filter_boxes (box.py:310-315, float32 arithmetic) -> label -2.
// cfg.ignore_thresh >= 0: rows covered by a GT of class <= 0 (bbox.pyx:96-135 against box.py:213-221's
// mask) -> label -1.  Rows whose image index is outside [0, B) -> label -2, match -1.
__global__ void __launch_bounds__(256)
match_boxes_kernel(const float *__restrict__ boxes, int stride, const int64_t *__restrict__ image_inds, long R,
                   const float *__restrict__ gt, const int32_t *__restrict__ gt_off, int B, detbox_match_cfg cfg,
                   int64_t *__restrict__ labels, int32_t *__restrict__ match, float *__restrict__ match_boxes)
{
    const long r = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const float *p = boxes + (size_t)r * stride;
    const float fx1 = p[0], fy1 = p[1], fx2 = p[2], fy2 = p[3];
    const int64_t img = image_inds[r];
    int64_t label = 0; int m = -1;
    float4 mb = make_float4(0.f, 0.f, 0.f, 0.f);
    bool dropped = img < 0 || img >= B;
    if (!dropped && cfg.min_size > 0.0f) {
        const float w = fx2 - fx1 + 1.0f, h = fy2 - fy1 + 1.0f;
        dropped = !(w >= cfg.min_size && h >= cfg.min_size);
    }
    if (dropped) {
        label = -2;
    } else {
        const int g0 = gt_off[img], g1 = gt_off[img + 1];
        if (g1 > g0) {
            const double x1 = fx1, y1 = fy1, x2 = fx2, y2 = fy2;
            const double area = (x2 - x1 + 1) * (y2 - y1 + 1);
            double best = -1.0; int arg = 0; bool ignored = false;
            for (int j = g0; j < g1; ++j) {
                const float *q = gt + (size_t)j * 5;
                const double qx1 = q[0], qy1 = q[1], qx2 = q[2], qy2 = q[3];
                double ov = 0.0;
                const double iw = fmin(x2, qx2) - fmax(x1, qx1) + 1;
                if (iw > 0) {
                    const double ih = fmin(y2, qy2) - fmax(y1, qy1) + 1;
                    if (ih > 0) {
                        const double qarea = (qx2 - qx1 + 1) * (qy2 - qy1 + 1);
                        ov = iw * ih / (area + qarea - iw * ih);
                        if (cfg.ignore_thresh >= 0.0 && q[4] <= 0.0f && iw * ih / area > cfg.ignore_thresh) ignored = true;
                    }
                }
                if (ov > best) { best = ov; arg = j; }
            }
            const float *q = gt + (size_t)arg * 5;
            mb = make_float4(q[0], q[1], q[2], q[3]);
            m = best < cfg.bg_thresh ? -1 : arg - g0;
            label = best < cfg.fg_thresh ? 0 : (int64_t)q[4];
            if (ignored) label = -1;
        }
    }
    if (labels) labels[r] = label;
    if (match) match[r] = m;
    if (match_boxes) reinterpret_cast<float4 *>(match_boxes)[r] = mb;
}

// ---- ROITargetForward ------------------------------------------------------------------------------
// The reference kernel's float32 arithmetic with its double-literal promotions (`+ 1.0`, `* 0.5`,
// `/ 0.1`, `/ 0.2`: roi_target_kernel.cu:31-48), `<=` tie rule (the LAST maximum wins, and a zero
// overlap still matches: best_iou starts at 0), inter/area(roi) for GT of class <= 0 (:25-27).
__device__ __forceinline__ float roi_iou(const float *a, const float *b, int cls) {
    const float left = fmaxf(a[0], b[0]), right = fminf(a[2], b[2]);
    const float top = fmaxf(a[1], b[1]), bottom = fminf(a[3], b[3]);
    const float width = fmaxf(right - left + 1, 0.f), height = fmaxf(bottom - top + 1, 0.f);
    const float inter = width * height;
    const float sa = (a[2] - a[0] + 1) * (a[3] - a[1] + 1);
    const float sb = (b[2] - b[0] + 1) * (b[3] - b[1] + 1);
    if (cls <= 0) return inter / sa;
    return inter / (sa + sb - inter);
}

__global__ void __launch_bounds__(256)
roi_target_kernel(const float *__restrict__ rois, int stride, const int64_t *__restrict__ roi_img, long R,
                  const float *__restrict__ gt, const int64_t *__restrict__ gt_img, int G,
                  float fg_overlap, int64_t *__restrict__ labels, float *__restrict__ deltas, float *__restrict__ bbw)
{
    const long r = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const float *a = rois + (size_t)r * stride;
    const int rid = (int)roi_img[r];
    float best = 0.f; int best_ind = -1, match_cls = -1;
    for (int j = 0; j < G; ++j) {
        if ((int)gt_img[j] != rid) continue;
        const float *b = gt + (size_t)j * 5;
        const int cls = (int)b[4];
        const float iou = roi_iou(a, b, cls);
        if (best <= iou) { best = iou; best_ind = j; match_cls = cls; }
    }
    labels[r] = best >= fg_overlap ? (match_cls > 0 ? match_cls : -1) : 0;
    if (best >= fminf(0.5f, fg_overlap) && match_cls > 0 && best_ind >= 0) {
        const float *b = gt + (size_t)best_ind * 5;
        const float bw = (float)((double)(b[2] - b[0]) + 1.0), bh = (float)((double)(b[3] - b[1]) + 1.0);
        const float bx = (float)((double)b[0] + (double)bw * 0.5), by = (float)((double)b[1] + (double)bh * 0.5);
        const float aw = (float)((double)(a[2] - a[0]) + 1.0), ah = (float)((double)(a[3] - a[1]) + 1.0);
        const float ax = (float)((double)a[0] + (double)aw * 0.5), ay = (float)((double)a[1] + (double)ah * 0.5);
        float4 d;
        d.x = (float)((double)((bx - ax) / aw) / 0.1);
        d.y = (float)((double)((by - ay) / ah) / 0.1);
        d.z = (float)((double)logf(bw / aw) / 0.2);
        d.w = (float)((double)logf(bh / ah) / 0.2);
        reinterpret_cast<float4 *>(deltas)[r] = d;
        reinterpret_cast<float4 *>(bbw)[r] = make_float4(1.f, 1.f, 1.f, 1.f);
    }
}

// ---- sample_rois: ordered compaction ---------------------------------------------------------------
// box.py:236-250: per image, the kept proposals (label >= 0) followed by that image's appended GT rows;
// images in order.  Then the positions of foreground (label > 0) and background (label == 0) rows in
// ascending order (np.where, box.py:253-254).  One block: R + E is a few thousand rows and every step
// is an ordered scan.  Proposal rows must be image-major (the mirror sorts them if they are not).
constexpr int SC_THREADS = 1024;
constexpr int SC_MAX_IMAGES = 2048;

__device__ __forceinline__ int block_excl_scan(int v, int *warp_sum, int &total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    __syncthreads();                         // warp_sum re-use across calls
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

__global__ void __launch_bounds__(SC_THREADS)
sample_compact_kernel(const float *__restrict__ boxes, int stride, const int64_t *__restrict__ image_inds,
                      const int64_t *__restrict__ labels, int R,
                      const float *__restrict__ extra_boxes, const int64_t *__restrict__ extra_labels,
                      const int32_t *__restrict__ extra_off, int B,
                      float *out_boxes, int64_t *out_labels, int64_t *out_img,
                      int32_t *__restrict__ fg_pos, int32_t *__restrict__ bg_pos, int32_t *__restrict__ counts)
{
    __shared__ int cnt[SC_MAX_IMAGES + 1];       // kept proposals per image -> exclusive prefix
    __shared__ int base[SC_MAX_IMAGES + 1];      // first output row of an image
    __shared__ int warp_sum[33];
    const int tid = threadIdx.x;
    for (int i = tid; i <= B; i += SC_THREADS) cnt[i] = 0;
    __syncthreads();
    for (int r = tid; r < R; r += SC_THREADS) {
        const int64_t img = image_inds[r];
        if (labels[r] >= 0 && img >= 0 && img < B) atomicAdd(&cnt[(int)img], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int kp = 0, ob = 0;
        for (int i = 0; i < B; ++i) {
            const int c = cnt[i], e = extra_off ? extra_off[i + 1] - extra_off[i] : 0;
            cnt[i] = kp; base[i] = ob;
            kp += c; ob += c + e;
        }
        cnt[B] = kp; base[B] = ob;
    }
    __syncthreads();
    int running = 0;
    for (int r0 = 0; r0 < R; r0 += SC_THREADS) {
        const int r = r0 + tid;
        int64_t img = -1, lab = -2;
        if (r < R) { img = image_inds[r]; lab = labels[r]; }
        const bool keep = lab >= 0 && img >= 0 && img < B;
        int total;
        const int rank = running + block_excl_scan(keep ? 1 : 0, warp_sum, total);
        if (keep) {
            const int i = (int)img, pos = base[i] + (rank - cnt[i]);
            const float *p = boxes + (size_t)r * stride;
            reinterpret_cast<float4 *>(out_boxes)[pos] = make_float4(p[0], p[1], p[2], p[3]);
            out_labels[pos] = lab; out_img[pos] = img;
        }
        running += total;
    }
    if (extra_off) {
        const int E = extra_off[B];
        for (int e = tid; e < E; e += SC_THREADS) {
            int lo = 0, hi = B - 1;                          // last image with extra_off[i] <= e
            while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (extra_off[mid] <= e) lo = mid; else hi = mid - 1; }
            const int pos = base[lo] + (cnt[lo + 1] - cnt[lo]) + (e - extra_off[lo]);
            reinterpret_cast<float4 *>(out_boxes)[pos] = reinterpret_cast<const float4 *>(extra_boxes)[e];
            out_labels[pos] = extra_labels[e]; out_img[pos] = lo;
        }
    }
    __threadfence_block();
    __syncthreads();
    const int total_rows = base[B];
    int nfg = 0, nbg = 0;
    for (int r0 = 0; r0 < total_rows; r0 += SC_THREADS) {
        const int r = r0 + tid;
        const int64_t lab = r < total_rows ? out_labels[r] : -1;
        int tf, tb;
        const int rf = block_excl_scan(lab > 0 ? 1 : 0, warp_sum, tf);
        const int rb = block_excl_scan(lab == 0 ? 1 : 0, warp_sum, tb);
        if (lab > 0) fg_pos[nfg + rf] = r;
        if (lab == 0) bg_pos[nbg + rb] = r;
        nfg += tf; nbg += tb;
    }
    if (tid == 0) { counts[0] = total_rows; counts[1] = nfg; counts[2] = nbg; }
}

// keep_inds = append(fg_pos[fg_choice], bg_pos[bg_choice]) (box.py:271-276), then the three gathers
__global__ void __launch_bounds__(256)
sample_gather_kernel(const float *__restrict__ boxes, const int64_t *__restrict__ labels, const int64_t *__restrict__ img,
                     const int32_t *__restrict__ fg_pos, const int32_t *__restrict__ bg_pos,
                     const int64_t *__restrict__ fg_choice, int n_fg, const int64_t *__restrict__ bg_choice, int n_bg,
                     float *__restrict__ out_boxes, int64_t *__restrict__ out_labels, int64_t *__restrict__ out_img)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_fg + n_bg) return;
    const int src = k < n_fg ? fg_pos[fg_choice[k]] : bg_pos[bg_choice[k - n_fg]];
    reinterpret_cast<float4 *>(out_boxes)[k] = reinterpret_cast<const float4 *>(boxes)[src];
    out_labels[k] = labels[src]; out_img[k] = img[src];
}

}  // namespace

extern "C" int detbox_match_boxes(const float *boxes, int box_stride, const int64_t *image_inds, long R,
                                  const float *gt, const int32_t *gt_offsets, int B, const detbox_match_cfg *cfg,
                                  int64_t *labels, int32_t *match_inds, float *match_boxes, void *stream)
{
    if (R < 0 || B < 0 || box_stride < 4 || !cfg) return DETBOX_ERR_INVALID;
    if (R == 0) return DETBOX_OK;
    if (!boxes || !image_inds || !gt_offsets) return DETBOX_ERR_INVALID;
    if (match_boxes && ((uintptr_t)match_boxes & 15)) return DETBOX_ERR_INVALID;
    match_boxes_kernel<<<(unsigned)((R + 255) / 256), 256, 0, as_stream(stream)>>>(
        boxes, box_stride, image_inds, R, gt, gt_offsets, B, *cfg, labels, match_inds, match_boxes);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

extern "C" int detbox_roi_target(const float *rois, int roi_stride, const int64_t *roi_batch_inds, long R,
                                 const float *gt, const int64_t *gt_batch_inds, long G,
                                 float bg_overlap, float fg_overlap,
                                 int64_t *labels, float *deltas, float *bbwght, void *stream)
{
    (void)bg_overlap;                            // accepted and unused, like the reference (roi_target_kernel.cu:52)
    if (R < 0 || G < 0 || G > 0x7fffffffL || roi_stride < 4) return DETBOX_ERR_INVALID;
    if (R == 0) return DETBOX_OK;
    if (!rois || !roi_batch_inds || !labels || !deltas || !bbwght || (G > 0 && (!gt || !gt_batch_inds))) return DETBOX_ERR_INVALID;
    if (((uintptr_t)deltas & 15) || ((uintptr_t)bbwght & 15)) return DETBOX_ERR_INVALID;
    cudaStream_t st = as_stream(stream);
    // the Python side of the reference zero-fills deltas / weights before the launch (functions/roi_target.py:20-22)
    DETBOX_CUDA_TRY(cudaMemsetAsync(deltas, 0, (size_t)R * 16, st));
    DETBOX_CUDA_TRY(cudaMemsetAsync(bbwght, 0, (size_t)R * 16, st));
    roi_target_kernel<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(rois, roi_stride, roi_batch_inds, R, gt, gt_batch_inds,
                                                                   (int)G, fg_overlap, labels, deltas, bbwght);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

extern "C" int detbox_sample_compact(const float *boxes, int box_stride, const int64_t *image_inds, const int64_t *labels, long R,
                                     const float *extra_boxes, const int64_t *extra_labels, const int32_t *extra_offsets, int B,
                                     float *out_boxes, int64_t *out_labels, int64_t *out_image_inds,
                                     int32_t *fg_pos, int32_t *bg_pos, int32_t *counts, void *stream)
{
    if (R < 0 || R > 0x3fffffffL || B < 1 || box_stride < 4) return DETBOX_ERR_INVALID;
    if (B > SC_MAX_IMAGES) return DETBOX_ERR_UNSUPPORTED;
    if (!out_boxes || !out_labels || !out_image_inds || !fg_pos || !bg_pos || !counts) return DETBOX_ERR_INVALID;
    if (R > 0 && (!boxes || !image_inds || !labels)) return DETBOX_ERR_INVALID;
    if (extra_offsets && (!extra_boxes || !extra_labels)) return DETBOX_ERR_INVALID;
    if (((uintptr_t)out_boxes & 15) || (extra_boxes && ((uintptr_t)extra_boxes & 15))) return DETBOX_ERR_INVALID;
    sample_compact_kernel<<<1, SC_THREADS, 0, as_stream(stream)>>>(boxes, box_stride, image_inds, labels, (int)R,
                                                                   extra_boxes, extra_labels, extra_offsets, B,
                                                                   out_boxes, out_labels, out_image_inds, fg_pos, bg_pos, counts);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

extern "C" int detbox_sample_gather(const float *boxes, const int64_t *labels, const int64_t *image_inds,
                                    const int32_t *fg_pos, const int32_t *bg_pos,
                                    const int64_t *fg_choice, int n_fg, const int64_t *bg_choice, int n_bg,
                                    float *out_boxes, int64_t *out_labels, int64_t *out_image_inds, void *stream)
{
    if (n_fg < 0 || n_bg < 0) return DETBOX_ERR_INVALID;
    const int n = n_fg + n_bg;
    if (n == 0) return DETBOX_OK;
    if (!boxes || !labels || !image_inds || !out_boxes || !out_labels || !out_image_inds) return DETBOX_ERR_INVALID;
    if ((n_fg > 0 && (!fg_pos || !fg_choice)) || (n_bg > 0 && (!bg_pos || !bg_choice))) return DETBOX_ERR_INVALID;
    if (((uintptr_t)boxes & 15) || ((uintptr_t)out_boxes & 15)) return DETBOX_ERR_INVALID;
    sample_gather_kernel<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(
        boxes, labels, image_inds, fg_pos, bg_pos, fg_choice, n_fg, bg_choice, n_bg, out_boxes, out_labels, out_image_inds);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}
