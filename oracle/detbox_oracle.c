/* detbox_oracle.c — CPU restatement of the reference's box-and-RoI hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the *checker* for the CUDA path in
 * detectron-pytorch_b200/csrc; it is never linked, imported or called by the product.  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs load it.
 *
 * Every function names the reference file:line (relative to the reference root) whose
 * arithmetic — operation order, dtype of every intermediate, comparator — it restates.
 * Parity of this file against the reference itself is pinned by tests/test_oracle_*.py:
 *   (a) against tests/golden/*.npz, produced by oracle/make_golden.py from the compiled,
 *       unmodified reference (oracle/_ref), and
 *   (b) when oracle/_ref is present, against that reference on fresh random inputs.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp -fPIC -shared (oracle/Makefile).
 * -ffp-contract=off matters: the reference is NumPy / Cython / gcc -O2 on x86-64, none of
 * which fuse multiply-add.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

static inline double dmin(double a, double b) { return a < b ? a : b; }
static inline double dmax(double a, double b) { return a > b ? a : b; }

/* ------------------------------------------------------------------------------------------
 * a1  base anchors — libs/boxes/anchor.py:125-189 (generate_anchors, _whctrs, _mkanchors,
 *     _ratio_enum, _scale_enum) and :101-103 (anchors(): scales cast to float32 first).
 *     out[(r*n_scales + s)*4 + c], ratio-major / scale-minor, float64.
 *     np.round is round-half-to-even == rint() in the default rounding mode.
 *     Promotion follows NumPy >= 2 (the NumPy the reference runs under in this image):
 *     float64 scalar * float32 array -> float64.
 * ---------------------------------------------------------------------------------------- */
ORC_API void orc_base_anchors(const double *scales, int n_scales, const double *ratios,
                              int n_ratios, long base_size, double *out)
{
    /* base_anchor = [-b//2, -b//2, b//2, b//2]  (Python floor division), anchor.py:132 */
    double lo = floor(-(double)base_size / 2.0), hi = floor((double)base_size / 2.0);
    double w = hi - lo + 1, h = hi - lo + 1;                  /* _whctrs :143-146 */
    double xc = lo + 0.5 * (w - 1), yc = lo + 0.5 * (h - 1);
    double size = w * h;                                       /* _ratio_enum :171-176 */
    for (int r = 0; r < n_ratios; ++r) {
        double ws = rint(sqrt(size / ratios[r]));
        double hs = rint(ws * ratios[r]);
        /* ratio anchor, _mkanchors :158-162 */
        double rx1 = xc - 0.5 * (ws - 1), ry1 = yc - 0.5 * (hs - 1);
        double rx2 = xc + 0.5 * (ws - 1), ry2 = yc + 0.5 * (hs - 1);
        /* _scale_enum :185-188 on that anchor */
        double rw = rx2 - rx1 + 1, rh = ry2 - ry1 + 1;
        double rxc = rx1 + 0.5 * (rw - 1), ryc = ry1 + 0.5 * (rh - 1);
        for (int s = 0; s < n_scales; ++s) {
            double sc = (double)(float)scales[s];              /* np.asarray(scales, np.float32) :103 */
            double sw = rw * sc, sh = rh * sc;
            double *o = out + ((size_t)r * n_scales + s) * 4;
            o[0] = rxc - 0.5 * (sw - 1);
            o[1] = ryc - 0.5 * (sh - 1);
            o[2] = rxc + 0.5 * (sw - 1);
            o[3] = ryc + 0.5 * (sh - 1);
        }
    }
}

/* a2  one pyramid level — libs/boxes/anchor.py:47-70 (anchor_one_plane).
 *     base: [n_base,4] f64; shifts: [n_shift,2] f64 (x,y) in units of stride.
 *     Order: y -> x -> (shift, base).  out float64 [height*width*n_shift*n_base, 4]. */
ORC_API void orc_anchor_plane(const double *base, int n_base, const double *shifts, int n_shift,
                              int height, int width, long stride, double *out)
{
    size_t A = (size_t)n_base * n_shift;
    double *anc = (double *)malloc(A * 4 * sizeof(double));
    for (int s = 0; s < n_shift; ++s)
        for (int b = 0; b < n_base; ++b)
            for (int c = 0; c < 4; ++c)                          /* hstack((sh,sh))*stride :54-55 */
                anc[((size_t)s * n_base + b) * 4 + c] = base[b * 4 + c] + shifts[s * 2 + (c & 1)] * (double)stride;
    for (int y = 0; y < height; ++y)
        for (int x = 0; x < width; ++x) {
            double loc[4] = {(double)(x * stride), (double)(y * stride),
                             (double)(x * stride), (double)(y * stride)};   /* :59-63 */
            double *o = out + ((size_t)y * width + x) * A * 4;
            for (size_t a = 0; a < A; ++a)
                for (int c = 0; c < 4; ++c) o[a * 4 + c] = anc[a * 4 + c] + loc[c];  /* :67 */
        }
    free(anc);
}

/* ------------------------------------------------------------------------------------------
 * a3  bbox_overlaps — libs/boxes/bbox.pyx:15-55.  float64, +1 pixel convention, k-outer /
 *     n-inner, row-major [N,K] output that must be zero-initialised by the caller.
 * ---------------------------------------------------------------------------------------- */
ORC_API void orc_bbox_overlaps(const double *boxes, long N, const double *query, long K, double *out)
{
    for (long k = 0; k < K; ++k) {
        const double *q = query + k * 4;
        double box_area = (q[2] - q[0] + 1) * (q[3] - q[1] + 1);           /* :34-37 */
        for (long n = 0; n < N; ++n) {
            const double *b = boxes + n * 4;
            double iw = dmin(b[2], q[2]) - dmax(b[0], q[0]) + 1;           /* :39-42 */
            if (iw > 0) {
                double ih = dmin(b[3], q[3]) - dmax(b[1], q[1]) + 1;       /* :44-47 */
                if (ih > 0) {
                    double ua = (b[2] - b[0] + 1) * (b[3] - b[1] + 1) + box_area - iw * ih; /* :49-53 */
                    out[n * K + k] = iw * ih / ua;                         /* :54 */
                }
            }
        }
    }
}

/* a4  bbox_intersections — bbox.pyx:57-94: inter / area(query_k). */
ORC_API void orc_bbox_intersections(const double *boxes, long N, const double *query, long K, double *out)
{
    for (long k = 0; k < K; ++k) {
        const double *q = query + k * 4;
        double box_area = (q[2] - q[0] + 1) * (q[3] - q[1] + 1);
        for (long n = 0; n < N; ++n) {
            const double *b = boxes + n * 4;
            double iw = dmin(b[2], q[2]) - dmax(b[0], q[0]) + 1;
            if (iw > 0) {
                double ih = dmin(b[3], q[3]) - dmax(b[1], q[1]) + 1;
                if (ih > 0) out[n * K + k] = iw * ih / box_area;           /* :93 */
            }
        }
    }
}

/* a4  bbox_exclude_ignored_areas — bbox.pyx:96-135: flag[n]=1 if inter/area(box_n) > thr for any k. */
ORC_API void orc_bbox_exclude_ignored_areas(const double *boxes, long N, const double *ign, long K,
                                            double thr, int64_t *out)
{
    for (long n = 0; n < N; ++n) {
        const double *b = boxes + n * 4;
        double box_area = (b[2] - b[0] + 1.0) * (b[3] - b[1] + 1.0);
        out[n] = 0;
        for (long k = 0; k < K; ++k) {
            const double *q = ign + k * 4;
            double iw = dmin(b[2], q[2]) - dmax(b[0], q[0]) + 1;
            if (iw > 0) {
                double ih = dmin(b[3], q[3]) - dmax(b[1], q[1]) + 1;
                if (ih > 0 && iw * ih / box_area > thr) out[n] = 1;
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * a7  bbox_transform — libs/boxes/bbox_transform.py:11-41 (== torch encoding_box
 *     libs/layers/box.py:112-132), float32 throughout, constants rounded to float32
 *     (NumPy: python-float scalar with a float32 array stays float32).
 * ---------------------------------------------------------------------------------------- */
static inline void enc_fastrcnn_f32(const float *ex, const float *gt, float *t)
{
    float ew = ex[2] - ex[0] + 1.0f, eh = ex[3] - ex[1] + 1.0f;       /* :18-19 */
    float ecx = ex[0] + 0.5f * ew, ecy = ex[1] + 0.5f * eh;           /* :20-21 */
    float gw = gt[2] - gt[0] + 1.0f, gh = gt[3] - gt[1] + 1.0f;       /* :27-28 */
    float gcx = gt[0] + 0.5f * gw, gcy = gt[1] + 0.5f * gh;           /* :29-30 */
    t[0] = (gcx - ecx) / ew / 0.1f;                                   /* :34 */
    t[1] = (gcy - ecy) / eh / 0.1f;
    t[2] = logf(gw / ew) / 0.2f;                                      /* :36 */
    t[3] = logf(gh / eh) / 0.2f;
}

/* mode 0 'fastrcnn' (bbox_transform.py:11-41), 1 'linear' (:93-116, alpha=10),
 * 2 'simple' (libs/layers/box.py:134-143) */
ORC_API void orc_box_encode(const float *anchors, const float *gt, long n, int mode, float *out)
{
    for (long i = 0; i < n; ++i) {
        const float *a = anchors + i * 4, *g = gt + i * 4;
        float *t = out + i * 4;
        if (mode == 0) {
            enc_fastrcnn_f32(a, g, t);
        } else {
            float w = a[2] - a[0] + 1.0f, h = a[3] - a[1] + 1.0f;
            for (int c = 0; c < 4; ++c) {
                float d = g[c] - a[c];
                float s = (c & 1) ? h : w;
                t[c] = mode == 1 ? d / s * 10.0f : d / s;
            }
        }
    }
}

/* a8  decode — bbox_transform.py:43-76 (== torch decoding_box libs/layers/box.py:72-109);
 *     float32; note exp(dw + log(w)) and the -1 on x2/y2.  clip: bbox_transform.py:78-91,
 *     applied when im_h > 0 (x to [0, W-1], y to [0, H-1]).
 *     mode 0 'fastrcnn', 1 'linear' (:118-144), 2 'simple' (box.py:102-107). */
ORC_API void orc_box_decode(const float *anchors, const float *deltas, long n, int mode,
                            int im_h, int im_w, float *out)
{
    for (long i = 0; i < n; ++i) {
        const float *a = anchors + i * 4, *d = deltas + i * 4;
        float *o = out + i * 4;
        float w = a[2] - a[0] + 1.0f, h = a[3] - a[1] + 1.0f;
        if (mode == 0) {
            float cx = a[0] + 0.5f * w, cy = a[1] + 0.5f * h;
            float dx = d[0] * 0.1f, dy = d[1] * 0.1f, dw = d[2] * 0.2f, dh = d[3] * 0.2f;
            float pcx = dx * w + cx, pcy = dy * h + cy;
            float pw = expf(dw + logf(w)), ph = expf(dh + logf(h));
            o[0] = pcx - 0.5f * pw;
            o[1] = pcy - 0.5f * ph;
            o[2] = pcx + 0.5f * pw - 1.0f;
            o[3] = pcy + 0.5f * ph - 1.0f;
        } else if (mode == 1) {
            for (int c = 0; c < 4; ++c) o[c] = d[c] * ((c & 1) ? h : w) / 10.0f + a[c];
        } else {
            for (int c = 0; c < 4; ++c) o[c] = d[c] * ((c & 1) ? h : w) + a[c];
        }
        if (im_h > 0) {
            float mx = (float)(im_w - 1), my = (float)(im_h - 1);
            o[0] = fmaxf(fminf(o[0], mx), 0.0f);
            o[1] = fmaxf(fminf(o[1], my), 0.0f);
            o[2] = fmaxf(fminf(o[2], mx), 0.0f);
            o[3] = fmaxf(fminf(o[3], my), 0.0f);
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * a5  anchor-target assignment — libs/layers/anchor.py:17-145 ('simple' strategy, 'fastrcnn'
 *     or 'linear' encoding), steps numbered as in SURVEY.md §8a row a5.
 *     gt [G,5] f32 (x1,y1,x2,y2,cls), anchors [A,4] f32.
 *     Outputs f32: labels[A], label_weights[A], bbox_targets[A,4], bbox_inside_weights[A,4].
 *     Diagnostics (may be NULL): max_overlaps[A] f64, gt_assignment[A] i32, ignore_sum[A] f64.
 * ---------------------------------------------------------------------------------------- */
ORC_API void orc_anchor_encode(const float *gt, int G, const float *anchors, long A,
                               double bg_thr, double fg_thr, int clobber_positives,
                               double ignore_frac, const float *bbweights, int box_encoding,
                               float *labels, float *label_weights, float *bbox_targets,
                               float *bbox_inside_weights,
                               double *max_overlaps_out, int32_t *assignment_out, double *ignore_sum_out)
{
    int32_t *flags = (int32_t *)calloc((size_t)A, sizeof(int32_t));
    memset(label_weights, 0, (size_t)A * sizeof(float));
    memset(bbox_targets, 0, (size_t)A * 4 * sizeof(float));
    memset(bbox_inside_weights, 0, (size_t)A * 4 * sizeof(float));
    if (G > 0) {
        double *ad = (double *)malloc((size_t)A * 4 * sizeof(double));
        double *gd = (double *)malloc((size_t)G * 4 * sizeof(double));
        double *ov = (double *)calloc((size_t)A * G, sizeof(double));
        int32_t *assign = (int32_t *)malloc((size_t)A * sizeof(int32_t));
        double *maxov = (double *)malloc((size_t)A * sizeof(double));
        double *gtmax = (double *)malloc((size_t)G * sizeof(double));
        for (long i = 0; i < A * 4; ++i) ad[i] = (double)anchors[i];          /* :39 */
        for (int g = 0; g < G; ++g)
            for (int c = 0; c < 4; ++c) gd[g * 4 + c] = (double)gt[g * 5 + c]; /* :40 */
        orc_bbox_overlaps(ad, A, gd, G, ov);                                   /* (1) :38 */
        for (long a = 0; a < A; ++a) {                                         /* (2) :42-43 argmax: first max */
            int best = 0;
            for (int g = 1; g < G; ++g) if (ov[a * G + g] > ov[a * G + best]) best = g;
            assign[a] = best; maxov[a] = ov[a * G + best];
        }
        for (int g = 0; g < G; ++g) {                                          /* (3) :44-46 */
            double m = ov[g];
            for (long a = 1; a < A; ++a) if (ov[a * G + g] > m) m = ov[a * G + g];
            gtmax[g] = m;
        }
        for (long a = 0; a < A; ++a) {                                         /* (4) :49-53 */
            float l = gt[assign[a] * 5 + 4];
            if (maxov[a] < bg_thr) l = 0.0f;
            if (maxov[a] < fg_thr && maxov[a] >= bg_thr) l = -1.0f;
            labels[a] = l;
            if (maxov[a] >= 0.5) flags[a] = 1;
        }
        for (long a = 0; a < A; ++a) {                                         /* (5) :47,56 */
            int hit = 0;
            for (int g = 0; g < G; ++g) if (ov[a * G + g] == gtmax[g]) { hit = 1; break; }
            if (hit) labels[a] = gt[assign[a] * 5 + 4];
        }
        for (long a = 0; a < A; ++a) {                                         /* (6) :60-62 */
            if (clobber_positives && maxov[a] < bg_thr) labels[a] = 0.0f;
            if (labels[a] >= 1.0f) flags[a] = 1;
        }
        /* (7) ignore regions :77-87 — bbox_intersections(ignored[D], anchors[A]) -> [D,A],
         * i.e. inter / area(anchor); summed over D in row order (np.sum(axis=0)). */
        int D = 0;
        for (int g = 0; g < G; ++g) if (gt[g * 5 + 4] < 0) ++D;
        if (D > 0) {
            double *ig = (double *)malloc((size_t)D * 4 * sizeof(double));
            double *inter = (double *)calloc((size_t)D * A, sizeof(double));
            int d = 0;
            for (int g = 0; g < G; ++g)
                if (gt[g * 5 + 4] < 0) { for (int c = 0; c < 4; ++c) ig[d * 4 + c] = (double)gt[g * 5 + c]; ++d; }
            orc_bbox_intersections(ig, D, ad, A, inter);
            for (long a = 0; a < A; ++a) {
                double s = 0.0;
                for (int dd = 0; dd < D; ++dd) s = s + inter[(size_t)dd * A + a];
                if (ignore_sum_out) ignore_sum_out[a] = s;
                if (s > ignore_frac) { labels[a] = -1.0f; flags[a] = 0; }
            }
            free(ig); free(inter);
        } else if (ignore_sum_out) {
            memset(ignore_sum_out, 0, (size_t)A * sizeof(double));
        }
        /* (9) :133-135 bbox_targets for EVERY anchor */
        for (long a = 0; a < A; ++a) {
            const float *g = gt + assign[a] * 5;
            if (box_encoding == 0) enc_fastrcnn_f32(anchors + a * 4, g, bbox_targets + a * 4);
            else orc_box_encode(anchors + a * 4, g, 1, 1, bbox_targets + a * 4);
        }
        if (max_overlaps_out) memcpy(max_overlaps_out, maxov, (size_t)A * sizeof(double));
        if (assignment_out) memcpy(assignment_out, assign, (size_t)A * sizeof(int32_t));
        free(ad); free(gd); free(ov); free(assign); free(maxov); free(gtmax);
    } else {
        memset(labels, 0, (size_t)A * sizeof(float));                          /* :89-91 */
        if (max_overlaps_out) memset(max_overlaps_out, 0, (size_t)A * sizeof(double));
        if (assignment_out) memset(assignment_out, 0, (size_t)A * sizeof(int32_t));
        if (ignore_sum_out) memset(ignore_sum_out, 0, (size_t)A * sizeof(double));
    }
    /* (8) 'simple' label weights :111-118 */
    long nfg = 0, nbg = 0;
    for (long a = 0; a < A; ++a) { if (labels[a] >= 1.0f) ++nfg; if (labels[a] == 0.0f) ++nbg; }
    double fgm = (double)nfg > 1.0 ? (double)nfg : 1.0;
    long mx = nbg > nfg ? nbg : nfg;
    double den = (double)mx > 1.0 ? (double)mx : 1.0;
    float wbg = (float)(3 * fgm / den);
    for (long a = 0; a < A; ++a) {
        if (labels[a] >= 1.0f) label_weights[a] = 1.0f;
        if (labels[a] == 0.0f) label_weights[a] = wbg;
        if (flags[a] == 1)                                                     /* (10) :136-138 */
            for (int c = 0; c < 4; ++c) bbox_inside_weights[a * 4 + c] = bbweights[c];
    }
    free(flags);
}

/* ------------------------------------------------------------------------------------------
 * a11-a14  greedy NMS.
 *   cmp_mode 0: libs/nms/cpu_nms.pyx:17-68    (double)ovr >= thresh   (thresh a Python float)
 *   cmp_mode 1: libs/layers/nms/src/nms_cpu.c:21-86   ovr >= (float)thresh
 *   cmp_mode 2: libs/nms/nms_kernel.cu:24-32,83 / py_cpu_nms.py:35   ovr > (float)thresh
 * IoU in float32, (area_i + area_j) - inter, IEEE division.  `order` = indices sorted by
 * descending score (ties: the reference's argsort/torch.sort are unstable; callers pass a
 * stable order, lowest index first).  Returns the number kept; keep[] holds ORIGINAL indices
 * in score order.
 * ---------------------------------------------------------------------------------------- */
ORC_API long orc_nms(const float *boxes, long box_stride, const int64_t *order, long n,
                     double thresh, int cmp_mode, int64_t *keep)
{
    unsigned char *sup = (unsigned char *)calloc((size_t)n + 1, 1);
    float *areas = (float *)malloc(((size_t)n + 1) * sizeof(float));
    for (long i = 0; i < n; ++i) {
        const float *b = boxes + i * box_stride;
        areas[i] = (b[2] - b[0] + 1) * (b[3] - b[1] + 1);                      /* cpu_nms.pyx:24 */
    }
    float thr32 = (float)thresh;
    long nk = 0;
    for (long _i = 0; _i < n; ++_i) {
        long i = order[_i];
        if (sup[i]) continue;
        keep[nk++] = i;
        const float *bi = boxes + i * box_stride;
        float ix1 = bi[0], iy1 = bi[1], ix2 = bi[2], iy2 = bi[3], iarea = areas[i];
        for (long _j = _i + 1; _j < n; ++_j) {
            long j = order[_j];
            if (sup[j]) continue;
            const float *bj = boxes + j * box_stride;
            float xx1 = ix1 >= bj[0] ? ix1 : bj[0];                            /* :57-60 */
            float yy1 = iy1 >= bj[1] ? iy1 : bj[1];
            float xx2 = ix2 <= bj[2] ? ix2 : bj[2];
            float yy2 = iy2 <= bj[3] ? iy2 : bj[3];
            float w = xx2 - xx1 + 1; w = 0.0f >= w ? 0.0f : w;                 /* :61-62 max(0.0, .) */
            float h = yy2 - yy1 + 1; h = 0.0f >= h ? 0.0f : h;
            float inter = w * h;
            float ovr = inter / (iarea + areas[j] - inter);                    /* :63-64 */
            int s;
            if (cmp_mode == 0) s = (double)ovr >= thresh;
            else if (cmp_mode == 1) s = ovr >= thr32;
            else s = ovr > thr32;
            if (s) sup[j] = 1;
        }
    }
    free(sup); free(areas);
    return nk;
}

/* ------------------------------------------------------------------------------------------
 * a10  objectness — libs/nets/model.py:54-61.  activation 0 softmax: 1.f - p[:,0];
 *      1 sigmoid: max_c p[:,c].  float32.
 * ---------------------------------------------------------------------------------------- */
ORC_API void orc_objectness(const float *prob, long n, int C, int activation, float *out)
{
    for (long i = 0; i < n; ++i) {
        const float *p = prob + i * C;
        if (activation == 0) out[i] = 1.0f - p[0];
        else { float m = p[0]; for (int c = 1; c < C; ++c) if (p[c] > m) m = p[c]; out[i] = m; }
    }
}

/* ------------------------------------------------------------------------------------------
 * a18  bin-centre RoIAlign forward — wrapper arithmetic of
 *      libs/layers/roi_align_tf/roi_align_tf.py:24-40 (torch float32 ops, this order) followed
 *      by libs/layers/roi_align_tf/src/crop_and_resize.c:6-112, extrapolation value 0.
 *      feat [B,C,H,W] f32 (NCHW), rois [R,5]=(b,x1,y1,x2,y2) image coords -> out [R,C,PH,PW].
 *      A RoI whose batch index is outside [0,B) yields zeros (the reference exits the process,
 *      crop_and_resize.c:39-42; the CUDA twin skips it, crop_and_resize_kernel.cu:35-38).
 * ---------------------------------------------------------------------------------------- */
static void tf_box(const float *roi, float s, int PH, int PW, int H, int W,
                   float *y1, float *x1, float *y2, float *x2)
{
    float bx1 = s * roi[1], by1 = s * roi[2], bx2 = s * roi[3], by2 = s * roi[4];   /* :26 */
    float sw = (bx2 - bx1) / (float)PW, sh = (by2 - by1) / (float)PH;               /* :30-31 */
    float nx0 = (bx1 + sw / 2 - 0.5f) / (float)(W - 1);                             /* :34 */
    float ny0 = (by1 + sh / 2 - 0.5f) / (float)(H - 1);
    float nw = sw * (float)(PW - 1) / (float)(W - 1);                               /* :37 */
    float nh = sh * (float)(PH - 1) / (float)(H - 1);
    *y1 = ny0; *x1 = nx0; *y2 = ny0 + nh; *x2 = nx0 + nw;                           /* :40 */
}

/* rois [R,5] -> normalised (y1,x1,y2,x2) boxes [R,4] + int32 batch index [R] */
ORC_API void orc_tf_boxes(const float *rois, long R, float spatial_scale, int PH, int PW,
                          int H, int W, float *boxes, int32_t *box_ind)
{
    for (long r = 0; r < R; ++r) {
        tf_box(rois + r * 5, spatial_scale, PH, PW, H, W,
               boxes + r * 4, boxes + r * 4 + 1, boxes + r * 4 + 2, boxes + r * 4 + 3);
        box_ind[r] = (int32_t)rois[r * 5];                                          /* .int() :25 */
    }
}

/* crop_and_resize.c:6-112 — boxes normalised (y1,x1,y2,x2); out [R,C,PH,PW] */
ORC_API void orc_crop_and_resize_fwd(const float *feat, int B, int C, int H, int W,
                                     const float *boxes, const int32_t *box_ind, long R,
                                     float extrapolation, int PH, int PW, float *out)
{
    #pragma omp parallel for schedule(dynamic, 8)
    for (long r = 0; r < R; ++r) {
        const float y1 = boxes[r * 4], x1 = boxes[r * 4 + 1], y2 = boxes[r * 4 + 2], x2 = boxes[r * 4 + 3];
        const int b_in = box_ind[r];
        float *o = out + (size_t)r * C * PH * PW;
        if (b_in < 0 || b_in >= B) { memset(o, 0, (size_t)C * PH * PW * sizeof(float)); continue; }
        const float hs = (PH > 1) ? (y2 - y1) * (H - 1) / (PH - 1) : 0;             /* c:44-50 */
        const float ws = (PW > 1) ? (x2 - x1) * (W - 1) / (PW - 1) : 0;
        for (int y = 0; y < PH; ++y) {
            const float in_y = (PH > 1) ? y1 * (H - 1) + y * hs : 0.5 * (y1 + y2) * (H - 1);  /* :54-56 */
            if (in_y < 0 || in_y > H - 1) {
                for (int x = 0; x < PW; ++x) for (int d = 0; d < C; ++d) o[(size_t)d * PH * PW + y * PW + x] = extrapolation;
                continue;
            }
            const int top = floorf(in_y), bot = ceilf(in_y);                        /* :70-72 */
            const float yl = in_y - top;
            for (int x = 0; x < PW; ++x) {
                const float in_x = (PW > 1) ? x1 * (W - 1) + x * ws : 0.5 * (x1 + x2) * (W - 1);
                if (in_x < 0 || in_x > W - 1) {
                    for (int d = 0; d < C; ++d) o[(size_t)d * PH * PW + y * PW + x] = extrapolation;
                    continue;
                }
                const int left = floorf(in_x), right = ceilf(in_x);
                const float xl = in_x - left;
                for (int d = 0; d < C; ++d) {
                    const float *p = feat + ((size_t)b_in * C + d) * H * W;
                    const float tl = p[top * W + left], tr = p[top * W + right];
                    const float bl = p[bot * W + left], br = p[bot * W + right];
                    const float t = tl + (tr - tl) * xl;                            /* :101-105 */
                    const float bo = bl + (br - bl) * xl;
                    o[(size_t)d * PH * PW + y * PW + x] = t + (bo - t) * yl;
                }
            }
        }
    }
}

/* a18 backward — crop_and_resize.c:164-268 (serial, box order r=0..R-1, so the float32
 *     accumulation order is defined: r -> y -> x -> d).  grad_feat must be zeroed by the caller
 *     (the reference zeroes it at :206). */
ORC_API void orc_crop_and_resize_bwd(const float *grad_out, int B, int C, int H, int W,
                                     const float *boxes, const int32_t *box_ind, long R,
                                     int PH, int PW, float *grad_feat)
{
    for (long r = 0; r < R; ++r) {
        const float y1 = boxes[r * 4], x1 = boxes[r * 4 + 1], y2 = boxes[r * 4 + 2], x2 = boxes[r * 4 + 3];
        const int b_in = box_ind[r];
        if (b_in < 0 || b_in >= B) continue;
        const float *g = grad_out + (size_t)r * C * PH * PW;
        const float hs = (PH > 1) ? (y2 - y1) * (H - 1) / (PH - 1) : 0;
        const float ws = (PW > 1) ? (x2 - x1) * (W - 1) / (PW - 1) : 0;
        for (int y = 0; y < PH; ++y) {
            const float in_y = (PH > 1) ? y1 * (H - 1) + y * hs : 0.5 * (y1 + y2) * (H - 1);
            if (in_y < 0 || in_y > H - 1) continue;
            const int top = floorf(in_y), bot = ceilf(in_y);
            const float yl = in_y - top;
            for (int x = 0; x < PW; ++x) {
                const float in_x = (PW > 1) ? x1 * (W - 1) + x * ws : 0.5 * (x1 + x2) * (W - 1);
                if (in_x < 0 || in_x > W - 1) continue;
                const int left = floorf(in_x), right = ceilf(in_x);
                const float xl = in_x - left;
                for (int d = 0; d < C; ++d) {
                    float *p = grad_feat + ((size_t)b_in * C + d) * H * W;
                    const float gv = g[(size_t)d * PH * PW + y * PW + x];
                    const float dtop = (1 - yl) * gv;                               /* :258-264 */
                    p[top * W + left] += (1 - xl) * dtop;
                    p[top * W + right] += xl * dtop;
                    const float dbot = yl * gv;
                    p[bot * W + left] += (1 - xl) * dbot;
                    p[bot * W + right] += xl * dbot;
                }
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * a19  grid-point RoIAlign — libs/layers/roi_align/src/cuda/roi_align_kernel.cu:15-68 (fwd),
 *      :97-153 (bwd).  CUDA-only upstream; this restates the kernel body one output element at
 *      a time, keeping the double-literal promotions (1. / 0.) of :38-41, :62-65, :147-150:
 *      weights and the 4-term sum are evaluated in double, then rounded to float.
 *      h = (float)ph*bin + start is left unfused here (nvcc may contract it in the reference
 *      build; the difference is below the 1e-5 tolerance except exactly on a tap boundary).
 * ---------------------------------------------------------------------------------------- */
typedef struct { int valid; int hs, ws; float hr, wr; } gp_tap;

static gp_tap gp_coords(const float *roi, float s, int H, int W, int PH, int PW, int ph, int pw)
{
    gp_tap t;
    float sw = roi[1] * s, sh = roi[2] * s, ew = roi[3] * s, eh = roi[4] * s;       /* :32-35 */
    float rw = fmaxf(ew - sw + 1., 0.), rh = fmaxf(eh - sh + 1., 0.);               /* :38-39 */
    float bh = rh / (PH - 1.), bw = rw / (PW - 1.);                                 /* :40-41 */
    float h = (float)ph * bh + sh, w = (float)pw * bw + sw;                         /* :43-44 */
    t.hs = fminf(floor(h), H - 2);                                                  /* :46-47 */
    t.ws = fminf(floor(w), W - 2);
    t.valid = !(h < 0 || h >= H || w < 0 || w >= W);                                /* :52 */
    t.hr = h - (float)t.hs; t.wr = w - (float)t.ws;                                 /* :55-56 */
    return t;
}

ORC_API void orc_roi_align_grid_fwd(const float *feat, int B, int C, int H, int W,
                                    const float *rois, long R, float spatial_scale,
                                    int PH, int PW, float *out)
{
    #pragma omp parallel for schedule(dynamic, 8)
    for (long r = 0; r < R; ++r) {
        int b = (int)rois[r * 5];
        for (int ph = 0; ph < PH; ++ph)
            for (int pw = 0; pw < PW; ++pw) {
                gp_tap t = gp_coords(rois + r * 5, spatial_scale, H, W, PH, PW, ph, pw);
                for (int c = 0; c < C; ++c) {
                    float *o = out + (((size_t)r * C + c) * PH + ph) * PW + pw;
                    if (!t.valid || b < 0 || b >= B) { *o = 0.0f; continue; }
                    const float *p = feat + (((size_t)b * C + c) * H + t.hs) * W + t.ws;
                    *o = p[0] * (1. - t.hr) * (1. - t.wr) + p[1] * (1. - t.hr) * t.wr   /* :62-65 */
                       + p[W] * t.hr * (1. - t.wr) + p[W + 1] * t.hr * t.wr;
                }
            }
    }
}

ORC_API void orc_roi_align_grid_bwd(const float *grad_out, int B, int C, int H, int W,
                                    const float *rois, long R, float spatial_scale,
                                    int PH, int PW, float *grad_feat)
{
    for (long r = 0; r < R; ++r) {
        int b = (int)rois[r * 5];
        if (b < 0 || b >= B) continue;
        for (int c = 0; c < C; ++c)
            for (int ph = 0; ph < PH; ++ph)
                for (int pw = 0; pw < PW; ++pw) {
                    gp_tap t = gp_coords(rois + r * 5, spatial_scale, H, W, PH, PW, ph, pw);
                    if (!t.valid) continue;
                    float g = grad_out[(((size_t)r * C + c) * PH + ph) * PW + pw];
                    float *p = grad_feat + (((size_t)b * C + c) * H + t.hs) * W + t.ws;
                    p[0] += (float)(g * (1. - t.hr) * (1 - t.wr));                       /* :147-150 */
                    p[1] += (float)(g * (1. - t.hr) * t.wr);
                    p[W] += (float)(g * t.hr * (1 - t.wr));
                    p[W + 1] += (float)(g * t.hr * t.wr);
                }
    }
}

/* a20  FPN level assignment — libs/layers/roi_align_tf/pyramid_roi_align.py:33-37 (same in
 *      roi_align/modules/pyramid_roi_align.py:36-40): float32,
 *      argmin_l |log(sqrt(area/ref_l)) / 0.6931472|, first minimum wins. */
ORC_API void orc_assign_levels(const float *boxes, long R, const float *ref_areas, int L, int32_t *levels)
{
    for (long r = 0; r < R; ++r) {
        const float *b = boxes + r * 4;
        float area = (b[2] - b[0] + 1.0f) * (b[3] - b[1] + 1.0f);
        int best = 0; float bv = 0;
        for (int l = 0; l < L; ++l) {
            float v = fabsf(logf(sqrtf(area / ref_areas[l])) / 0.6931472f);
            if (l == 0 || v < bv) { bv = v; best = l; }
        }
        levels[r] = best;
    }
}
