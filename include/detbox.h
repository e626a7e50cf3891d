/* detbox.h — C ABI of libdetbox.so: the box-and-RoI hot path of CharlesShang/Detectron-PYTORCH
 * as hand-written sm_100a CUDA kernels.
 *
 * Conventions (all entry points):
 *   - extern "C", plain pointers and sizes; no torch / TH types.
 *   - pointers named *_dev are DEVICE pointers on the current CUDA device; *_host are host
 *     pointers.  Arrays are dense, row-major, float32 unless the name says otherwise.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Device
 *     entry points only enqueue work: they never synchronise and never allocate; scratch
 *     memory comes from the caller, sized by the matching *_workspace_bytes() query.
 *   - return value: DETBOX_OK (0) or a negative DETBOX_ERR_* code; detbox_error_string()
 *     describes it.  Nothing ever calls exit() (the reference's launchers do:
 *     roi_align_kernel.cu:87-91, crop_and_resize_kernel.cu:186-191).
 *   - "reference" file:line citations are relative to the Detectron-PYTORCH tree.
 *
 * The *_host entry points at the end take HOST buffers and do the copies themselves; they are
 * what the reference's own FFI for this path binds (gpu_nms.hpp:1-2) and what bench.py's
 * end-to-end leg times.
 */
#ifndef DETBOX_H
#define DETBOX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DETBOX_OK               0
#define DETBOX_ERR_INVALID     -1   /* bad argument (null pointer, negative size, unknown mode) */
#define DETBOX_ERR_WORKSPACE   -2   /* workspace missing or smaller than *_workspace_bytes()   */
#define DETBOX_ERR_CUDA        -3   /* a CUDA runtime call or launch failed                    */
#define DETBOX_ERR_UNSUPPORTED -4   /* size beyond a documented limit                          */

/* box encodings: reference cfg.rpn_box_encoding (libs/configs/config.py:106) + 'simple' of
 * libs/layers/box.py:102,134 */
#define DETBOX_ENC_FASTRCNN 0
#define DETBOX_ENC_LINEAR   1
#define DETBOX_ENC_SIMPLE   2

/* class activation: libs/nets/model.py:54-61 */
#define DETBOX_ACT_SOFTMAX 0    /* objectness = 1 - p[:,0]   */
#define DETBOX_ACT_SIGMOID 1    /* objectness = max_c p[:,c] */

/* NMS suppression comparator */
#define DETBOX_NMS_GE_F64 0     /* (double)iou >= thr        libs/nms/cpu_nms.pyx:65          */
#define DETBOX_NMS_GE_F32 1     /* iou >= (float)thr         libs/layers/nms/src/nms_cpu.c:76 */
#define DETBOX_NMS_GT_F32 2     /* iou >  (float)thr         libs/nms/nms_kernel.cu:83        */
/* OR-ed into cmp_mode: the rows already are in processing order — segment ids non-decreasing and
 * scores non-increasing inside a segment (what detbox_topk_decode(per_image=1) emits) — so the
 * sort and the gather are skipped.  Needs box_stride == 4 and 16-byte aligned boxes. */
#define DETBOX_NMS_PRESORTED 0x100
/* OR-able into cmp_mode: seg_ids_dev points at int64 ids (what detbox_topk_decode writes as img_ids_dev);
 * they are narrowed inside the segment-count pass, so the caller needs no conversion kernel. */
#define DETBOX_NMS_SEG_I64 0x200
/* suppression rule variants of libs/boxes/nms.pyx (detbox_nms_variant) */
#define DETBOX_NMS_VARIANT_PLAIN       0   /* nms             nms.pyx:17-68   IoU only                                    */
#define DETBOX_NMS_VARIANT_NEW         1   /* nms_new         nms.pyx:70-123  + inter/area_i > 0.95 or inter/area_j > 0.95 */
#define DETBOX_NMS_VARIANT_MULTIANCHOR 2   /* nms_multianchor nms.pyx:125-195 (inter+0.1)/(union+0.1), centre-distance rule */

/* RoIAlign sampling convention / level mode / fused 2x2 pooling / layout */
#define DETBOX_ROI_BIN_CENTRE 0 /* roi_align_tf (crop_and_resize.c + roi_align_tf.py:24-40)   */
#define DETBOX_ROI_GRID_POINT 1 /* roi_align/src/cuda/roi_align_kernel.cu:15-68              */
#define DETBOX_ROI_CROP_NORM  2 /* raw crop_and_resize (crop_and_resize.c:6-112): rois rows are
                                   (box_ind, y1, x1, y2, x2) with NORMALISED coordinates          */
#define DETBOX_POOL_NONE 0
#define DETBOX_POOL_AVG2 1      /* RoIAlignAvg roi_align/modules/roi_align.py:23-33           */
#define DETBOX_POOL_MAX2 2      /* RoIAlignMax roi_align/modules/roi_align.py:36-46           */
#define DETBOX_LAYOUT_NCHW 0
#define DETBOX_LAYOUT_NHWC 1    /* torch.channels_last memory, same logical [B,C,H,W] tensor  */
#define DETBOX_MAX_LEVELS 8

const char *detbox_error_string(int code);
int detbox_version(void);
/* last cudaError_t seen by this library on the calling thread (0 if none) */
int detbox_last_cuda_error(void);

/* Measurement hook (bench.py's roofline leg): record the two CUDA events (cudaEvent_t passed as
 * void*) on the launching stream immediately before and after every launch of ONE kernel,
 * chosen by id, made from the calling thread.  kernel_id 0 switches the hook off.  Two host-side
 * cudaEventRecord calls, nothing else; results are unaffected. */
#define DETBOX_K_NONE            0
#define DETBOX_K_ANCHORS         1   /* anchor_pyramid_kernel                       */
#define DETBOX_K_TARGET_COLMAX   2   /* colmax_kernel            per-GT column max  */
#define DETBOX_K_TARGET_ASSIGN   3   /* anchor_target_kernel     labels + targets   */
#define DETBOX_K_TARGET_WEIGHTS  4   /* label_weight_kernel                         */
#define DETBOX_K_OBJECTNESS      5   /* objectness_kernel (score keys + histogram)  */
#define DETBOX_K_NMS_MASK        6   /* mask_kernel                                 */
#define DETBOX_K_NMS_SCAN        7   /* nms_resolve_kernel (fixed-point keep set)    */
#define DETBOX_K_ROI_FWD         8   /* roi_fwd_kernel                              */
#define DETBOX_K_ROI_BWD         9   /* roi_bwd_kernel (without the zero fill)      */
#define DETBOX_K_ROI_ZERO       10   /* zero fill of the dense gradient maps        */
int detbox_trace_kernel(int kernel_id, void *start_event, void *stop_event);
/* Diagnostics: the in-range reciprocal of the encode (box_math.cuh::rcp_rn_inrange) against the IEEE
 * round-to-nearest reciprocal for every float whose bit pattern lies in [bits_lo, bits_hi];
 * *mismatches_dev (device, uint64) receives the number of differing results. */
int detbox_selftest_rcp(uint32_t bits_lo, uint32_t bits_hi, unsigned long long *mismatches_dev, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Anchors — replaces libs/boxes/anchor.py:101-189 (anchors / generate_anchors) and :30-70
 * (anchor_pyramid / anchor_one_plane), and cython_anchor.anchors_plane (cython_anchor.pyx:14-48).
 * ------------------------------------------------------------------------------------------- */
/* HOST: base anchors at one cell, float64 [n_ratios*n_scales,4], ratio-major / scale-minor. */
int detbox_base_anchors_host(const double *scales, int n_scales, const double *ratios, int n_ratios,
                             long base_size, double *out_host);

/* DEVICE: expand per-level cell anchors over the level grids.  cell_dev: float64
 * [sum_l n_cell[l], 4] = base anchors (+ shift*stride), level after level.  Output rows are
 * ordered level -> y -> x -> cell anchor; written as float32 (out_f64 == 0) or float64. */
int detbox_anchor_pyramid(const double *cell_dev, int n_levels, const int *n_cell_host,
                          const int *heights_host, const int *widths_host, const int *strides_host,
                          void *out_dev, int out_f64, void *stream);

/* ---------------------------------------------------------------------------------------------
 * IoU matrices — replaces libs/boxes/bbox.pyx:15-55 / :57-94 / :96-135.  float64 in, float64
 * out, identical operation order (bit-exact).
 * ------------------------------------------------------------------------------------------- */
int detbox_bbox_overlaps_f64(const double *boxes_dev, long n, const double *query_dev, long k,
                             double *out_dev /*[n,k]*/, void *stream);
int detbox_bbox_intersections_f64(const double *boxes_dev, long n, const double *query_dev, long k,
                                  double *out_dev /*[n,k]*/, void *stream);
int detbox_bbox_exclude_ignored_f64(const double *boxes_dev, long n, const double *ignored_dev, long k,
                                    double threshold, int64_t *out_dev /*[n]*/, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Box encode / decode — replaces libs/boxes/bbox_transform.py:11-41,43-76,78-91,93-144 and
 * libs/layers/box.py:72-145 (decoding_box / encoding_box).  float32.
 * decode clips to [0,im_w-1]x[0,im_h-1] when im_h > 0 (clip_boxes), else no clip (torch path).
 * ------------------------------------------------------------------------------------------- */
int detbox_box_encode(const float *anchors_dev, const float *gt_dev, long n, int encoding,
                      float *out_dev, void *stream);
int detbox_box_decode(const float *anchors_dev, const float *deltas_dev, long n, int encoding,
                      int im_h, int im_w, float *out_dev, void *stream);
/* in-place clip_boxes (bbox_transform.py:78-91) on [n, 4*k] rows */
int detbox_clip_boxes(float *boxes_dev, long n, int k, int im_h, int im_w, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Anchor-target assignment — replaces libs/layers/anchor.py:17-145 (encode, 'simple' sampling)
 * and its batch wrapper libs/layers/data_layer.py:228-239; supersedes the disabled
 * libs/layers/anchor_target CUDA layer.  Fused: no [A,G] overlap matrix is materialised.
 * IoU in float64 (bit-exact labels), targets in float32.
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    double bg_threshold;        /* cfg.rpn_bg_threshold */
    double fg_threshold;        /* cfg.rpn_fg_threshold */
    double ignore_fraction;     /* cfg.ignored_area_intersection_fraction */
    int    clobber_positives;   /* cfg.rpn_clobber_positives */
    int    box_encoding;        /* DETBOX_ENC_FASTRCNN | DETBOX_ENC_LINEAR */
    float  bbweights[4];        /* cfg.bbweights */
} detbox_target_cfg;

size_t detbox_anchor_target_workspace_bytes(long A, int batch, long total_gt);
/* anchors_dev [A,4] shared by the batch; gt_dev [total_gt,5] = (x1,y1,x2,y2,cls), images
 * concatenated; gt_offsets_dev int32 [batch+1].  Outputs [batch,A], [batch,A], [batch,A,4],
 * [batch,A,4].  counts_dev (optional, may be NULL): int32 [batch,2] = (#fg, #bg). */
int detbox_anchor_target(const float *anchors_dev, long A, const float *gt_dev,
                         const int32_t *gt_offsets_dev, int batch, long total_gt,
                         const detbox_target_cfg *cfg,
                         float *labels_dev, float *label_weights_dev, float *bbox_targets_dev,
                         float *bbox_inside_weights_dev, int32_t *counts_dev,
                         void *workspace_dev, size_t workspace_bytes, void *stream);

/* The same assignment with two additions (round 2):
 *
 * pyr (optional): the grid geometry the anchors were expanded over — rows ordered level -> y -> x -> cell anchor, what
 *   detbox_anchor_pyramid / libs/boxes/anchor.py:30-70 produce; the descriptor must describe exactly A rows.  The
 *   fastrcnn encode (bbox_transform.py:18-37) is separable per axis and anchor (l, y, x, c) shares its x-interval with
 *   (l, 0, x, c) and its y-interval with (l, y, 0, c), so the targets of every anchor matched to GT 0 of its image (all
 *   anchors that overlap nothing: np.argmax of an all-zero row) come from per-image tables of sum_l n_cell * (W_l + H_l)
 *   half-encodes.  Separability is VERIFIED on the device on every call; if the check fails the generic path runs, so
 *   the results never depend on the descriptor being right.  Bit-identical to pyr == NULL.
 *
 * format: DETBOX_TARGETS_REFERENCE — the four float32 arrays of libs/layers/anchor.py:17-145 (packed_dev unused);
 *   DETBOX_TARGETS_COMPACT — bbox_targets [batch,A,4] float32 as before, plus packed_dev uint8 [batch,A] =
 *   (label + 1) | flag << 7 and counts_dev int32 [batch,2] = (#fg, #bg) (mandatory): bbox_inside_weights is
 *   flag ? cfg.bbweights : 0 (anchor.py:136-138) and label_weights a function of the label and the image's counts
 *   (anchor.py:111-118), so 17 instead of 40 bytes per anchor carry the same information
 *   (detbox_expand_targets_host rebuilds the reference's arrays bit for bit).  GT classes must be integers in
 *   [-1, 125]; any other class is packed as label code 127, which the expander reports. */
typedef struct {
    int n_levels;
    int n_cell[DETBOX_MAX_LEVELS];          /* anchors per grid cell (shifts x ratios x scales) */
    int height[DETBOX_MAX_LEVELS], width[DETBOX_MAX_LEVELS];
} detbox_pyramid_desc;
#define DETBOX_TARGETS_REFERENCE 0
#define DETBOX_TARGETS_COMPACT   1
size_t detbox_anchor_target_ex_workspace_bytes(long A, int batch, long total_gt, const detbox_pyramid_desc *pyr);
int detbox_anchor_target_ex(const float *anchors_dev, long A, const detbox_pyramid_desc *pyr,
                            const float *gt_dev, const int32_t *gt_offsets_dev, int batch, long total_gt,
                            const detbox_target_cfg *cfg, int format,
                            float *labels_dev, float *label_weights_dev, float *bbox_targets_dev,
                            float *bbox_inside_weights_dev, uint8_t *packed_dev, int32_t *counts_dev,
                            void *workspace_dev, size_t workspace_bytes, void *stream);
/* The collate contract of a device-side target pipeline (libs/datasets/citypersons2.py:1045-1070 pads per-image GT;
 * libs/layers/data_layer.py:228-239 loops over the list on the host): gt_padded_dev [batch, gmax, 5] + gt_counts_dev
 * int32 [batch] -> the concatenated rows gt_out_dev [<= batch*gmax, 5] and gt_offsets_out_dev int32 [batch+1] that
 * detbox_anchor_target takes.  Counts are clamped to [0, gmax].  No host round trip. */
int detbox_pack_padded_gt(const float *gt_padded_dev, const int32_t *gt_counts_dev, int batch, int gmax,
                          float *gt_out_dev, int32_t *gt_offsets_out_dev, void *stream);
/* HOST: DETBOX_TARGETS_COMPACT -> labels [batch,A], label_weights [batch,A], bbox_inside_weights [batch,A,4]
 * (any may be NULL) with n_threads host threads.  DETBOX_ERR_UNSUPPORTED if a label code 127 is met. */
int detbox_expand_targets_host(const uint8_t *packed_host, const int32_t *counts_host, long A, int batch,
                               const float *bbweights_host /*[4]*/, float *labels_host, float *label_weights_host,
                               float *bbox_inside_weights_host, int n_threads);

/* ---------------------------------------------------------------------------------------------
 * Objectness + top-k + gather + decode — replaces detection_model._decode_and_choose_top_n_stage1
 * (libs/nets/model.py:308-321) incl. _objectness (:54-61).  One stable descending selection
 * over batch x levels x anchors (per_image == 0, the reference's behaviour) or inside every
 * image (per_image == 1, the sharded mode).  Ties: lower flat index first.
 * Outputs are score-descending (image-major when per_image): boxes [K,4], probs [K,C],
 * img_ids int64 [K], sel_anchors [K,4], flat_idx int64 [K], scores [K];  K = min(top_n, batch*A)
 * (per_image: K = batch * min(top_n, A)).  Any output pointer except boxes may be NULL.
 * ------------------------------------------------------------------------------------------- */
size_t detbox_topk_decode_workspace_bytes(int batch, long A, long top_n, int per_image);
int detbox_topk_decode(const float *prob_dev, int C, const float *deltas_dev, const float *anchors_dev,
                       int batch, long A, int activation, long top_n, int per_image, int encoding,
                       float *boxes_dev, float *probs_dev, int64_t *img_ids_dev, float *sel_anchors_dev,
                       int64_t *flat_idx_dev, float *scores_dev,
                       void *workspace_dev, size_t workspace_bytes, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Stable radix sort of float32 scores, descending (ties: lower index first).  Exposed because
 * the reference's callers sort explicitly (libs/layers/box.py:160, libs/nms/gpu_nms.pyx:25).
 * order_dev int32 [n] receives the permutation (one segment; segmented selection is detbox_topk_decode's job).
 * ------------------------------------------------------------------------------------------- */
size_t detbox_sort_desc_workspace_bytes(long n);
int detbox_sort_desc(const float *scores_dev, long n, int32_t *order_dev,
                     void *workspace_dev, size_t workspace_bytes, void *stream);

/* ---------------------------------------------------------------------------------------------
 * NMS — replaces libs/nms/gpu_nms.pyx + nms_kernel.cu:46-155, libs/nms/cpu_nms.pyx:17-68,
 * libs/layers/box.py:148-173 (apply_nms) + libs/layers/nms/src/{nms_cpu.c,nms_cuda.c,
 * nms_cuda_kernel.cu}, and the per-image loop of libs/nets/model.py:343-355
 * (_apply_nms_in_batch) through segments.  Sorting, bitmask and the greedy scan all run on the
 * device; nothing is copied to the host.
 * boxes_dev [n, box_stride] (x1,y1,x2,y2 first), scores_dev [n].  seg_ids_dev: int32 [n], the
 * segment (image id) of every row, values in [0,n_seg); rows of different segments never
 * suppress each other (NULL or n_seg <= 1: one segment).  Rows need not be grouped.  An id outside
 * [0,n_seg) is a caller error: it is never used as an index — the row is treated as part of the nearest
 * valid segment (0 or n_seg-1), memory-safe, its own keep decision unspecified.
 * keep_dev int64 [n]: kept ORIGINAL row indices, segment-major, score-descending inside a
 * segment.  n_keep_dev int32 [n_seg+1]: [0..n_seg) per-segment counts, [n_seg] the total
 * (n_seg == 1: both entries hold the total).
 * ------------------------------------------------------------------------------------------- */
size_t detbox_nms_workspace_bytes(long n, int n_seg);
int detbox_nms(const float *boxes_dev, int box_stride, const float *scores_dev, long n,
               const int32_t *seg_ids_dev, int n_seg, double thresh, int cmp_mode,
               int64_t *keep_dev, int32_t *n_keep_dev,
               void *workspace_dev, size_t workspace_bytes, void *stream);
/* The NMS variants of libs/boxes/nms.pyx:70-195 (`nms_new`, `nms_multianchor`; used by
 * single_shot.nms_advanced / nms_stitch) on the same mask + resolve pipeline: `variant` selects the
 * pair rule (DETBOX_NMS_VARIANT_*), cmp_mode the IoU comparison (upstream: DETBOX_NMS_GE_F64).
 * clusters_dev (optional, int64 [n]): nms_multianchor's second result — a kept box's own index, a
 * suppressed box's first kept suppressor (original indices). */
int detbox_nms_variant(const float *boxes_dev, int box_stride, const float *scores_dev, long n,
                       const int32_t *seg_ids_dev, int n_seg, double thresh, int cmp_mode, int variant,
                       int64_t *keep_dev, int32_t *n_keep_dev, int64_t *clusters_dev,
                       void *workspace_dev, size_t workspace_bytes, void *stream);

/* detbox_nms with its scratch memory built into the library — NO counterpart in the reference: added
 * for callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  Same arguments and results as detbox_nms minus the workspace; the scratch is a file-scope
 * __device__ array of 2.25 MiB (csrc/no_alloc.cu) that every GPU loading the library holds until the process exits.
 * Limits: n <= DETBOX_NMS_STATIC_MAX_BOXES and n_seg <= DETBOX_NMS_STATIC_MAX_SEGMENTS, else DETBOX_ERR_UNSUPPORTED
 * ("size beyond a documented limit") and nothing is enqueued.  Calls no cudaMalloc / cudaFree / allocating routine.
 * All calls share the one array (it keeps the last call's contents; what must start at zero is cleared per call), so
 * calls that OVERLAP on the GPU — different streams, concurrent graph branches — race: serialise them.  The stream must
 * belong to the CURRENT device; the first call on a device looks up the array's device address
 * (cudaGetSymbolAddress) — make that call once outside graph capture. */
#define DETBOX_NMS_STATIC_MAX_BOXES    4096
#define DETBOX_NMS_STATIC_MAX_SEGMENTS 64
int detbox_nms_static(const float *boxes_dev, int box_stride, const float *scores_dev, long n,
                      const int32_t *seg_ids_dev, int n_seg, double thresh, int cmp_mode,
                      int64_t *keep_dev, int32_t *n_keep_dev, void *stream);

/* ---------------------------------------------------------------------------------------------
 * RoIAlign over FPN levels, forward and backward — replaces
 *   roi_align_tf: crop_and_resize.c:6-268, crop_and_resize_kernel.cu:10-165, roi_align_tf.py:16-42
 *   roi_align   : roi_align_kernel.cu:15-153, modules/roi_align.py:11-46 (Avg/Max fused)
 *   pyramid     : pyramid_roi_align.py:30-60 (level-assigned) and :82-91 (all levels)
 * One launch covers all levels: level[r] comes from levels_dev (int32 [R]) or, when NULL and
 * n_levels > 1, is computed in-kernel from ref_areas (pyramid_roi_align.py:33-37); output is
 * written in the caller's RoI order.  all_levels != 0: every RoI on every level, channel-
 * concatenated ([R, n_levels*C, PH, PW]).
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    int n_levels;
    const float *feat[DETBOX_MAX_LEVELS];   /* device, [B,C,H_l,W_l] in `layout`          */
    float *grad_feat[DETBOX_MAX_LEVELS];    /* device, same shapes (backward only)         */
    int height[DETBOX_MAX_LEVELS], width[DETBOX_MAX_LEVELS];
    float spatial_scale[DETBOX_MAX_LEVELS];
    float ref_area[DETBOX_MAX_LEVELS];      /* (mean(anchor_scales[l])*anchor_base)^2       */
    int batch, channels;
    int layout;                             /* DETBOX_LAYOUT_*  of feat / grad_feat         */
    int out_layout;                         /* DETBOX_LAYOUT_*  of out / grad_out           */
    int pooled_h, pooled_w;                 /* OUTPUT size (after the optional 2x2 pool)    */
    int mode;                               /* DETBOX_ROI_*                                 */
    int pool;                               /* DETBOX_POOL_*  (grid-point mode only)        */
    int all_levels;
    float extrapolation;                    /* value of samples outside the map (crop modes; the
                                               RoIAlign wrapper passes 0, roi_align_tf.py:42)    */
    const int32_t *roi_order;               /* device, int32 [R] permutation or NULL: block i processes RoI
                                               roi_order[i] (results are written to the RoI's own slot, so they do
                                               not depend on it); see detbox_roi_locality_order                  */
} detbox_roi_cfg;

int detbox_roi_align_forward(const detbox_roi_cfg *cfg, const float *rois_dev /*[R,5] b,x1,y1,x2,y2*/,
                             long R, const int32_t *levels_dev, float *out_dev, void *stream);
/* grad_feat buffers are ACCUMULATED into; zero_grad != 0 clears them first (the reference's
 * contract: functions/roi_align.py:43, crop_and_resize_gpu.c:57). */
int detbox_roi_align_backward(const detbox_roi_cfg *cfg, const float *rois_dev, long R,
                              const int32_t *levels_dev, const float *grad_out_dev, int zero_grad,
                              void *stream);
/* The same gradient as detbox_roi_align_backward(..., zero_grad = 1) — crop_and_resize.c:207-268 (CropAndResizeBackprop),
 * crop_and_resize_kernel.cu:119-165, functions/crop_and_resize.py:34-47 — computed tile by tile of the OUTPUT: every grad_feat
 * buffer is written exactly once, zeros included (no memset, no atomics, no read of the gradient; deterministic summation order:
 * RoI index, bin row, bin column).  NCHW, bin-centre or normalised-crop mode, no pool, one level per RoI, pooled_h * pooled_w <= 320;
 * anything else returns DETBOX_ERR_UNSUPPORTED and the caller uses detbox_roi_align_backward.  Workspace: 16 + 8 * (ph + pw) bytes per RoI. */
size_t detbox_roi_align_backward_tiled_workspace_bytes(const detbox_roi_cfg *cfg, long R);
int detbox_roi_align_backward_tiled(const detbox_roi_cfg *cfg, const float *rois_dev, long R, const int32_t *levels_dev,
                                    const float *grad_out_dev, void *workspace_dev, size_t workspace_bytes, void *stream);
/* A processing order for detbox_roi_cfg.roi_order that keeps overlapping RoIs close together in time: RoIs sorted by
 * (level, image, row band, column band) of their centre on an im_h x im_w canvas, so that the feature rows neighbouring
 * RoIs share are L2 hits instead of DRAM re-reads (the reference launches one thread per output element in RoI order,
 * roi_align_kernel.cu:15-30 — no counterpart).  levels_dev / ref_areas as for the forward.  order_dev int32 [R]. */
size_t detbox_roi_locality_order_workspace_bytes(long R);
int detbox_roi_locality_order(const float *rois_dev, long R, const int32_t *levels_dev, const float *ref_areas_host, int n_levels,
                              float im_h, float im_w, int32_t *order_dev, void *workspace_dev, size_t workspace_bytes, void *stream);
/* level assignment alone (pyramid_roi_align.py:33-37) */
int detbox_assign_levels(const float *boxes_dev /*[R,4]*/, long R, const float *ref_areas_host, int n_levels,
                         int32_t *levels_dev, void *stream);

/* ---------------------------------------------------------------------------------------------
 * RoI -> ground-truth matching, RoI targets, RoI sampling (the MaskRCNN training
 * step right after NMS).  All pointers are device pointers; nothing round-trips through the host.
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    double bg_thresh;      /* match index = -1 below this max overlap   libs/layers/box.py:44        */
    double fg_thresh;      /* label = 0 below this max overlap          libs/layers/box.py:51, 210   */
    double ignore_thresh;  /* < 0: off.  Else rows with inter/area(row) > thresh against a GT of
                              class <= 0 get label -1                   libs/layers/box.py:213-221   */
    float  min_size;       /* 0: off.  Else rows with w or h (+1) below it get label -2
                              (filter_boxes)                            libs/layers/box.py:310-315   */
} detbox_match_cfg;

/* libs/layers/box.py:12-69 matching_box and the labelling half of sample_rois (:193-221), batched:
 * boxes [R, box_stride>=4] f32, image_inds [R] i64, gt [sum G, 5] f32 (x1,y1,x2,y2,cls) concatenated
 * image-major with gt_offsets [B+1] i32.  IoU in float64 with bbox.pyx:15-55's arithmetic, first
 * maximum wins.  Outputs in INPUT order (any may be NULL): labels [R] i64, match_inds [R] i32 (index
 * inside the image's GT list, -1 = unmatched), match_boxes [R,4] f32 (16-byte aligned).  Rows whose
 * image index is outside [0,B) get label -2 / match -1 (the reference drops them). */
int detbox_match_boxes(const float *boxes_dev, int box_stride, const int64_t *image_inds_dev, long R,
                       const float *gt_dev, const int32_t *gt_offsets_dev, int B, const detbox_match_cfg *cfg,
                       int64_t *labels_dev, int32_t *match_inds_dev, float *match_boxes_dev, void *stream);

/* The reference's FFI `roi_target_forward_cuda` (libs/layers/roi_target/src/roi_target_cuda.h,
 * kernel roi_target_kernel.cu:49-90) with raw pointers: float32 IoU (inter/area(roi) for GT of class
 * <= 0), last maximum wins, labels i64, deltas / bbwght [R,4] zero-filled here then written for
 * matched rows.  bg_overlap is accepted and unused, as upstream. */
int detbox_roi_target(const float *rois_dev, int roi_stride, const int64_t *roi_batch_inds_dev, long R,
                      const float *gt_dev /*[G,5]*/, const int64_t *gt_batch_inds_dev, long G,
                      float bg_overlap, float fg_overlap,
                      int64_t *labels_dev, float *deltas_dev, float *bbwght_dev, void *stream);

/* sample_rois, ordered compaction (libs/layers/box.py:236-254): per image the proposals with
 * label >= 0 (rows must be image-major) followed by that image's extra rows (the jittered GT boxes,
 * extra_offsets [B+1]); then the ascending positions of foreground (label > 0) and background
 * (label == 0) rows.  Outputs sized R + E; counts = (rows, n_fg, n_bg).  B <= 2048. */
int detbox_sample_compact(const float *boxes_dev, int box_stride, const int64_t *image_inds_dev,
                          const int64_t *labels_dev, long R,
                          const float *extra_boxes_dev, const int64_t *extra_labels_dev,
                          const int32_t *extra_offsets_dev, int B,
                          float *out_boxes_dev, int64_t *out_labels_dev, int64_t *out_image_inds_dev,
                          int32_t *fg_pos_dev, int32_t *bg_pos_dev, int32_t *counts_dev, void *stream);
/* sample_rois, final gather (libs/layers/box.py:271-280): rows fg_pos[fg_choice[..]] then
 * bg_pos[bg_choice[..]]; the choices are the caller's random draws (np.random.choice positions). */
int detbox_sample_gather(const float *boxes_dev, const int64_t *labels_dev, const int64_t *image_inds_dev,
                         const int32_t *fg_pos_dev, const int32_t *bg_pos_dev,
                         const int64_t *fg_choice_dev, int n_fg, const int64_t *bg_choice_dev, int n_bg,
                         float *out_boxes_dev, int64_t *out_labels_dev, int64_t *out_image_inds_dev, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Stage-one epilogue on the device: score thresholding without a host round trip
 * and the `to_Dets*` conversions of libs/nets/model.py:357-469.
 * ------------------------------------------------------------------------------------------- */
/* libs/nets/model.py:326-333 (_decoding_and_thresholding_stage1): rows with score >= threshold stay
 * in their image's segment; if NO row passes, only the single best row does (lowest index on ties).
 * All other rows (and rows with an image index outside [0, n_images)) go to the dummy segment
 * n_images.  Writes the segment ids twice: int32 for detbox_nms, int64 for detbox_to_dets. */
int detbox_threshold_segments(const float *scores_dev, const int64_t *img_ids_dev, long n, int n_images,
                              float score_threshold, int32_t *seg32_dev, int64_t *seg64_dev, void *stream);
/* to_Dets (:358-378), to_Dets_sigmoid (:381-402), to_Dets2 (:405-436), to_Dets2_sigmoid (:439-469):
 * rows keep[0..m) (or 0..m) of boxes [.,4] / probs [.,C] / img_ids, image-major, m = *m_dev when given
 * (e.g. detbox_nms's n_keep[n_seg]) else m_max.  multi_label = 0: one detection per row, arg-max class;
 * multi_label = 1: every (row, class) with prob > score_threshold, in row-major order, and the arg-max
 * rule for an image where no pair passes.  softmax: column 0 is background; C == 2 (softmax) or C == 1
 * (sigmoid): one detection per row, class 1, no threshold.  Rows whose image id is outside
 * [0, n_images) are dropped.  dets [cap,6] float64 = (x1,y1,x2,y2,prob,class) (NumPy's result type of
 * the reference's concatenate), det_offsets int32 [n_images+1]; det_offsets[n_images] > cap means the
 * output was truncated.  n_images <= 2048. */
int detbox_to_dets(const float *boxes_dev, const float *probs_dev, int C, const int64_t *img_ids_dev,
                   const int64_t *keep_dev, const int32_t *m_dev, long m_max, int n_images,
                   int activation, int multi_label, float score_threshold,
                   double *dets_dev, long cap, int32_t *det_offsets_dev, void *stream);

/* Fixed-size detection buffers for the optional final all-gather of the per-image sharded mode (the
 * reference gathers DataParallel outputs in one process, libs/nets/data_parallel.py:19-91, and truncates to max_det_num,
 * libs/postprocessings/single_shot.py:59): from detbox_nms's keep / n_keep (n_seg images, segment-major, best first) the
 * first max_dets kept rows of every image as out [n_seg, max_dets, 6] float32 = (x1,y1,x2,y2,score,class), zero padded,
 * and counts int32 [n_seg].  classes_dev may be NULL (class 1).  boxes [.,4] 16-byte aligned.  No host round trip. */
int detbox_pack_detections(const float *boxes_dev, const float *scores_dev, const float *classes_dev,
                           const int64_t *keep_dev, const int32_t *n_keep_dev, int n_seg, int max_dets,
                           float *out_dev, int32_t *counts_dev, void *stream);

/* libs/postprocessings/single_shot.py:401-415 dets_cluster — score-weighted box voting over the clusters
 * detbox_nms_variant(MULTIANCHOR) returns: out row k (float32 [5]) = sum over the members i of cluster
 * keep[k] (clusters[i] == keep[k], ascending i) of dets[i,:4] * score_i / sum(score), score = the members'
 * maximum; a cluster of one copies its row.  Sums follow NumPy's float32 order (pairwise score sum,
 * sequential row sum).  n_keep = *n_keep_dev when given (clamped to n_keep_max), else n_keep_max.
 * dets [n,5] float32 contiguous, keep / clusters int64 (original indices), out [n_keep_max,5]. */
int detbox_dets_cluster(const float *dets_dev, long n, const int64_t *keep_dev, const int32_t *n_keep_dev,
                        long n_keep_max, const int64_t *clusters_dev, float *out_dev, void *stream);

/* ---------------------------------------------------------------------------------------------
 * HOST-buffer entry points (copies inside; synchronous; scratch cached per device).
 * ------------------------------------------------------------------------------------------- */
/* Exactly the reference's FFI symbol `_nms` (libs/nms/gpu_nms.hpp:1-2, bound by gpu_nms.pyx:14):
 * boxes_host [n, boxes_dim>=5] ALREADY SORTED by descending score; keep_out receives indices
 * into that sorted array; comparator '>' with a float threshold. */
void _nms(int *keep_out, int *num_out, const float *boxes_host, int boxes_num, int boxes_dim,
          float nms_overlap_thresh, int device_id);
/* Unsorted dets [n,5] in, kept ORIGINAL indices out (what gpu_nms.pyx / cpu_nms.pyx return). */
int detbox_nms_host(const float *dets_host, int n, double thresh, int cmp_mode,
                    int *keep_out_host, int *num_out_host, int device_id);
/* libs/layers/anchor.py:17 encode() for one image with host arrays in and out. */
int detbox_anchor_target_host(const float *anchors_host, long A, const float *gt_host, int G,
                              const detbox_target_cfg *cfg, float *labels_host, float *label_weights_host,
                              float *bbox_targets_host, float *bbox_inside_weights_host, int device_id);

/* Device-side address of PINNED host memory (cudaHostAlloc / cudaHostRegister), for passing a host array where a
 * *_dev pointer is expected and the kernel reads only a few rows of it: detbox_topk_decode's deltas_dev (the gather
 * reads top_n rows per image — the reference indexes rpn_box[order] on the host, libs/nets/model.py:314-318). */
int detbox_host_device_pointer(const void *host_ptr, void **dev_ptr_out);

#ifdef __cplusplus
}
#endif
#endif /* DETBOX_H */
