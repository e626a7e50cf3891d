// anchor_target.cu — fused anchor-target assignment (reference libs/layers/anchor.py:17-145,
// 'simple' sampling; batch wrapper libs/layers/data_layer.py:228-239).
//
// The reference materialises the [A,G] float64 overlap matrix (25-207 MB per image) and makes
// three NumPy passes over it.  Here nothing of size A*G ever exists:
//
//   pass 1  column maxima: per-GT max IoU over all anchors, as order-preserving u64 bit patterns
//           (IoU >= 0) accumulated with shared-memory atomicMax, one global atomicMax per
//           (block, GT) that has a non-zero value.
//   pass 2  per anchor: recompute its row (cheap, see culling), first-max argmax, thresholds,
//           per-GT promotion (ov == colmax, all ties, incl. all-zero columns), clobber, ignore
//           regions, float32 regression targets; fg/bg counts by block reduction.
//   pass 3  label_weights from the per-image counts (needs the global fg/bg totals).
//
// HBM bytes per anchor: 36 written in pass 2 (labels 4 + targets 16 + inside weights 16),
// 4 written + 4 read in pass 3; the 16-byte anchor rows are shared by the whole batch and stay
// L2-resident.  No A*G traffic.
//
// Culling keeps the float64 work off the critical path: GT boxes are staged in shared memory
// (float64 copies + areas); a block owns 1024 consecutive anchors, discards every GT that cannot
// touch the block's bounding box, and each warp discards the survivors that cannot touch the
// bounding box of its own 32 anchors (a warp-uniform branch).  Exact float64 IoU (reference
// bbox.pyx:33-54, same operation order, --fmad=false) is evaluated only for pairs that may
// intersect, so labels are bit-exact while ~99% of pairs cost one predicate.
//
// Monotonicity makes the culling exact: min/max are exact, and a rounded (x - y + 1) is a
// monotone function of its operands, so iw(anchor, gt) <= iw(bbox, gt) also after rounding.
#include "box_math.cuh"

using namespace detbox;

namespace {

constexpr int AT_THREADS = 256;
constexpr int AT_ITEMS = 4;                         // anchors per thread
constexpr int AT_TILE = AT_THREADS * AT_ITEMS;      // anchors per block
constexpr int GCH = 128;                            // GT boxes staged per chunk

struct GtChunk {
    double x1[GCH], y1[GCH], x2[GCH], y2[GCH], area[GCH];
    unsigned long long colmax[GCH];
    float cls[GCH];
    int cand[GCH];
    int warp_cnt[GCH / 32];
    int n_cand;
    int zero_noncand;
    unsigned bb[4];                                  // block bbox as ordered uints
};

__device__ __forceinline__ unsigned long long iou_bits(double v) {
    return (unsigned long long)__double_as_longlong(v);
}

struct TargetParams {
    const float4 *anchors; long A;
    const float *gt; const int32_t *gt_off;
    unsigned long long *colmax; int32_t *counts;
    double bg, fg, ign;
    int clobber, enc;
    float4 bbw;
    float *labels; float4 *targets; float4 *inside;
};

template <int PASS>
__global__ void __launch_bounds__(AT_THREADS)
anchor_target_kernel(const TargetParams p)
{
    __shared__ GtChunk sm;
    const int tid = threadIdx.x, lane = tid & 31;
    const int b = blockIdx.y;
    const int g0 = p.gt_off[b];
    const int G = p.gt_off[b + 1] - g0;
    const long tile0 = (long)blockIdx.x * AT_TILE;
    const float INF = __int_as_float(0x7f800000);

    float4 anc[AT_ITEMS];
    bool valid[AT_ITEMS];
#pragma unroll
    for (int i = 0; i < AT_ITEMS; ++i) {
        const long a = tile0 + i * AT_THREADS + tid;
        valid[i] = a < p.A;
        anc[i] = valid[i] ? p.anchors[a] : make_float4(INF, INF, -INF, -INF);
    }

    if (G == 0) {                                    // anchor.py:89-91,133: all background, zero targets
        if (PASS == 2) {
            int nbg = 0;
#pragma unroll
            for (int i = 0; i < AT_ITEMS; ++i) {
                if (!valid[i]) continue;
                const long o = (long)b * p.A + tile0 + i * AT_THREADS + tid;
                p.labels[o] = 0.0f;
                p.targets[o] = make_float4(0, 0, 0, 0);
                p.inside[o] = make_float4(0, 0, 0, 0);
                ++nbg;
            }
            nbg = __reduce_add_sync(0xffffffffu, nbg);
            if (lane == 0 && nbg) atomicAdd(&p.counts[b * 2 + 1], nbg);
        }
        return;
    }

    // ---- bounding boxes: per warp and sub-tile (registers), per block (shared) ---------------
    unsigned wbb[AT_ITEMS][4];
    if (tid < 4) sm.bb[tid] = (tid < 2) ? 0xffffffffu : 0u;
    if (tid == 0) sm.zero_noncand = 0;
    __syncthreads();
    {
        unsigned m0 = 0xffffffffu, m1 = 0xffffffffu, m2 = 0u, m3 = 0u;
#pragma unroll
        for (int i = 0; i < AT_ITEMS; ++i) {
            wbb[i][0] = __reduce_min_sync(0xffffffffu, float_to_ordered(anc[i].x));
            wbb[i][1] = __reduce_min_sync(0xffffffffu, float_to_ordered(anc[i].y));
            wbb[i][2] = __reduce_max_sync(0xffffffffu, float_to_ordered(anc[i].z));
            wbb[i][3] = __reduce_max_sync(0xffffffffu, float_to_ordered(anc[i].w));
            m0 = min(m0, wbb[i][0]); m1 = min(m1, wbb[i][1]);
            m2 = max(m2, wbb[i][2]); m3 = max(m3, wbb[i][3]);
        }
        if (lane == 0) {
            atomicMin(&sm.bb[0], m0); atomicMin(&sm.bb[1], m1);
            atomicMax(&sm.bb[2], m2); atomicMax(&sm.bb[3], m3);
        }
    }

    // per-anchor running state
    double best[AT_ITEMS], isum[AT_ITEMS], aarea[AT_ITEMS];
    int best_g[AT_ITEMS];
    bool promoted[AT_ITEMS];
#pragma unroll
    for (int i = 0; i < AT_ITEMS; ++i) {
        best[i] = 0.0; isum[i] = 0.0; best_g[i] = 0; promoted[i] = false;
        aarea[i] = ((double)anc[i].z - (double)anc[i].x + 1) * ((double)anc[i].w - (double)anc[i].y + 1);
    }

    for (int c0 = 0; c0 < G; c0 += GCH) {
        const int gn = min(GCH, G - c0);
        __syncthreads();                             // previous chunk fully consumed; bb complete
        bool is_cand = false;
        if (tid < gn) {
            const float *q = p.gt + (size_t)(g0 + c0 + tid) * 5;
            const double x1 = q[0], y1 = q[1], x2 = q[2], y2 = q[3];
            sm.x1[tid] = x1; sm.y1[tid] = y1; sm.x2[tid] = x2; sm.y2[tid] = y2;
            sm.area[tid] = (x2 - x1 + 1) * (y2 - y1 + 1);                 // bbox.pyx:34-37
            sm.cls[tid] = q[4];
            const unsigned long long cm = (PASS == 1) ? 0ull : p.colmax[g0 + c0 + tid];
            sm.colmax[tid] = cm;
            const double bx1 = ordered_to_float(sm.bb[0]), by1 = ordered_to_float(sm.bb[1]);
            const double bx2 = ordered_to_float(sm.bb[2]), by2 = ordered_to_float(sm.bb[3]);
            const double iw = fmin(bx2, x2) - fmax(bx1, x1) + 1;
            const double ih = fmin(by2, y2) - fmax(by1, y1) + 1;
            is_cand = (iw > 0) && (ih > 0);
            if (PASS == 2 && !is_cand && cm == 0ull) sm.zero_noncand = 1;   // all-zero column: every anchor ties
        }
        if (tid < GCH) {                             // ordered compaction of the candidates
            const unsigned m = __ballot_sync(0xffffffffu, is_cand);
            if (lane == 0) sm.warp_cnt[tid >> 5] = __popc(m);
            __syncwarp();
        }
        __syncthreads();
        if (tid < GCH) {
            const unsigned m = __ballot_sync(0xffffffffu, is_cand);
            int off = 0;
            for (int w = 0; w < (tid >> 5); ++w) off += sm.warp_cnt[w];
            if (is_cand) sm.cand[off + __popc(m & ((1u << lane) - 1))] = tid;
            if (tid == GCH - 1) sm.n_cand = off + __popc(m);
        }
        __syncthreads();
        const int nc = sm.n_cand;

#pragma unroll
        for (int i = 0; i < AT_ITEMS; ++i) {
            const double wx1 = ordered_to_float(wbb[i][0]), wy1 = ordered_to_float(wbb[i][1]);
            const double wx2 = ordered_to_float(wbb[i][2]), wy2 = ordered_to_float(wbb[i][3]);
            const double ax1 = anc[i].x, ay1 = anc[i].y, ax2 = anc[i].z, ay2 = anc[i].w;
            for (int k = 0; k < nc; ++k) {
                const int j = sm.cand[k];
                const double gx1 = sm.x1[j], gy1 = sm.y1[j], gx2 = sm.x2[j], gy2 = sm.y2[j];
                // warp-uniform cull against the warp's bounding box
                const double wiw = fmin(wx2, gx2) - fmax(wx1, gx1) + 1;
                const double wih = fmin(wy2, gy2) - fmax(wy1, gy1) + 1;
                if (!(wiw > 0 && wih > 0)) {
                    if (PASS == 2 && sm.colmax[j] == 0ull) promoted[i] = true;
                    continue;
                }
                double ov = 0.0;
                const double iw = fmin(ax2, gx2) - fmax(ax1, gx1) + 1;          // bbox.pyx:39-42
                if (iw > 0) {
                    const double ih = fmin(ay2, gy2) - fmax(ay1, gy1) + 1;      // bbox.pyx:44-47
                    if (ih > 0) {
                        const double inter = iw * ih;
                        const double ua = aarea[i] + sm.area[j] - inter;        // bbox.pyx:49-53
                        ov = inter / ua;                                        // bbox.pyx:54
                        if (PASS == 2 && sm.cls[j] < 0.0f)
                            isum[i] = isum[i] + inter / aarea[i];              // anchor.py:81-85 (bbox.pyx:93)
                    }
                }
                if (PASS == 1) {
                    if (ov > 0.0) atomicMax(&sm.colmax[j], iou_bits(ov));
                } else {
                    if (ov > best[i]) { best[i] = ov; best_g[i] = c0 + j; }     // first max wins (anchor.py:42)
                    if (valid[i] && iou_bits(ov) == sm.colmax[j]) promoted[i] = true;   // anchor.py:46
                }
            }
        }
        if (PASS == 1) {
            __syncthreads();
            if (tid < gn && sm.colmax[tid] != 0ull) atomicMax(&p.colmax[g0 + c0 + tid], sm.colmax[tid]);
        }
    }
    if (PASS == 1) return;

    // ---- pass 2 epilogue: labels, flags, targets ---------------------------------------------
    const bool zero_nc = sm.zero_noncand != 0;      // written before the last __syncthreads above
    int nfg = 0, nbg = 0;
#pragma unroll
    for (int i = 0; i < AT_ITEMS; ++i) {
        if (!valid[i]) continue;
        const float *q = p.gt + (size_t)(g0 + best_g[i]) * 5;
        const float4 gbox = make_float4(q[0], q[1], q[2], q[3]);
        const float gcls = q[4];
        const double mo = best[i];
        float l = gcls;                                                         // anchor.py:49
        if (mo < p.bg) l = 0.0f;                                                // :50
        if (mo < p.fg && mo >= p.bg) l = -1.0f;                                 // :51-52
        bool flag = mo >= 0.5;                                                  // :53 (hard-coded 0.5)
        if (promoted[i] || zero_nc) l = gcls;                                   // :56 own-argmax class
        if (p.clobber && mo < p.bg) l = 0.0f;                                   // :60-61
        if (l >= 1.0f) flag = true;                                             // :62
        if (isum[i] > p.ign) { l = -1.0f; flag = false; }                       // :86-87
        nfg += (l >= 1.0f); nbg += (l == 0.0f);
        const long o = (long)b * p.A + tile0 + i * AT_THREADS + tid;
        p.labels[o] = l;
        p.targets[o] = encode_box(anc[i], gbox, p.enc);                         // :133-135 every anchor
        p.inside[o] = flag ? p.bbw : make_float4(0, 0, 0, 0);                   // :136-138
    }
    nfg = __reduce_add_sync(0xffffffffu, nfg);
    nbg = __reduce_add_sync(0xffffffffu, nbg);
    if (lane == 0) {
        if (nfg) atomicAdd(&p.counts[b * 2 + 0], nfg);
        if (nbg) atomicAdd(&p.counts[b * 2 + 1], nbg);
    }
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
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < A4; i += (long)gridDim.x * blockDim.x) {
        const float4 v = reinterpret_cast<const float4 *>(l)[i];
        float4 o;
        o.x = v.x >= 1.0f ? 1.0f : (v.x == 0.0f ? wbg : 0.0f);
        o.y = v.y >= 1.0f ? 1.0f : (v.y == 0.0f ? wbg : 0.0f);
        o.z = v.z >= 1.0f ? 1.0f : (v.z == 0.0f ? wbg : 0.0f);
        o.w = v.w >= 1.0f ? 1.0f : (v.w == 0.0f ? wbg : 0.0f);
        reinterpret_cast<float4 *>(w)[i] = o;
    }
    for (long i = A4 * 4 + (long)blockIdx.x * blockDim.x + threadIdx.x; i < A; i += (long)gridDim.x * blockDim.x) {
        const float v = l[i];
        w[i] = v >= 1.0f ? 1.0f : (v == 0.0f ? wbg : 0.0f);
    }
}

}  // namespace

extern "C" size_t detbox_anchor_target_workspace_bytes(int batch, long total_gt)
{
    if (batch < 0 || total_gt < 0) return 0;
    return align_up((size_t)(total_gt > 0 ? total_gt : 1) * sizeof(unsigned long long)) +
           align_up((size_t)(batch > 0 ? batch : 1) * 2 * sizeof(int32_t));
}

extern "C" int detbox_anchor_target(const float *anchors, long A, const float *gt, const int32_t *gt_off,
                                    int batch, long total_gt, const detbox_target_cfg *cfg,
                                    float *labels, float *label_weights, float *bbox_targets,
                                    float *bbox_inside_weights, int32_t *counts_out,
                                    void *ws, size_t ws_bytes, void *stream)
{
    if (A < 0 || batch < 0 || total_gt < 0 || !cfg) return DETBOX_ERR_INVALID;
    if (cfg->box_encoding != DETBOX_ENC_FASTRCNN && cfg->box_encoding != DETBOX_ENC_LINEAR) return DETBOX_ERR_INVALID;
    if (A == 0 || batch == 0) return DETBOX_OK;
    if (!anchors || !gt_off || !labels || !label_weights || !bbox_targets || !bbox_inside_weights) return DETBOX_ERR_INVALID;
    if (total_gt > 0 && !gt) return DETBOX_ERR_INVALID;
    if (!ws || ws_bytes < detbox_anchor_target_workspace_bytes(batch, total_gt)) return DETBOX_ERR_WORKSPACE;
    if (batch > 65535) return DETBOX_ERR_UNSUPPORTED;
    cudaStream_t st = as_stream(stream);
    Carver cv(ws);
    unsigned long long *colmax = cv.take<unsigned long long>(total_gt > 0 ? total_gt : 1);
    int32_t *counts = cv.take<int32_t>((size_t)batch * 2);
    DETBOX_CUDA_TRY(cudaMemsetAsync(ws, 0, cv.off, st));

    TargetParams p;
    p.anchors = (const float4 *)anchors; p.A = A; p.gt = gt; p.gt_off = gt_off;
    p.colmax = colmax; p.counts = counts;
    p.bg = cfg->bg_threshold; p.fg = cfg->fg_threshold; p.ign = cfg->ignore_fraction;
    p.clobber = cfg->clobber_positives; p.enc = cfg->box_encoding;
    p.bbw = make_float4(cfg->bbweights[0], cfg->bbweights[1], cfg->bbweights[2], cfg->bbweights[3]);
    p.labels = labels; p.targets = (float4 *)bbox_targets; p.inside = (float4 *)bbox_inside_weights;

    dim3 grid((unsigned)((A + AT_TILE - 1) / AT_TILE), (unsigned)batch);
    if (total_gt > 0) {
        DETBOX_TRACE_BEGIN(DETBOX_K_TARGET_COLMAX, st);
        anchor_target_kernel<1><<<grid, AT_THREADS, 0, st>>>(p);
        DETBOX_TRACE_END(DETBOX_K_TARGET_COLMAX, st);
        DETBOX_LAUNCH_CHECK();
    }
    DETBOX_TRACE_BEGIN(DETBOX_K_TARGET_ASSIGN, st);
    anchor_target_kernel<2><<<grid, AT_THREADS, 0, st>>>(p);
    DETBOX_TRACE_END(DETBOX_K_TARGET_ASSIGN, st);
    DETBOX_LAUNCH_CHECK();
    long wb = (A / 4 + 255) / 256;
    if (wb < 1) wb = 1;
    if (wb > 4 * kNumSMs) wb = 4 * kNumSMs;
    DETBOX_TRACE_BEGIN(DETBOX_K_TARGET_WEIGHTS, st);
    label_weight_kernel<<<dim3((unsigned)wb, (unsigned)batch), 256, 0, st>>>(labels, A, counts, label_weights, counts_out);
    DETBOX_TRACE_END(DETBOX_K_TARGET_WEIGHTS, st);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}
