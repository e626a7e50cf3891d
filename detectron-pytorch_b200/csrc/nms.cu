// nms.cu — greedy non-maximum suppression entirely on the device.
// Replaces: libs/nms/gpu_nms.pyx + nms_kernel.cu:46-155 (bitmask kernel + HOST scan + per-call
// cudaMalloc/Memcpy), libs/nms/cpu_nms.pyx:17-68, libs/layers/box.py:148-173 + layers/nms/src/*,
// and — through segments — the per-image Python loop of libs/nets/model.py:343-355.
//
// Pipeline (one stream, no host round trip, no allocation):
//   1. keys = ~ordered(score), stable radix sort (sort.cu) -> score-descending order, ties by index;
//      with segments one more stable pass on the segment id -> segment-major order.
//   2. gather boxes into sorted float4 rows (+ segment id).
//   3. bitmask: 64x64 tiles of the UPPER triangle only (the reference computes the full square,
//      nms_kernel.cu:51); column boxes and their areas staged in shared memory; one thread
//      produces one 64-bit word.  IoU arithmetic is the reference's float32 sequence
//      (cpu_nms.pyx:57-64 == nms_kernel.cu:24-32): (area_i + area_j) - inter, IEEE division,
//      --fmad=false.  The division is skipped when inter == 0 and thr > 0 (the quotient is 0
//      or NaN, neither can suppress) — most pairs in the sparse regime.
//   4. greedy scan, one block per segment: walk the 64-row tiles in order; warp 0 resolves the
//      diagonal word chain with shuffles (64 dependent steps of register work), then all threads
//      OR the mask rows of the newly kept boxes into the shared-memory `removed` bitset.  Exact
//      greedy order => exact keep indices.
//   5. ordered compaction of the kept bits into original indices + per-segment counts.
#include <mutex>
#include <cooperative_groups.h>
#include "sort.cuh"

namespace cg = cooperative_groups;

using namespace detbox;
typedef unsigned long long u64;

namespace {

struct NmsCmp { int mode; float thr_f; double thr_d; bool nonpos; };

__device__ __forceinline__ bool suppresses(const NmsCmp &c, float ovr) {
    if (c.mode == DETBOX_NMS_GE_F64) return (double)ovr >= c.thr_d;      // cpu_nms.pyx:65
    if (c.mode == DETBOX_NMS_GE_F32) return ovr >= c.thr_f;             // nms_cpu.c:76
    return ovr > c.thr_f;                                               // nms_kernel.cu:83
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

// segment sizes -> offsets in the (segment-major) processing order
__global__ void __launch_bounds__(256)
nms_seg_count_kernel(const int32_t *__restrict__ seg_ids, long n, int n_seg, int32_t *__restrict__ seg_off)
{
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const int s = seg_ids[i];
        if (s >= 0 && s < n_seg) atomicAdd(&seg_off[s + 1], 1);
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
// j/64 iff j would suppress i.  Blocks own (row tile, group of MASK_CG column tiles); groups right
// of the diagonal, or left of the row tile's segment, exit at once.
constexpr int MASK_CT = 4;     // column tiles staged per iteration (one per 64-thread group)
constexpr int MASK_CG = 32;    // column tiles per block

__global__ void __launch_bounds__(64 * MASK_CT)
nms_mask_kernel(const float4 *__restrict__ boxes, const int32_t *__restrict__ seg, const int32_t *__restrict__ seg_off,
                int n, int ncol, NmsCmp cmp, u64 *__restrict__ mask)
{
    const int rt = blockIdx.y;
    const int ct_lo = blockIdx.x * MASK_CG;
    if (ct_lo > rt) return;                                    // strictly right of the diagonal
    const int row0 = rt * 64, row_last = min(n, row0 + 64) - 1;
    int first_ct = 0;                                          // first column tile any row of this tile can see
    if (seg) first_ct = seg_off[seg[row0]] >> 6;
    if (ct_lo + MASK_CG <= first_ct) return;
    __shared__ float4 cb[64 * MASK_CT];
    __shared__ float carea[64 * MASK_CT];
    __shared__ int cseg[64 * MASK_CT];
    const int tid = threadIdx.x, sub = tid >> 6, rl = tid & 63;
    const int r = row0 + rl;
    float4 a = make_float4(0, 0, 0, 0); float aarea = 0.0f; int aseg = 0;
    if (r < n) { a = boxes[r]; aarea = (a.z - a.x + 1) * (a.w - a.y + 1); aseg = seg ? seg[r] : 0; }
    const int ct_hi = min(rt, ct_lo + MASK_CG - 1);
    for (int c0 = max(ct_lo, first_ct & ~(MASK_CT - 1)); c0 <= ct_hi; c0 += MASK_CT) {
        __syncthreads();
        {
            const int c = c0 * 64 + tid;
            if (c <= row_last) {
                const float4 b = boxes[c];
                cb[tid] = b;
                carea[tid] = (b.z - b.x + 1) * (b.w - b.y + 1);    // cpu_nms.pyx:24
                cseg[tid] = seg ? seg[c] : 0;
            }
        }
        __syncthreads();
        const int ct = c0 + sub;
        if (r >= n || ct > rt || ct < ct_lo) continue;
        const int cbase = sub * 64;
        const int cnt = (ct == rt) ? rl : 64;                      // diagonal tile: only j < i
        u64 word = 0;
        for (int i = 0; i < cnt; ++i) {
            const float4 b = cb[cbase + i];
            const float xx1 = fmaxf(a.x, b.x), yy1 = fmaxf(a.y, b.y);
            const float xx2 = fminf(a.z, b.z), yy2 = fminf(a.w, b.w);
            const float w = fmaxf(0.0f, xx2 - xx1 + 1), h = fmaxf(0.0f, yy2 - yy1 + 1);
            const float inter = w * h;
            if (inter > 0.0f || cmp.nonpos) {
                // the earlier box j plays cpu_nms.pyx's `i`, this row its `j`: (area_i + area_j) - inter
                const float ovr = inter / (carea[cbase + i] + aarea - inter);
                if (suppresses(cmp, ovr) && aseg == cseg[cbase + i]) word |= 1ull << i;
            }
        }
        mask[(size_t)r * ncol + ct] = word;
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

__global__ void __launch_bounds__(1024)
nms_compact_kernel(const u64 *__restrict__ kept, int n, int ncol, const uint32_t *__restrict__ order,
               const int32_t *__restrict__ seg_sorted, int n_seg, int64_t *__restrict__ keep, int32_t *__restrict__ n_keep)
{
    __shared__ int warp_sum[32];
    __shared__ int running;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) running = 0;
    __syncthreads();
    for (int base = 0; base < ncol; base += 1024) {
        const int wi = base + tid;
        u64 word = wi < ncol ? kept[wi] : 0ull;
        const int c = __popcll(word);
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        if (lane == 31) warp_sum[w] = incl;
        __syncthreads();
        if (w == 0) {
            const int v = warp_sum[lane]; int iv = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, iv, o); if (lane >= o) iv += t; }
            warp_sum[lane] = iv - v;
        }
        __syncthreads();
        int pos = running + warp_sum[w] + incl - c;
        while (word) {
            const int bit = __ffsll((long long)word) - 1;
            word &= word - 1;
            const int sp = wi * 64 + bit;
            keep[pos++] = (int64_t)(order ? order[sp] : (uint32_t)sp);
            if (n_seg > 1) atomicAdd(&n_keep[seg_sorted[sp]], 1);
        }
        __syncthreads();
        if (tid == 1023) running = pos;          // last thread's end == running + block total
        __syncthreads();
    }
    if (tid == 0) { n_keep[n_seg] = running; if (n_seg == 1) n_keep[0] = running; }
}

struct NmsBuffers {
    uint32_t *ka, *kb, *va, *vb, *counts;
    float4 *sorted; int32_t *seg_sorted, *seg_off; u64 *mask, *kept, *supp; unsigned *cnt;
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
    b.seg_off = cv.take<int32_t>(n_seg + 2);
    b.kept = cv.take<u64>(2 * ncol + 2);           // kept | supp | round counters, cleared together
    b.supp = b.kept + ncol;
    b.cnt = reinterpret_cast<unsigned *>(b.supp + ncol);
    b.mask = cv.take<u64>((size_t)n * ncol);
    b.total = cv.off;
    return b;
}

NmsCmp make_cmp(double thresh, int mode) {
    NmsCmp c; c.mode = mode; c.thr_d = thresh; c.thr_f = (float)thresh;
    c.nonpos = (mode == DETBOX_NMS_GE_F64) ? !(thresh > 0.0) : !(c.thr_f > 0.0f);
    return c;
}

// mask + scan + compact over boxes that are already gathered in processing order
int nms_sorted(const NmsBuffers &b, const uint32_t *order, long n, int n_seg,
               NmsCmp cmp, int64_t *keep, int32_t *n_keep, cudaStream_t st)
{
    const int32_t *seg_off = b.seg_off;
    const int ncol = (int)((n + 63) / 64);
    const int32_t *seg_sorted = n_seg > 1 ? b.seg_sorted : nullptr;
    DETBOX_CUDA_TRY(cudaMemsetAsync(b.kept, 0, (size_t)(2 * ncol + 2) * sizeof(u64), st));
    DETBOX_CUDA_TRY(cudaMemsetAsync(n_keep, 0, (size_t)(n_seg + 1) * sizeof(int32_t), st));
    dim3 grid((unsigned)((ncol + MASK_CG - 1) / MASK_CG), (unsigned)ncol);
    DETBOX_TRACE_BEGIN(DETBOX_K_NMS_MASK, st);
    nms_mask_kernel<<<grid, 64 * MASK_CT, 0, st>>>(b.sorted, seg_sorted, seg_sorted ? seg_off : nullptr, (int)n, ncol, cmp, b.mask);
    DETBOX_TRACE_END(DETBOX_K_NMS_MASK, st);
    DETBOX_LAUNCH_CHECK();
    // cooperative launch: every block must be resident (grid.sync between rounds)
    static thread_local int coop_blocks[64] = {0};
    int dev = 0;
    DETBOX_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return DETBOX_ERR_UNSUPPORTED;
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
    nms_compact_kernel<<<1, 1024, 0, st>>>(b.kept, (int)n, ncol, order, seg_sorted, n_seg, keep, n_keep);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

}  // namespace

extern "C" size_t detbox_nms_workspace_bytes(long n, int n_seg)
{
    if (n <= 0) return 256;
    return carve_nms(nullptr, n, n_seg > 1 ? n_seg : 1).total;
}

extern "C" int detbox_nms(const float *boxes, int box_stride, const float *scores, long n,
                          const int32_t *seg_ids, int n_seg, double thresh, int cmp_mode,
                          int64_t *keep, int32_t *n_keep, void *ws, size_t ws_bytes, void *stream)
{
    const bool presorted = (cmp_mode & DETBOX_NMS_PRESORTED) != 0;
    cmp_mode &= ~DETBOX_NMS_PRESORTED;
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
    if (presorted) {                              // rows are in processing order: no sort, no gather
        if (n_seg > 1) {
            DETBOX_CUDA_TRY(cudaMemsetAsync(b.seg_off, 0, (size_t)(n_seg + 2) * sizeof(int32_t), st));
            nms_seg_count_kernel<<<(unsigned)gb0, 256, 0, st>>>(seg_ids, n, n_seg, b.seg_off);
            DETBOX_LAUNCH_CHECK();
            nms_seg_scan_kernel<<<1, 32, 0, st>>>(b.seg_off, n_seg);
            DETBOX_LAUNCH_CHECK();
        }
        b.sorted = const_cast<float4 *>(reinterpret_cast<const float4 *>(boxes));
        b.seg_sorted = const_cast<int32_t *>(seg_ids);
        return nms_sorted(b, nullptr, n, n_seg, make_cmp(thresh, cmp_mode), keep, n_keep, st);
    }
    int rc = make_score_keys(scores, n, b.ka, b.va, st);
    if (rc != DETBOX_OK) return rc;
    rc = radix_sort_pairs(b.ka, b.va, b.kb, b.vb, n, b.counts, st);
    if (rc != DETBOX_OK) return rc;
    uint32_t *order = b.va;
    long gb = (n + 255) / 256; if (gb > kNumSMs * 8) gb = kNumSMs * 8;
    if (n_seg > 1) {                              // stable grouping by segment: 1 pass per 8 bits of id
        DETBOX_CUDA_TRY(cudaMemsetAsync(b.seg_off, 0, (size_t)(n_seg + 2) * sizeof(int32_t), st));
        nms_seg_count_kernel<<<(unsigned)gb, 256, 0, st>>>(seg_ids, n, n_seg, b.seg_off);
        DETBOX_LAUNCH_CHECK();
        nms_seg_scan_kernel<<<1, 32, 0, st>>>(b.seg_off, n_seg);
        DETBOX_LAUNCH_CHECK();
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
    return nms_sorted(b, order, n, n_seg, make_cmp(thresh, cmp_mode), keep, n_keep, st);
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
