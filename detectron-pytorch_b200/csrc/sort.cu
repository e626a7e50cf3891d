// sort.cu — stable LSD radix sort (8-bit digits) of (uint32 key, uint32 value) pairs.
//
// One pass = three launches:
//   count   every block histograms the digits of its tile in shared memory and writes
//           counts[digit][block] (digit-major, so one flat exclusive scan yields every
//           (digit, block) destination offset);
//   scan    one 1024-thread block scans the <= 128 Ki counters;
//   scatter every block re-reads its tile in 2048-element sub-tiles, ranks the elements stably
//           (warp-striped order; __match_any_sync groups equal digits inside a warp step, a
//           per-warp digit counter carries the rank across the 8 steps, a per-digit scan over
//           the 8 warps orders the warps) and writes them to their final positions.
// Tiles grow with n so that the block count stays <= 512 (bounded scan cost); work per pass is
// 20 B/element of HBM traffic (4 B counted + 8 B read + 8 B written).
//
// The digit can come from the key, from value / divisor (image id of a flat anchor index) or from
// the segment a value falls in — that is how "image-major, score-descending inside an image"
// orderings are produced: sort by score, then one more stable pass on the image id.
#include "sort.cuh"

namespace detbox {
namespace {

__global__ void __launch_bounds__(SORT_THREADS)
count_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, long n, long tile,
             DigitSrc d, uint32_t *__restrict__ counts)
{
    __shared__ uint32_t hist[256];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const long begin = (long)blockIdx.x * tile;
    const long end = min(n, begin + tile);
    for (long i = begin + threadIdx.x; i < end; i += SORT_THREADS) {
        const uint32_t k = keys[i];
        const uint32_t v = d.mode == 0 ? 0u : vals[i];
        atomicAdd(&hist[digit_of(d, k, v)], 1u);
    }
    __syncthreads();
    counts[(size_t)threadIdx.x * gridDim.x + blockIdx.x] = hist[threadIdx.x];
}

__global__ void __launch_bounds__(1024)
scan_kernel(uint32_t *__restrict__ counts, int m)
{
    __shared__ uint32_t warp_sum[32];
    const int per = (m + 1023) / 1024;
    const int begin = min(m, (int)threadIdx.x * per), end = min(m, begin + per);
    uint32_t s = 0;
    for (int i = begin; i < end; ++i) s += counts[i];
    // block-wide exclusive scan of s
    uint32_t incl = s;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sum[w] = incl;
    __syncthreads();
    if (w == 0) {
        uint32_t v = warp_sum[lane], iv = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, iv, o);
            if (lane >= o) iv += t;
        }
        warp_sum[lane] = iv - v;
    }
    __syncthreads();
    uint32_t run = warp_sum[w] + incl - s;
    for (int i = begin; i < end; ++i) { const uint32_t c = counts[i]; counts[i] = run; run += c; }
}

__global__ void __launch_bounds__(SORT_THREADS)
scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
               uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, long n, long tile,
               DigitSrc d, const uint32_t *__restrict__ offsets)
{
    __shared__ uint32_t wcnt[SORT_THREADS / 32][256];
    __shared__ uint32_t goff[256];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    goff[tid] = offsets[(size_t)tid * gridDim.x + blockIdx.x];
    const long begin = (long)blockIdx.x * tile;
    const long end = min(n, begin + tile);
    for (long sub = begin; sub < end; sub += SORT_SUB) {
#pragma unroll
        for (int i = 0; i < SORT_THREADS / 32; ++i) wcnt[i][tid] = 0;
        __syncthreads();
        uint32_t key[SORT_ITEMS], val[SORT_ITEMS], rank[SORT_ITEMS];
        int dg[SORT_ITEMS];
        const long wbase = sub + (long)w * 32 * SORT_ITEMS;
#pragma unroll
        for (int j = 0; j < SORT_ITEMS; ++j) {
            const long idx = wbase + j * 32 + lane;
            const bool ok = idx < end;
            key[j] = ok ? keys_in[idx] : 0xffffffffu;
            val[j] = ok ? vals_in[idx] : 0u;
            dg[j] = ok ? (int)digit_of(d, key[j], val[j]) : 256;            // 256: padding, ranked nowhere
            const unsigned peers = __match_any_sync(0xffffffffu, dg[j]);
            uint32_t prior = 0;
            if (ok) prior = wcnt[w][dg[j]];
            __syncwarp();
            if (ok && lane == (__ffs(peers) - 1)) wcnt[w][dg[j]] = prior + __popc(peers);
            __syncwarp();
            rank[j] = prior + __popc(peers & ((1u << lane) - 1u));
        }
        __syncthreads();
        {   // thread == digit: order the 8 warps, advance the block's running offset
            uint32_t run = goff[tid];
#pragma unroll
            for (int i = 0; i < SORT_THREADS / 32; ++i) { const uint32_t c = wcnt[i][tid]; wcnt[i][tid] = run; run += c; }
            goff[tid] = run;
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < SORT_ITEMS; ++j) {
            if (dg[j] < 256) {
                const uint32_t pos = wcnt[w][dg[j]] + rank[j];
                keys_out[pos] = key[j];
                vals_out[pos] = val[j];
            }
        }
        __syncthreads();
    }
}

}  // namespace

int radix_pass(const uint32_t *keys_in, const uint32_t *vals_in, uint32_t *keys_out, uint32_t *vals_out,
               long n, DigitSrc d, uint32_t *counts, cudaStream_t st)
{
    if (n <= 0) return DETBOX_OK;
    const SortPlan p = make_sort_plan(n);
    count_kernel<<<p.blocks, SORT_THREADS, 0, st>>>(keys_in, vals_in, n, p.tile, d, counts);
    DETBOX_LAUNCH_CHECK();
    scan_kernel<<<1, 1024, 0, st>>>(counts, 256 * p.blocks);
    DETBOX_LAUNCH_CHECK();
    scatter_kernel<<<p.blocks, SORT_THREADS, 0, st>>>(keys_in, vals_in, keys_out, vals_out, n, p.tile, d, counts);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

int radix_sort_pairs(uint32_t *keys_a, uint32_t *vals_a, uint32_t *keys_b, uint32_t *vals_b,
                     long n, uint32_t *counts, cudaStream_t st)
{
    DigitSrc d{0, 0, 1u, nullptr};
    for (int pass = 0; pass < 4; ++pass) {
        d.shift = 8 * pass;
        const bool fwd = (pass & 1) == 0;
        const int rc = radix_pass(fwd ? keys_a : keys_b, fwd ? vals_a : vals_b,
                                  fwd ? keys_b : keys_a, fwd ? vals_b : vals_a, n, d, counts, st);
        if (rc != DETBOX_OK) return rc;
    }
    return DETBOX_OK;
}

// ------------------------------------------------------------------------------------------------
// public: descending stable argsort of float scores
// ------------------------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256)
score_keys_kernel(const float *__restrict__ scores, long n, uint32_t *__restrict__ keys, uint32_t *__restrict__ vals)
{
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        keys[i] = ~float_to_ordered(scores[i]);      // ascending key == descending score
        vals[i] = (uint32_t)i;
    }
}
}  // namespace

namespace {
// n <= SMALL_SORT_MAX: one block, bitonic network in shared memory over the unique 64-bit composites
// (~ordered(score) << 32 | index) — ascending composite == descending score, ties by lower index.
__global__ void __launch_bounds__(1024)
small_sort_kernel(const float *__restrict__ scores, int n, int N /*pow2 >= n*/, uint32_t *__restrict__ order)
{
    __shared__ unsigned long long v[SMALL_SORT_MAX];
    const int tid = threadIdx.x;
    for (int i = tid; i < N; i += blockDim.x)
        v[i] = i < n ? ((unsigned long long)(~float_to_ordered(scores[i])) << 32) | (unsigned)i : ~0ull;
    __syncthreads();
    for (int size = 2; size <= N; size <<= 1) {
        for (int j = size >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < (N >> 1); t += blockDim.x) {
                const int i = 2 * t - (t & (j - 1)), q = i + j;
                const unsigned long long a = v[i], b = v[q];
                const bool up = (i & size) == 0;
                if ((a > b) == up) { v[i] = b; v[q] = a; }
            }
            __syncthreads();
        }
    }
    for (int i = tid; i < n; i += blockDim.x) order[i] = (uint32_t)v[i];
}
}  // namespace

int small_sort_scores_desc(const float *scores, long n, uint32_t *order, cudaStream_t st)
{
    if (n <= 0) return DETBOX_OK;
    if (n > SMALL_SORT_MAX) return DETBOX_ERR_UNSUPPORTED;
    int N = 2; while (N < n) N <<= 1;
    const int threads = N / 2 < 32 ? 32 : (N / 2 > 1024 ? 1024 : N / 2);
    small_sort_kernel<<<1, threads, 0, st>>>(scores, (int)n, N, order);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

int make_score_keys(const float *scores, long n, uint32_t *keys, uint32_t *vals, cudaStream_t st)
{
    if (n <= 0) return DETBOX_OK;
    long b = (n + 255) / 256; if (b > kNumSMs * 8) b = kNumSMs * 8;
    score_keys_kernel<<<(unsigned)b, 256, 0, st>>>(scores, n, keys, vals);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}

}  // namespace detbox

using namespace detbox;

extern "C" size_t detbox_sort_desc_workspace_bytes(long n)
{
    if (n <= 0) return 256;
    return 3 * align_up((size_t)n * 4) + sort_counts_bytes(n);
}

extern "C" int detbox_sort_desc(const float *scores, long n, int32_t *order, void *ws, size_t ws_bytes, void *stream)
{
    if (n < 0) return DETBOX_ERR_INVALID;
    if (n == 0) return DETBOX_OK;
    if (n >= (1L << 31)) return DETBOX_ERR_UNSUPPORTED;
    if (!scores || !order) return DETBOX_ERR_INVALID;
    if (!ws || ws_bytes < detbox_sort_desc_workspace_bytes(n)) return DETBOX_ERR_WORKSPACE;
    Carver cv(ws);
    uint32_t *ka = cv.take<uint32_t>(n), *kb = cv.take<uint32_t>(n), *vb = cv.take<uint32_t>(n);
    uint32_t *counts = cv.take<uint32_t>(sort_counts_bytes(n) / 4);
    cudaStream_t st = as_stream(stream);
    if (n <= SMALL_SORT_MAX) return small_sort_scores_desc(scores, n, (uint32_t *)order, st);
    int rc = make_score_keys(scores, n, ka, (uint32_t *)order, st);
    if (rc != DETBOX_OK) return rc;
    return radix_sort_pairs(ka, (uint32_t *)order, kb, vb, n, counts, st);
}
