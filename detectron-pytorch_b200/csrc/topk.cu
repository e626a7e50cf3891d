// topk.cu — objectness + segmented radix top-k + gather + decode.
// Replaces detection_model._decode_and_choose_top_n_stage1 (reference libs/nets/model.py:308-321,
// objectness :54-61): the reference runs a full torch.sort over batch x levels x anchors to keep
// <= 1% of the keys.  Here:
//
//   1. objectness  one read of the class probabilities (4*C*A bytes per image — the compulsory
//                  traffic of this stage), score -> order-preserving descending uint32 key,
//                  keys written once (4 B/anchor), level-0 histogram (top 11 bits) fused.
//   2. radix select (MSB first, 11+11+10 bits): per segment find the exact key T of the k-th
//                  best anchor, the number of strictly better keys and how many keys equal to T
//                  are needed.  Segment = whole batch (reference behaviour, "global") or one
//                  image ("per_image", the sharded mode).
//   3. compaction  ORDER-PRESERVING (index order) copy of the winners: better keys + the first
//                  `need` ties by index -> deterministic tie rule "lower flat index first".
//   4. sort        stable radix sort of the k winners only (sort.cu); per-image mode adds one
//                  stable pass on the image id.
//   5. gather      deltas / probs / anchors of the winners, decode (box_math.cuh), image ids.
//
// Segments are equal-sized (A anchors per image), so a segment is identified by blockIdx.y.
#include <cstdlib>
#include "box_math.cuh"
#include "sort.cuh"

using namespace detbox;

namespace {

constexpr int NB0 = 2048;            // bins of level 0 / 1 (11 bits); level 2 uses 1024 (10 bits)
constexpr int TK_THREADS = 256;
constexpr int TK_WROWS = 8;          // sigmoid, C > 4: rows per warp and trip
constexpr int TK_WCH = 4;            // ... and 32-float chunks of a row loaded side by side

struct SegState { uint32_t prefix; uint32_t k_rem; uint32_t n_better; uint32_t pad; };

// ---- 1. objectness -> keys + level-0 histogram -------------------------------------------------
template <int ACT>
__global__ void __launch_bounds__(TK_THREADS)
objectness_kernel(const float *__restrict__ prob, int C, long seg_len, uint32_t *__restrict__ keys,
                  uint32_t *__restrict__ hist /*[n_seg][NB0]*/)
{
    __shared__ uint32_t sh[NB0];
    const int tid = threadIdx.x;
    for (int i = tid; i < NB0; i += TK_THREADS) sh[i] = 0;
    __syncthreads();
    const long seg0 = (long)blockIdx.y * seg_len;
    if (C == 2 && (seg_len & 1) == 0 && ((size_t)prob & 15) == 0) {
        // two-class heads (RPN / RetinaNet-on-CityPersons: 134 MB of scores per 64-image step): one 16-byte load covers two
        // rows, and four of them are issued per thread before the first use — the one-row-per-trip loop below keeps 8 useful
        // bytes per thread in flight and ran at 0.46 of the HBM peak
        const float4 *p4 = reinterpret_cast<const float4 *>(prob) + (seg0 >> 1);
        uint2 *k2 = reinterpret_cast<uint2 *>(keys) + (seg0 >> 1);
        const long np = seg_len >> 1, stride = (long)gridDim.x * TK_THREADS;
        for (long q0 = (long)blockIdx.x * TK_THREADS + tid; q0 < np; q0 += 4 * stride) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { const long q = q0 + u * stride; v[u] = q < np ? __ldcs(p4 + q) : make_float4(0, 0, 0, 0); }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long q = q0 + u * stride;
                if (q >= np) continue;
                float s0, s1;
                if (ACT == DETBOX_ACT_SOFTMAX) { s0 = 1.0f - v[u].x; s1 = 1.0f - v[u].z; }        // model.py:57
                else { s0 = fmaxf(v[u].x, v[u].y); s1 = fmaxf(v[u].z, v[u].w); }                  // model.py:59
                const uint32_t ka = ~float_to_ordered(s0), kb = ~float_to_ordered(s1);
                k2[q] = make_uint2(ka, kb);
                atomicAdd(&sh[ka >> 21], 1u);
                atomicAdd(&sh[kb >> 21], 1u);
            }
        }
    } else if (ACT == DETBOX_ACT_SOFTMAX || C <= 4) {
        for (long r = (long)blockIdx.x * TK_THREADS + tid; r < seg_len; r += (long)gridDim.x * TK_THREADS) {
            const float *p = prob + (size_t)(seg0 + r) * C;
            float s;
            if (ACT == DETBOX_ACT_SOFTMAX) s = 1.0f - p[0];                    // model.py:57
            else { s = p[0]; for (int c = 1; c < C; ++c) s = fmaxf(s, p[c]); } // model.py:59
            const uint32_t k = ~float_to_ordered(s);
            keys[seg0 + r] = k;
            atomicAdd(&sh[k >> 21], 1u);
        }
    } else {
        // max over C classes: one warp per row, TK_WROWS rows per warp and trip.  A row is C contiguous floats, so the
        // lanes' 4-byte loads are full 128-byte requests; TK_WROWS * ceil(C/32) independent loads per lane are in flight
        // before the first maximum, and a row's result is one warp reduction on the order-preserving key (the maximum
        // of the floats is the minimum of the descending keys).  No shared-memory staging, no barrier.
        const int lane = tid & 31, w = tid >> 5;
        const float NINF = -__int_as_float(0x7f800000);
        for (long r0 = ((long)blockIdx.x * (TK_THREADS / 32) + w) * TK_WROWS; r0 < seg_len;
             r0 += (long)gridDim.x * (TK_THREADS / 32) * TK_WROWS) {
            float m[TK_WROWS];
#pragma unroll
            for (int j = 0; j < TK_WROWS; ++j) m[j] = NINF;
            const float *p0 = prob + (size_t)(seg0 + r0) * C;
            for (int c0 = lane; c0 < C; c0 += 32 * TK_WCH) {     // one trip for C <= 128
                float v[TK_WROWS][TK_WCH];                       // every load of the trip is issued before the first use
#pragma unroll
                for (int j = 0; j < TK_WROWS; ++j)
#pragma unroll
                    for (int q = 0; q < TK_WCH; ++q) {
                        const int c = c0 + 32 * q;
                        v[j][q] = (r0 + j < seg_len && c < C) ? __ldcs(p0 + (size_t)j * C + c) : NINF;   // read once: streaming
                    }
#pragma unroll
                for (int j = 0; j < TK_WROWS; ++j)
#pragma unroll
                    for (int q = 0; q < TK_WCH; ++q) m[j] = fmaxf(m[j], v[j][q]);
            }
            uint32_t mine = 0;
#pragma unroll
            for (int j = 0; j < TK_WROWS; ++j) {
                const uint32_t k = __reduce_min_sync(0xffffffffu, ~float_to_ordered(m[j]));
                if (lane == j) mine = k;
            }
            if (lane < TK_WROWS && r0 + lane < seg_len) {
                keys[seg0 + r0 + lane] = mine;
                atomicAdd(&sh[mine >> 21], 1u);
            }
        }
    }
    __syncthreads();
    uint32_t *h = hist + (size_t)blockIdx.y * NB0;
    for (int i = tid; i < NB0; i += TK_THREADS) if (sh[i]) atomicAdd(&h[i], sh[i]);
}

// ---- 2. radix select ---------------------------------------------------------------------------
// one block per segment: find the bin that holds the k_rem-th key (bins ascending = best first)
__global__ void __launch_bounds__(1024)
select_kernel(const uint32_t *__restrict__ hist, int nbins, int bits, SegState *__restrict__ state)
{
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t found_bin, found_before;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t *h = hist + (size_t)blockIdx.x * NB0;
    const int per = nbins / 1024 > 0 ? nbins / 1024 : 1;          // 2 (2048 bins) or 1
    uint32_t c[2] = {0, 0};
    uint32_t s = 0;
    for (int i = 0; i < per; ++i) { const int b = tid * per + i; c[i] = b < nbins ? h[b] : 0; s += c[i]; }
    uint32_t incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) wsum[w] = incl;
    __syncthreads();
    if (w == 0) {
        const uint32_t v = wsum[lane]; uint32_t iv = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, iv, o); if (lane >= o) iv += t; }
        wsum[lane] = iv - v;
    }
    __syncthreads();
    const uint32_t k = state[blockIdx.x].k_rem;
    uint32_t before = wsum[w] + incl - s;                        // keys in bins before this thread's
    for (int i = 0; i < per; ++i) {
        if (k > before && k <= before + c[i]) { found_bin = tid * per + i; found_before = before; }
        before += c[i];
    }
    __syncthreads();
    if (tid == 0 && k > 0) {
        SegState st = state[blockIdx.x];
        st.prefix = (st.prefix << bits) | found_bin;
        st.k_rem = k - found_before;
        st.n_better += found_before;
        state[blockIdx.x] = st;
    }
}

// histogram of the next digit among the keys that match the prefix found so far
__global__ void __launch_bounds__(TK_THREADS)
refine_kernel(const uint32_t *__restrict__ keys, long seg_len, const SegState *__restrict__ state,
              int prefix_shift, int bin_shift, uint32_t bin_mask, uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[NB0];
    const int tid = threadIdx.x;
    for (int i = tid; i < NB0; i += TK_THREADS) sh[i] = 0;
    __syncthreads();
    const uint32_t prefix = state[blockIdx.y].prefix;
    const uint32_t *k = keys + (size_t)blockIdx.y * seg_len;
    for (long r = (long)blockIdx.x * TK_THREADS + tid; r < seg_len; r += (long)gridDim.x * TK_THREADS) {
        const uint32_t v = k[r];
        if ((v >> prefix_shift) == prefix) atomicAdd(&sh[(v >> bin_shift) & bin_mask], 1u);
    }
    __syncthreads();
    uint32_t *h = hist + (size_t)blockIdx.y * NB0;
    for (int i = tid; i < NB0; i += TK_THREADS) if (sh[i]) atomicAdd(&h[i], sh[i]);
}

// ---- 3. order-preserving compaction --------------------------------------------------------------
constexpr int CP_ITEMS = 8;
constexpr int CP_TILE = TK_THREADS * CP_ITEMS;

__global__ void __launch_bounds__(TK_THREADS)
cand_count_kernel(const uint32_t *__restrict__ keys, long seg_len, const SegState *__restrict__ state,
                  uint32_t *__restrict__ blk_lt, uint32_t *__restrict__ blk_eq)
{
    const uint32_t T = state[blockIdx.y].prefix;
    const uint32_t *k = keys + (size_t)blockIdx.y * seg_len;
    const long t0 = (long)blockIdx.x * CP_TILE;
    int lt = 0, eq = 0;
    for (int j = 0; j < CP_ITEMS; ++j) {
        const long i = t0 + j * TK_THREADS + threadIdx.x;
        if (i < seg_len) { const uint32_t v = k[i]; lt += v < T; eq += v == T; }
    }
    __shared__ int s_lt, s_eq;
    if (threadIdx.x == 0) { s_lt = 0; s_eq = 0; }
    __syncthreads();
    lt = __reduce_add_sync(0xffffffffu, lt);
    eq = __reduce_add_sync(0xffffffffu, eq);
    if ((threadIdx.x & 31) == 0) { atomicAdd(&s_lt, lt); atomicAdd(&s_eq, eq); }
    __syncthreads();
    if (threadIdx.x == 0) {
        blk_lt[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = s_lt;
        blk_eq[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = s_eq;
    }
}

// exclusive scan of the per-block counts inside every segment (one block per segment)
__global__ void __launch_bounds__(1024)
cand_scan_kernel(uint32_t *__restrict__ blk_lt, uint32_t *__restrict__ blk_eq, int nblk)
{
    __shared__ uint32_t wsum[2][32];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    uint32_t *a = blk_lt + (size_t)blockIdx.x * nblk, *b = blk_eq + (size_t)blockIdx.x * nblk;
    const int per = (nblk + 1023) / 1024;
    const int begin = min(nblk, tid * per), end = min(nblk, begin + per);
    uint32_t sa = 0, sb = 0;
    for (int i = begin; i < end; ++i) { sa += a[i]; sb += b[i]; }
    uint32_t ia = sa, ib = sb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t ta = __shfl_up_sync(0xffffffffu, ia, o), tb = __shfl_up_sync(0xffffffffu, ib, o);
        if (lane >= o) { ia += ta; ib += tb; }
    }
    if (lane == 31) { wsum[0][w] = ia; wsum[1][w] = ib; }
    __syncthreads();
    if (w < 2) {
        const uint32_t v = wsum[w][lane]; uint32_t iv = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, iv, o); if (lane >= o) iv += t; }
        wsum[w][lane] = iv - v;
    }
    __syncthreads();
    uint32_t ra = wsum[0][w] + ia - sa, rb = wsum[1][w] + ib - sb;
    for (int i = begin; i < end; ++i) {
        const uint32_t ca = a[i], cb = b[i];
        a[i] = ra; b[i] = rb; ra += ca; rb += cb;
    }
}

__global__ void __launch_bounds__(TK_THREADS)
cand_write_kernel(const uint32_t *__restrict__ keys, long seg_len, const SegState *__restrict__ state,
                  const uint32_t *__restrict__ blk_lt, const uint32_t *__restrict__ blk_eq, long K,
                  uint32_t *__restrict__ ckeys, uint32_t *__restrict__ cvals)
{
    __shared__ uint32_t w_lt[TK_THREADS / 32], w_eq[TK_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const SegState st = state[blockIdx.y];
    const uint32_t T = st.prefix, need = st.k_rem;
    const uint32_t *k = keys + (size_t)blockIdx.y * seg_len;
    const long base = (long)blockIdx.x * CP_TILE + (long)w * 32 * CP_ITEMS;   // warp-striped: index order = (warp, step, lane)
    uint32_t key[CP_ITEMS], m_lt[CP_ITEMS], m_eq[CP_ITEMS];
    uint32_t tot_lt = 0, tot_eq = 0;
#pragma unroll
    for (int j = 0; j < CP_ITEMS; ++j) {
        const long i = base + j * 32 + lane;
        key[j] = i < seg_len ? k[i] : 0xffffffffu;
        const bool ok = i < seg_len;
        m_lt[j] = __ballot_sync(0xffffffffu, ok && key[j] < T);
        m_eq[j] = __ballot_sync(0xffffffffu, ok && key[j] == T);
        tot_lt += __popc(m_lt[j]); tot_eq += __popc(m_eq[j]);
    }
    if (lane == 0) { w_lt[w] = tot_lt; w_eq[w] = tot_eq; }
    __syncthreads();
    uint32_t lt_before = blk_lt[(size_t)blockIdx.y * gridDim.x + blockIdx.x];
    uint32_t eq_before = blk_eq[(size_t)blockIdx.y * gridDim.x + blockIdx.x];
    for (int i = 0; i < w; ++i) { lt_before += w_lt[i]; eq_before += w_eq[i]; }
    const uint32_t below = (1u << lane) - 1u;
    const long seg0 = (long)blockIdx.y * seg_len;
#pragma unroll
    for (int j = 0; j < CP_ITEMS; ++j) {
        const long i = base + j * 32 + lane;
        const uint32_t my_lt = lt_before + __popc(m_lt[j] & below);
        const uint32_t my_eq = eq_before + __popc(m_eq[j] & below);
        const bool is_lt = (m_lt[j] >> lane) & 1u, is_eq = (m_eq[j] >> lane) & 1u;
        if (is_lt || (is_eq && my_eq < need)) {
            const uint32_t pos = my_lt + min(my_eq, need);
            ckeys[(size_t)blockIdx.y * K + pos] = key[j];
            cvals[(size_t)blockIdx.y * K + pos] = (uint32_t)(seg0 + i);
        }
        lt_before += __popc(m_lt[j]); eq_before += __popc(m_eq[j]);
    }
}

// ---- 3'. partition + candidate select (K <= SEG_SORT_MAX) -----------------------------------------
// When the winners are sorted by (key, index) afterwards (seg_sort_kernel), their order of arrival does not
// matter, so ONE more pass over the keys is enough after the level-0 select: keys in better bins than the
// threshold bin are winners outright, keys inside it become candidates (appended to per-segment lists with
// warp-aggregated atomics); a block per segment then finds the k_rem best candidates by radix select over
// the 53-bit composite (low 21 key bits, flat index) — unique values, so the tie rule "lower flat index
// first" is part of the order.  This replaces two refine passes and the count / scan / write compaction
// (four reads of the key array) by one.
constexpr int PT_ITEMS = 8;            // keys per thread and trip (grid sizing)

// VEC: 16-byte key loads (segment start and length multiples of 4).  Nearly all keys are neither winners nor
// candidates, so a warp first asks with ONE ballot whether any of its 128 keys is; only then the per-key code runs.
template <bool VEC>
__global__ void __launch_bounds__(TK_THREADS)
partition_kernel(const uint32_t *__restrict__ keys, long seg_len, const SegState *__restrict__ state, long K,
                 uint32_t *__restrict__ ckeys, uint32_t *__restrict__ cvals,
                 uint32_t *__restrict__ cand_keys, uint32_t *__restrict__ cand_idx, uint32_t *__restrict__ cnt /*[n_seg][2]*/)
{
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t b0 = state[blockIdx.y].prefix;

    const long seg0 = (long)blockIdx.y * seg_len;
    const uint32_t *k = keys + seg0;
    uint32_t *win_cnt = cnt + 2 * blockIdx.y, *cand_cnt = win_cnt + 1;
    const uint32_t below = (1u << lane) - 1u;
    auto emit = [&](uint32_t key, long r, bool ok) {    // warp-collective: every lane calls it
        const uint32_t bin = key >> 21;
        const bool better = ok && bin < b0, cand = ok && bin == b0;
        const unsigned mb = __ballot_sync(0xffffffffu, better), mc = __ballot_sync(0xffffffffu, cand);
        if (mb) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(win_cnt, (uint32_t)__popc(mb));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (better) {
                const size_t pos = (size_t)blockIdx.y * K + base + __popc(mb & below);
                ckeys[pos] = key; cvals[pos] = (uint32_t)(seg0 + r);
            }
        }
        if (mc) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(cand_cnt, (uint32_t)__popc(mc));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (cand) {
                const size_t pos = (size_t)seg0 + base + __popc(mc & below);
                cand_keys[pos] = key; cand_idx[pos] = (uint32_t)(seg0 + r);
            }
        }
    };
    if (VEC) {
        // 8 keys per thread and trip.  The thread classifies its own keys, ONE packed warp scan (winners in the low
        // half, candidates in the high half) turns the counts into positions, and ONE 64-bit atomic per warp and trip
        // reserves both ranges — the pair of 32-bit counters is one aligned word.  (A ballot + atomic per key made
        // every trip wait for up to 16 round trips to L2.)  The next trip's loads are issued before that wait.
        // Any segment length / start: the <= 3 keys before the first 16-byte boundary and the <= 3 after the last whole
        // vector go through emit() in one warp at the end (COCO-R's 132,450 anchors per image are not a multiple of 4: the
        // scalar fallback — a ballot and an L2 round trip per key — cost 28 us there against 6 us for this path).
        const int head = (int)min((long)((4 - (((size_t)k >> 2) & 3)) & 3), seg_len);
        const long n4 = (seg_len - head) >> 2;
        const uint4 *k4 = reinterpret_cast<const uint4 *>(k + head);
        unsigned long long *cnt64 = reinterpret_cast<unsigned long long *>(win_cnt);
        const long stride = (long)gridDim.x * TK_THREADS * 2;
        long q0 = (long)blockIdx.x * TK_THREADS * 2;
        uint4 v[2], nv[2];
        auto load = [&](long base, uint4 *dst) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const long q = base + j * TK_THREADS + tid;
                dst[j] = q < n4 ? __ldcs(k4 + q) : make_uint4(~0u, ~0u, ~0u, ~0u);      // bin 2047 > b0 unless b0 == 2047: see `ok`
            }
        };
        if (q0 < n4) load(q0, v);
        for (; q0 < n4; q0 += stride) {
            if (q0 + stride < n4) load(q0 + stride, nv);
            uint32_t kk[8] = {v[0].x, v[0].y, v[0].z, v[0].w, v[1].x, v[1].y, v[1].z, v[1].w};
            unsigned packed = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const bool ok = q0 + (j >> 2) * TK_THREADS + tid < n4;
                const uint32_t bin = kk[j] >> 21;
                packed += (ok && bin < b0 ? 1u : 0u) + (ok && bin == b0 ? 0x10000u : 0u);
            }
            if (__any_sync(0xffffffffu, packed != 0)) {
                unsigned incl = packed;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
                unsigned long long base = 0;
                if (lane == 31) base = atomicAdd(cnt64, (unsigned long long)(incl & 0xffffu) | ((unsigned long long)(incl >> 16) << 32));
                base = __shfl_sync(0xffffffffu, base, 31);
                size_t wpos = (size_t)blockIdx.y * K + (uint32_t)base + ((incl - packed) & 0xffffu);
                size_t cpos = (size_t)seg0 + (uint32_t)(base >> 32) + ((incl - packed) >> 16);
                if (packed) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const long q = q0 + (j >> 2) * TK_THREADS + tid;
                        const uint32_t bin = kk[j] >> 21;
                        if (q < n4 && bin <= b0) {
                            const uint32_t idx = (uint32_t)(seg0 + head + (q << 2) + (j & 3));
                            if (bin < b0) { ckeys[wpos] = kk[j]; cvals[wpos] = idx; ++wpos; }
                            else { cand_keys[cpos] = kk[j]; cand_idx[cpos] = idx; ++cpos; }
                        }
                    }
                }
            }
            v[0] = nv[0]; v[1] = nv[1];
        }
        if (blockIdx.x == 0 && tid < 32) {               // head and tail keys, warp-collective
            const long tail0 = head + (n4 << 2);
            const int n_tail = (int)(seg_len - tail0);
            const long r = lane < head ? lane : tail0 + (lane - head);
            const bool ok = lane < head + n_tail;
            emit(ok ? k[r] : 0xffffffffu, r, ok);
        }
    } else {
        for (long r0 = (long)blockIdx.x * TK_THREADS; r0 < seg_len; r0 += (long)gridDim.x * TK_THREADS) {
            const long r = r0 + tid;
            emit(r < seg_len ? k[r] : 0xffffffffu, r, r < seg_len);
        }
    }
}

__device__ __forceinline__ void
cand_select_body(const unsigned seg, const uint32_t *__restrict__ cand_keys, const uint32_t *__restrict__ cand_idx, long seg_len,
                   const SegState *__restrict__ state, const uint32_t *__restrict__ cnt, long K,
                   uint32_t *__restrict__ ckeys, uint32_t *__restrict__ cvals)
{
    typedef unsigned long long u64;
    __shared__ uint32_t hist[NB0];
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t s_bin, s_before, s_out;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const SegState st = state[seg];
    const uint32_t c0 = cnt[2 * seg + 1];
    const uint32_t *ck = cand_keys + (size_t)seg * seg_len, *ci = cand_idx + (size_t)seg * seg_len;
    uint32_t need = st.k_rem;                    // 1 <= need <= c0
    // composite = (key & 0x1fffff) << 32 | idx, 53 bits, digits MSB first: 11, 10 (key), 11, 11, 10 (index)
    const int shifts[5] = {42, 32, 21, 10, 0};
    const int bits[5] = {11, 10, 11, 11, 10};
    u64 prefix = 0;                              // value of the digits fixed so far, already shifted into place
    int fixed_shift = 53;                        // composites with (comp >> fixed_shift) == (prefix >> fixed_shift) still compete
    u64 T = ~0ull;
    uint32_t remaining = c0;                     // competitors under the current prefix
    for (int lv = 0; lv < 5; ++lv) {
        if (need == remaining) { T = prefix | ((1ull << fixed_shift) - 1ull); break; }   // everything left wins
        const int sh = shifts[lv], nb = 1 << bits[lv];
        for (int i = tid; i < nb; i += 1024) hist[i] = 0;
        __syncthreads();
        for (uint32_t i = tid; i < c0; i += 1024) {
            const u64 comp = ((u64)(ck[i] & 0x1fffffu) << 32) | ci[i];
            if (fixed_shift == 53 || (comp >> fixed_shift) == (prefix >> fixed_shift)) atomicAdd(&hist[(comp >> sh) & (nb - 1)], 1u);
        }
        __syncthreads();
        // block scan over the bins (2 per thread for 2048 bins): find the bin holding the need-th competitor
        const int per = nb > 1024 ? 2 : 1;
        uint32_t c[2] = {0, 0}, sum = 0;
        for (int j = 0; j < per; ++j) { const int b = tid * per + j; c[j] = b < nb ? hist[b] : 0; sum += c[j]; }
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        if (lane == 31) wsum[w] = incl;
        __syncthreads();
        if (w == 0) {
            const uint32_t x = wsum[lane]; uint32_t ix = x;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, ix, o); if (lane >= o) ix += t; }
            wsum[lane] = ix - x;
        }
        __syncthreads();
        uint32_t before = wsum[w] + incl - sum;
        for (int j = 0; j < per; ++j) {
            if (need > before && need <= before + c[j]) { s_bin = tid * per + j; s_before = before; }
            before += c[j];
        }
        __syncthreads();
        prefix |= (u64)s_bin << sh;
        need -= s_before;
        remaining = hist[s_bin];
        fixed_shift = sh;
        __syncthreads();
        if (lv == 4) T = prefix;                 // all 53 bits fixed: the need-th (== only) competitor itself
    }
    // emit: every candidate with composite <= T (exactly k_rem of them), after the segment's outright winners
    if (tid == 0) s_out = 0;
    __syncthreads();
    const size_t out0 = (size_t)seg * K + st.n_better;
    for (uint32_t i0 = 0; i0 < c0; i0 += 1024) {
        const uint32_t i = i0 + tid;
        bool take = false; uint32_t key = 0, idx = 0;
        if (i < c0) { key = ck[i]; idx = ci[i]; take = (((u64)(key & 0x1fffffu) << 32) | idx) <= T; }
        const unsigned m = __ballot_sync(0xffffffffu, take);
        if (m) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(&s_out, (uint32_t)__popc(m));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (take) { const size_t pos = out0 + base + __popc(m & ((1u << lane) - 1u)); ckeys[pos] = key; cvals[pos] = idx; }
        }
    }
}

__global__ void __launch_bounds__(1024)
cand_select_kernel(const uint32_t *__restrict__ cand_keys, const uint32_t *__restrict__ cand_idx, long seg_len,
                   const SegState *__restrict__ state, const uint32_t *__restrict__ cnt, long K,
                   uint32_t *__restrict__ ckeys, uint32_t *__restrict__ cvals)
{
    cand_select_body(blockIdx.x, cand_keys, cand_idx, seg_len, state, cnt, K, ckeys, cvals);
}

// ---- 4'. sort of small segments (K <= 2048 winners each): one block per segment, bitonic network in
// shared memory over the 64-bit composites (key << 32 | flat index) — unique, so the order is the
// stable one (ties: lower flat index first).  Segments already sit image-major in ckeys / cvals.
// Up to 16,384 (key, index) composites per segment = 128 KB of the SM's 227 KB shared memory, so the reference's own
// global budget (900 * batch = 14,400 at COCO-R batch 16, libs/nets/model.py:311-312) and MaskRCNN's 6,000 per image
// (maskrcnn.py:97) take the partition + candidate-select + ONE-block sort route instead of a three-pass radix select and
// a 12-launch global radix sort (round 2: COCO-R global_14400 top-k + decode 0.270 -> see profiles/r02_coco_path.txt).
constexpr int SEG_SORT_MAX = 16384;                 // capacity of the one-block sort
constexpr int SEG_SORT_THREADS = 1024;
// up to how many winners per segment the one-block route is TAKEN (a bitonic network costs N log^2 N on one SM; above the
// crossover the three-pass radix select + global radix sort wins).  Measured on B200: profiles/r02_coco_path.txt.
static int seg_sort_limit() {
    static const int v = []() { const char *e = getenv("DETBOX_SEG_SORT_LIMIT"); int x = e ? atoi(e) : 4096; return x < 2 ? 2 : (x > SEG_SORT_MAX ? SEG_SORT_MAX : x); }();
    return v;
}
__global__ void __launch_bounds__(SEG_SORT_THREADS)
seg_sort_kernel(uint32_t *__restrict__ ckeys, uint32_t *__restrict__ cvals, int K, int N /*pow2 >= K*/)
{
    extern __shared__ unsigned long long v[];
    const int tid = threadIdx.x;
    uint32_t *k = ckeys + (size_t)blockIdx.x * K, *x = cvals + (size_t)blockIdx.x * K;
    for (int i = tid; i < N; i += blockDim.x)
        v[i] = i < K ? ((unsigned long long)k[i] << 32) | x[i] : ~0ull;
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
    for (int i = tid; i < K; i += blockDim.x) { k[i] = (uint32_t)(v[i] >> 32); x[i] = (uint32_t)v[i]; }
}

// ---- 5. gather + decode ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gather_decode_kernel(const uint32_t *__restrict__ ckeys, const uint32_t *__restrict__ cvals, long n,
                     const float *__restrict__ prob, int C, const float4 *__restrict__ deltas,
                     const float4 *__restrict__ anchors, long A, int enc,
                     float4 *__restrict__ boxes, float *__restrict__ probs_out, int64_t *__restrict__ img_ids,
                     float4 *__restrict__ sel_anchors, int64_t *__restrict__ flat_idx, float *__restrict__ scores)
{
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const uint32_t idx = cvals[i];
        const long a = idx % A;                                               // model.py:316
        const float4 anc = anchors[a];
        boxes[i] = decode_box(anc, deltas[idx], enc);                         // model.py:318 (no clip)
        if (img_ids) img_ids[i] = (int64_t)(idx / A);                         // model.py:319
        if (sel_anchors) sel_anchors[i] = anc;
        if (flat_idx) flat_idx[i] = (int64_t)idx;
        if (scores) scores[i] = ordered_to_float(~ckeys[i]);
        if (probs_out) for (int c = 0; c < C; ++c) probs_out[(size_t)i * C + c] = prob[(size_t)idx * C + c];
    }
}

// Tail of the one-block route for K <= TAIL_MAX winners per segment, as ONE launch instead of three: select among the
// threshold-bin candidates (cand_select_body), sort the segment's K winners in shared memory (the network of
// seg_sort_kernel), then gather + decode straight from the sorted list (the body of gather_decode_kernel).  One block per
// segment; the winners never travel back to HBM between the three steps.  Same results by construction; the separate
// kernels stay for TAIL_MAX < K <= seg_sort_limit() and behind DETBOX_TOPK_UNFUSED_TAIL (A/B timing).
constexpr int TAIL_MAX = 2048;
__global__ void __launch_bounds__(1024)
topk_tail_kernel(const uint32_t *__restrict__ cand_keys, const uint32_t *__restrict__ cand_idx, long seg_len,
                 const SegState *__restrict__ state, const uint32_t *__restrict__ cnt, int K, int N /*pow2 >= K*/,
                 uint32_t *ckeys, uint32_t *cvals,
                 const float *__restrict__ prob, int C, const float4 *__restrict__ deltas,
                 const float4 *__restrict__ anchors, long A, int enc,
                 float4 *__restrict__ boxes, float *__restrict__ probs_out, int64_t *__restrict__ img_ids,
                 float4 *__restrict__ sel_anchors, int64_t *__restrict__ flat_idx, float *__restrict__ scores)
{
    __shared__ unsigned long long v[TAIL_MAX];
    const int tid = threadIdx.x;
    cand_select_body(blockIdx.x, cand_keys, cand_idx, seg_len, state, cnt, (long)K, ckeys, cvals);
    __syncthreads();                                  // this block's own winners are now visible to all of its threads
    const uint32_t *k = ckeys + (size_t)blockIdx.x * K, *x = cvals + (size_t)blockIdx.x * K;
    for (int i = tid; i < N; i += 1024)
        v[i] = i < K ? ((unsigned long long)k[i] << 32) | x[i] : ~0ull;
    __syncthreads();
    for (int size = 2; size <= N; size <<= 1) {
        for (int j = size >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < (N >> 1); t += 1024) {
                const int i = 2 * t - (t & (j - 1)), q = i + j;
                const unsigned long long a = v[i], b = v[q];
                const bool up = (i & size) == 0;
                if ((a > b) == up) { v[i] = b; v[q] = a; }
            }
            __syncthreads();
        }
    }
    for (int t = tid; t < K; t += 1024) {
        const uint32_t key = (uint32_t)(v[t] >> 32), idx = (uint32_t)v[t];
        const size_t i = (size_t)blockIdx.x * K + t;
        const long a = idx % A;
        const float4 anc = anchors[a];
        boxes[i] = decode_box(anc, deltas[idx], enc);
        if (img_ids) img_ids[i] = (int64_t)(idx / A);
        if (sel_anchors) sel_anchors[i] = anc;
        if (flat_idx) flat_idx[i] = (int64_t)idx;
        if (scores) scores[i] = ordered_to_float(~key);
        if (probs_out && C <= 4) for (int c = 0; c < C; ++c) probs_out[i * C + c] = prob[(size_t)idx * C + c];
    }
    if (probs_out && C > 4) {                         // wide rows (81 COCO classes): one warp per row, lanes across the classes
        const int lane = tid & 31;
        for (int t = tid >> 5; t < K; t += 32) {
            const uint32_t idx = (uint32_t)v[t];
            const float *src = prob + (size_t)idx * C;
            float *dst = probs_out + ((size_t)blockIdx.x * K + t) * C;
            for (int c = lane; c < C; c += 32) dst[c] = src[c];
        }
    }
}

__global__ void init_state_kernel(SegState *state, int n_seg, uint32_t k)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_seg) { state[i].prefix = 0; state[i].k_rem = k; state[i].n_better = 0; state[i].pad = 0; }
}

struct TopkPlan {
    int n_seg; long seg_len; long K; long n_cand; int nblk_c;
    size_t keys_b, hist_b, state_b, blk_b, cand_b, counts_b, part_b, cnt_b, total;
};

TopkPlan make_plan(int batch, long A, long top_n, int per_image) {
    TopkPlan p;
    p.n_seg = per_image ? batch : 1;
    p.seg_len = per_image ? A : (long)batch * A;
    p.K = top_n < p.seg_len ? top_n : p.seg_len;
    if (p.K < 0) p.K = 0;
    p.n_cand = p.K * p.n_seg;
    p.nblk_c = (int)((p.seg_len + CP_TILE - 1) / CP_TILE);
    p.keys_b = align_up((size_t)batch * A * 4);
    p.hist_b = align_up((size_t)3 * p.n_seg * NB0 * 4);
    p.state_b = align_up((size_t)p.n_seg * sizeof(SegState));
    p.blk_b = align_up((size_t)p.n_seg * p.nblk_c * 4);
    p.cand_b = align_up((size_t)(p.n_cand > 0 ? p.n_cand : 1) * 4);
    p.counts_b = sort_counts_bytes(p.n_cand > 0 ? p.n_cand : 1);
    p.part_b = p.K <= seg_sort_limit() ? align_up((size_t)batch * A * 4) : 0;       // candidate lists of the partition path
    p.cnt_b = align_up((size_t)p.n_seg * 2 * 4);
    p.total = p.keys_b + p.hist_b + p.cnt_b + p.state_b + 2 * p.blk_b + 4 * p.cand_b + p.counts_b + 2 * p.part_b;
    return p;
}

}  // namespace

extern "C" size_t detbox_topk_decode_workspace_bytes(int batch, long A, long top_n, int per_image)
{
    if (batch <= 0 || A <= 0) return 256;
    return make_plan(batch, A, top_n, per_image).total;
}

extern "C" int detbox_topk_decode(const float *prob, int C, const float *deltas, const float *anchors,
                                  int batch, long A, int activation, long top_n, int per_image, int enc,
                                  float *boxes, float *probs_out, int64_t *img_ids, float *sel_anchors,
                                  int64_t *flat_idx, float *scores, void *ws, size_t ws_bytes, void *stream)
{
    if (batch < 0 || A < 0 || C <= 0 || top_n < 0 || enc < 0 || enc > 2) return DETBOX_ERR_INVALID;
    if (activation != DETBOX_ACT_SOFTMAX && activation != DETBOX_ACT_SIGMOID) return DETBOX_ERR_INVALID;
    if (batch == 0 || A == 0 || top_n == 0) return DETBOX_OK;
    if ((double)batch * (double)A >= 2147483648.0) return DETBOX_ERR_UNSUPPORTED;
    if (!prob || !deltas || !anchors || !boxes) return DETBOX_ERR_INVALID;
    const TopkPlan p = make_plan(batch, A, top_n, per_image);
    if (!ws || ws_bytes < p.total) return DETBOX_ERR_WORKSPACE;
    if (p.n_seg > 65535) return DETBOX_ERR_UNSUPPORTED;
    cudaStream_t st = as_stream(stream);
    char *w = (char *)ws;
    uint32_t *keys = (uint32_t *)w; w += p.keys_b;
    uint32_t *hist = (uint32_t *)w; w += p.hist_b;
    uint32_t *part_cnt = (uint32_t *)w; w += p.cnt_b;            // right after hist: cleared by the same memset
    SegState *state = (SegState *)w; w += p.state_b;
    uint32_t *blk_lt = (uint32_t *)w; w += p.blk_b;
    uint32_t *blk_eq = (uint32_t *)w; w += p.blk_b;
    uint32_t *cka = (uint32_t *)w; w += p.cand_b;
    uint32_t *ckb = (uint32_t *)w; w += p.cand_b;
    uint32_t *cva = (uint32_t *)w; w += p.cand_b;
    uint32_t *cvb = (uint32_t *)w; w += p.cand_b;
    uint32_t *counts = (uint32_t *)w; w += p.counts_b;
    uint32_t *cand_keys = (uint32_t *)w; w += p.part_b;
    uint32_t *cand_idx = (uint32_t *)w;

    DETBOX_CUDA_TRY(cudaMemsetAsync(hist, 0, p.hist_b + p.cnt_b, st));
    init_state_kernel<<<(p.n_seg + 255) / 256, 256, 0, st>>>(state, p.n_seg, (uint32_t)p.K);
    DETBOX_LAUNCH_CHECK();

    // 1. objectness + level-0 histogram
    {
        const bool staged = activation == DETBOX_ACT_SIGMOID && C > 4;
        const long rows_per_block = staged ? (TK_THREADS / 32) * TK_WROWS : TK_THREADS;
        long gx = (p.seg_len + rows_per_block - 1) / rows_per_block;
        const long cap = (long)kNumSMs * 8 / (p.n_seg < 8 ? 1 : 1);
        long capx = cap / p.n_seg; if (capx < 1) capx = 1; if (p.n_seg == 1) capx = cap;
        if (gx > capx) gx = capx;
        dim3 grid((unsigned)gx, (unsigned)p.n_seg);
        DETBOX_TRACE_BEGIN(DETBOX_K_OBJECTNESS, st);
        if (activation == DETBOX_ACT_SOFTMAX) objectness_kernel<DETBOX_ACT_SOFTMAX><<<grid, TK_THREADS, 0, st>>>(prob, C, p.seg_len, keys, hist);
        else objectness_kernel<DETBOX_ACT_SIGMOID><<<grid, TK_THREADS, 0, st>>>(prob, C, p.seg_len, keys, hist);
        DETBOX_TRACE_END(DETBOX_K_OBJECTNESS, st);
        DETBOX_LAUNCH_CHECK();
    }
    // 2. radix select: 11 + 11 + 10 bits
    long gx = (p.seg_len + TK_THREADS * 8 - 1) / (TK_THREADS * 8);
    { long capx = (long)kNumSMs * 8 / p.n_seg; if (capx < 1) capx = 1; if (gx > capx) gx = capx; }
    dim3 rgrid((unsigned)gx, (unsigned)p.n_seg);
    uint32_t *h0 = hist, *h1 = hist + (size_t)p.n_seg * NB0, *h2 = hist + (size_t)2 * p.n_seg * NB0;
    select_kernel<<<p.n_seg, 1024, 0, st>>>(h0, 2048, 11, state);
    DETBOX_LAUNCH_CHECK();
    if (p.K <= seg_sort_limit()) {
        // 3'. one pass: outright winners + candidates of the threshold bin; then select among the candidates
        long px = (p.seg_len + TK_THREADS * PT_ITEMS - 1) / (TK_THREADS * PT_ITEMS);
        { long capx = (long)kNumSMs * 8 / p.n_seg; if (capx < 1) capx = 1; if (px > capx) px = capx; }
        const dim3 pgrid((unsigned)px, (unsigned)p.n_seg);
        static const bool scalar_part = getenv("DETBOX_TOPK_SCALAR_PARTITION") != nullptr;    // the old fallback, for A/B timing only
        if (!scalar_part) partition_kernel<true><<<pgrid, TK_THREADS, 0, st>>>(keys, p.seg_len, state, p.K, cka, cva, cand_keys, cand_idx, part_cnt);
        else partition_kernel<false><<<pgrid, TK_THREADS, 0, st>>>(keys, p.seg_len, state, p.K, cka, cva, cand_keys, cand_idx, part_cnt);
        DETBOX_LAUNCH_CHECK();
        static const bool unfused_tail = getenv("DETBOX_TOPK_UNFUSED_TAIL") != nullptr;
        if (p.K <= TAIL_MAX && !unfused_tail) {
            int N = 2; while (N < p.K) N <<= 1;
            topk_tail_kernel<<<p.n_seg, 1024, 0, st>>>(cand_keys, cand_idx, p.seg_len, state, part_cnt, (int)p.K, N, cka, cva,
                                                      prob, C, (const float4 *)deltas, (const float4 *)anchors, A, enc,
                                                      (float4 *)boxes, probs_out, img_ids, (float4 *)sel_anchors, flat_idx, scores);
            DETBOX_LAUNCH_CHECK();
            return DETBOX_OK;
        }
        cand_select_kernel<<<p.n_seg, 1024, 0, st>>>(cand_keys, cand_idx, p.seg_len, state, part_cnt, p.K, cka, cva);
        DETBOX_LAUNCH_CHECK();
    } else {
    refine_kernel<<<rgrid, TK_THREADS, 0, st>>>(keys, p.seg_len, state, 21, 10, 0x7ffu, h1);
        DETBOX_LAUNCH_CHECK();
        select_kernel<<<p.n_seg, 1024, 0, st>>>(h1, 2048, 11, state);
        DETBOX_LAUNCH_CHECK();
        refine_kernel<<<rgrid, TK_THREADS, 0, st>>>(keys, p.seg_len, state, 10, 0, 0x3ffu, h2);
        DETBOX_LAUNCH_CHECK();
        select_kernel<<<p.n_seg, 1024, 0, st>>>(h2, 1024, 10, state);
        DETBOX_LAUNCH_CHECK();
        // 3. ordered compaction
        dim3 cgrid((unsigned)p.nblk_c, (unsigned)p.n_seg);
        cand_count_kernel<<<cgrid, TK_THREADS, 0, st>>>(keys, p.seg_len, state, blk_lt, blk_eq);
        DETBOX_LAUNCH_CHECK();
        cand_scan_kernel<<<p.n_seg, 1024, 0, st>>>(blk_lt, blk_eq, p.nblk_c);
        DETBOX_LAUNCH_CHECK();
        cand_write_kernel<<<cgrid, TK_THREADS, 0, st>>>(keys, p.seg_len, state, blk_lt, blk_eq, p.K, cka, cva);
        DETBOX_LAUNCH_CHECK();
    }
    // 4. sort the winners (stable): small segments in shared memory, else by key and then (per-image
    //    mode) by image id with the global radix sort
    int rc = DETBOX_OK;
    uint32_t *fk = cka, *fv = cva;
    if (p.K <= seg_sort_limit()) {
        int N = 2; while (N < p.K) N <<= 1;
        const int threads = N / 2 < 32 ? 32 : (N / 2 > SEG_SORT_THREADS ? SEG_SORT_THREADS : N / 2);
        const size_t smem = (size_t)N * sizeof(unsigned long long);
        if (smem > 48 * 1024) {                       // opt in to the large carve-out once per device
            static thread_local int ready_dev = -1;
            int dev = 0;
            DETBOX_CUDA_TRY(cudaGetDevice(&dev));
            if (ready_dev != dev) {
                DETBOX_CUDA_TRY(cudaFuncSetAttribute(seg_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SEG_SORT_MAX * (int)sizeof(unsigned long long)));
                ready_dev = dev;
            }
        }
        seg_sort_kernel<<<p.n_seg, threads, smem, st>>>(cka, cva, (int)p.K, N);
        DETBOX_LAUNCH_CHECK();
    } else {
        rc = radix_sort_pairs(cka, cva, ckb, cvb, p.n_cand, counts, st);
        if (rc != DETBOX_OK) return rc;
    }
    if (p.K > seg_sort_limit() && p.n_seg > 1) {
        uint32_t *ki = cka, *vi = cva, *ko = ckb, *vo = cvb;
        for (int shift = 0; (p.n_seg - 1) >> shift; shift += 8) {
            DigitSrc d{1, shift, (uint32_t)A, nullptr};
            rc = radix_pass(ki, vi, ko, vo, p.n_cand, d, counts, st);
            if (rc != DETBOX_OK) return rc;
            uint32_t *t = ki; ki = ko; ko = t; t = vi; vi = vo; vo = t;
        }
        fk = ki; fv = vi;
    }
    // 5. gather + decode
    long gb = (p.n_cand + 255) / 256; if (gb > kNumSMs * 8) gb = kNumSMs * 8;
    gather_decode_kernel<<<(unsigned)gb, 256, 0, st>>>(fk, fv, p.n_cand, prob, C, (const float4 *)deltas,
                                                      (const float4 *)anchors, A, enc, (float4 *)boxes, probs_out,
                                                      img_ids, (float4 *)sel_anchors, flat_idx, scores);
    DETBOX_LAUNCH_CHECK();
    return DETBOX_OK;
}
