// sort.cuh — hand-written stable LSD radix sort of (uint32 key, uint32 value) pairs, ascending.
// Used by top-k (candidate ordering), NMS (score ordering) and the segment grouping passes.
#pragma once
#include "common.cuh"

namespace detbox {

// What a pass extracts its 8-bit digit from.
struct DigitSrc {
    int mode;                  // 0: key bits; 1: value / divisor (image id); 2: table[value] (segment id)
    int shift;                 // bit offset of the digit inside the chosen quantity
    uint32_t divisor;          // mode 1
    const int32_t *table;      // mode 2: int32 per ORIGINAL row, on the device
};

__device__ __forceinline__ uint32_t digit_of(const DigitSrc &d, uint32_t key, uint32_t val) {
    uint32_t q;
    if (d.mode == 0) q = key;
    else if (d.mode == 1) q = val / d.divisor;
    else q = (uint32_t)d.table[val];
    return (q >> d.shift) & 0xffu;
}

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 8;
constexpr int SORT_SUB = SORT_THREADS * SORT_ITEMS;     // elements ranked per sub-tile (2048)
constexpr int SORT_MAX_BLOCKS = 512;

struct SortPlan { long n; int blocks; long tile; };     // tile: elements per block (multiple of SORT_SUB)

inline SortPlan make_sort_plan(long n) {
    SortPlan p; p.n = n;
    long subs = (n + SORT_SUB - 1) / SORT_SUB;
    if (subs < 1) subs = 1;
    long per = (subs + SORT_MAX_BLOCKS - 1) / SORT_MAX_BLOCKS;     // sub-tiles per block
    p.tile = per * SORT_SUB;
    p.blocks = (int)((n + p.tile - 1) / p.tile);
    if (p.blocks < 1) p.blocks = 1;
    return p;
}

// scratch for the per-pass block histograms: uint32 [256 * blocks]
inline size_t sort_counts_bytes(long n) { return align_up((size_t)256 * make_sort_plan(n).blocks * sizeof(uint32_t)); }

// One stable pass: (keys_in, vals_in) -> (keys_out, vals_out), ordered by digit.
int radix_pass(const uint32_t *keys_in, const uint32_t *vals_in, uint32_t *keys_out, uint32_t *vals_out,
               long n, DigitSrc d, uint32_t *counts, cudaStream_t st);

// Full 32-bit key sort (4 passes).  Result ends in (keys_a, vals_a) again.
int radix_sort_pairs(uint32_t *keys_a, uint32_t *vals_a, uint32_t *keys_b, uint32_t *vals_b,
                     long n, uint32_t *counts, cudaStream_t st);

// Descending, stable argsort of n <= SMALL_SORT_MAX scores in ONE launch (single block, shared memory).
constexpr int SMALL_SORT_MAX = 4096;
int small_sort_scores_desc(const float *scores, long n, uint32_t *order, cudaStream_t st);

// keys[i] = ~ordered(scores[i]) (ascending key == descending score), vals[i] = i
int make_score_keys(const float *scores, long n, uint32_t *keys, uint32_t *vals, cudaStream_t st);

}  // namespace detbox
