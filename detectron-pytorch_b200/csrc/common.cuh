// common.cuh — shared helpers for libdetbox.so (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/detbox.h"

namespace detbox {

extern thread_local int g_last_cuda_error;

inline int cuda_fail(cudaError_t e) {
    g_last_cuda_error = (int)e;
    return DETBOX_ERR_CUDA;
}

#define DETBOX_CUDA_TRY(expr)                                             \
    do {                                                                  \
        cudaError_t _e = (expr);                                          \
        if (_e != cudaSuccess) return ::detbox::cuda_fail(_e);            \
    } while (0)

#define DETBOX_LAUNCH_CHECK() DETBOX_CUDA_TRY(cudaGetLastError())

// detbox_trace_kernel(): events recorded around the launches of one kernel id (bench roofline leg)
struct KernelTrace { int id; cudaEvent_t start, stop; };
extern thread_local KernelTrace g_trace;
#define DETBOX_TRACE_BEGIN(kid, st) do { if (::detbox::g_trace.id == (kid)) cudaEventRecord(::detbox::g_trace.start, (st)); } while (0)
#define DETBOX_TRACE_END(kid, st)   do { if (::detbox::g_trace.id == (kid)) cudaEventRecord(::detbox::g_trace.stop, (st)); } while (0)

inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

// carve typed arrays out of a caller-provided workspace
struct Carver {
    char *base; size_t off;
    explicit Carver(void *p) : base(reinterpret_cast<char *>(p)), off(0) {}
    template <typename T> T *take(size_t n) {
        T *r = reinterpret_cast<T *>(base + off);
        off += align_up(n * sizeof(T));
        return r;
    }
};

constexpr int kNumSMs = 148;   // B200

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ int warp_id() { return threadIdx.x >> 5; }

// order-preserving map float32 -> uint32 (ascending); ~key gives descending
__device__ __forceinline__ uint32_t float_to_ordered(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t k) {
    uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    return __uint_as_float(u);
}

}  // namespace detbox
