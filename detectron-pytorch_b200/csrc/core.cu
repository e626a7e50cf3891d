// core.cu — library bookkeeping (error strings, version).
#include "common.cuh"

namespace detbox {
thread_local int g_last_cuda_error = 0;
thread_local KernelTrace g_trace = {0, nullptr, nullptr};
}

extern "C" const char *detbox_error_string(int code) {
    switch (code) {
        case DETBOX_OK: return "ok";
        case DETBOX_ERR_INVALID: return "invalid argument";
        case DETBOX_ERR_WORKSPACE: return "workspace missing or too small";
        case DETBOX_ERR_CUDA: return cudaGetErrorString((cudaError_t)detbox::g_last_cuda_error);
        case DETBOX_ERR_UNSUPPORTED: return "size beyond a documented limit";
        default: return "unknown detbox error";
    }
}
extern "C" int detbox_version(void) { return 100; }
extern "C" int detbox_last_cuda_error(void) { return detbox::g_last_cuda_error; }
extern "C" int detbox_trace_kernel(int kernel_id, void *start_event, void *stop_event) {
    if (kernel_id < 0 || (kernel_id > 0 && (!start_event || !stop_event))) return DETBOX_ERR_INVALID;
    detbox::g_trace.id = kernel_id;
    detbox::g_trace.start = (cudaEvent_t)start_event;
    detbox::g_trace.stop = (cudaEvent_t)stop_event;
    return DETBOX_OK;
}
