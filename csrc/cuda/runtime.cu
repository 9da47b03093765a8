// Device runtime behind the C ABI: error strings, device facts, peer-memory arena (CUDA IPC),
// small utilities. No torch dependency: Python passes raw pointers and the stream handle.

#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {

namespace {
thread_local std::string tl_error;
}

int cuda_fail(char const* message) {
    tl_error = message;
    return 1;
}

int cuda_check(cudaError_t err, char const* what) {
    if (err == cudaSuccess) return 0;
    cudaGetLastError();  // clear the (non-sticky) error so it is not misattributed to a later call
    tl_error = std::string(what) + ": " + cudaGetErrorName(err) + " (" + cudaGetErrorString(err) + ")";
    return 1;
}

int device_sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
        cached[dev] = sms;
    }
    return cached[dev];
}

}  // namespace gb

extern "C" {

GB_API char const* gb_cuda_last_error(void) { return gb::tl_error.c_str(); }

GB_API int gb_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

GB_API int gb_cuda_sm_count(void) { return gb::device_sm_count(); }

// ---- peer-memory arena: cudaMalloc'd buffers shared between the ranks of one node via CUDA IPC.

GB_API int gb_ipc_handle_size(void) { return static_cast<int>(sizeof(cudaIpcMemHandle_t)); }

GB_API int gb_arena_alloc(int64_t bytes, void** ptr) {
    return gb::cuda_check(cudaMalloc(ptr, static_cast<size_t>(bytes)), "arena cudaMalloc");
}

GB_API int gb_arena_free(void* ptr) { return gb::cuda_check(cudaFree(ptr), "arena cudaFree"); }

GB_API int gb_ipc_export(void* ptr, void* handle_out) {
    cudaIpcMemHandle_t h;
    if (gb::cuda_check(cudaIpcGetMemHandle(&h, ptr), "cudaIpcGetMemHandle")) return 1;
    std::memcpy(handle_out, &h, sizeof(h));
    return 0;
}

GB_API int gb_ipc_open(void const* handle, void** ptr) {
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    return gb::cuda_check(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
}

GB_API int gb_ipc_close(void* ptr) { return gb::cuda_check(cudaIpcCloseMemHandle(ptr), "cudaIpcCloseMemHandle"); }

// Same-process multi-device use: enable peer access from the current device to `peer`.
GB_API int gb_enable_peer(int peer) {
    cudaError_t err = cudaDeviceEnablePeerAccess(peer, 0);
    if (err == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        return 0;
    }
    return gb::cuda_check(err, "cudaDeviceEnablePeerAccess");
}

}  // extern "C"
