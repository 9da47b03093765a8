// Programmatic dependent launch for the small kernels between the convolution / BatchNorm launches of a gradient pass (pooling,
// loss, patch matrices, zero-insertion, ReLU, gradient sums). Launched with the attribute, a kernel's CTAs are scheduled while its
// predecessor drains and wait in pdl_enter() until that predecessor has completed and flushed; without it the two device calls are
// no-ops. OPT-IN (GARFIELD_PDL_SMALL=1): measured on the graphed ResNet-50 pass it buys 10 µs of 1.33 ms with one worker per GPU and
// LOSES 45 µs of 1.96 ms with eight (together with the own gradient-sum kernel) — inside a CUDA graph the launch gaps around these
// ≈40 kernels are already small, and early-scheduled CTAs of the next kernel take slots from the one that is finishing.
#pragma once

#include <cuda_runtime.h>

#include <cstdlib>
#include <utility>

namespace gb {

inline bool small_kernel_pdl_enabled() {
    static int const on = [] {
        char const* v = std::getenv("GARFIELD_PDL");
        char const* w = std::getenv("GARFIELD_PDL_SMALL");
        return (w != nullptr && w[0] == '1' && !(v != nullptr && v[0] == '0')) ? 1 : 0;
    }();
    return on != 0;
}

// First statement of a kernel launched through launch_pdl: wait for the producer grid, then let the consumer grid be scheduled.
__device__ __forceinline__ void pdl_enter() {
    cudaGridDependencySynchronize();
    cudaTriggerProgrammaticLaunchCompletion();
}

template <typename... Params, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(Params...), dim3 grid, dim3 block, cudaStream_t stream, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = small_kernel_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<Params>(std::forward<Args>(args))...);
}

}  // namespace gb
