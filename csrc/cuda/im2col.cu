// im2col for channels-last activations: cols[(n, ho, wo), (kh, kw, c)] = x[n, ho·s − p + kh, wo·s − p + kw, c] (0 outside).
//
// Used by the worker-fusion path (garfield_b200/parallel/grouped.py): with the patch matrix in (kh, kw, c) order, the
// per-worker weight gradients of a k×k convolution are ONE batched GEMM  dW[g] = dY[g]ᵀ · cols[g]  whose output
// [G, Cout, KH·KW·Cin] is exactly the OHWI layout of the gradient rows — G library weight-gradient launches plus a
// scatter copy become one gather and one GEMM that writes in place. float4 over channels (C % 4 == 0).
// Reference counterpart: none (cuDNN's wgrad behind nn.Conv2d).

#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

__global__ void __launch_bounds__(256) im2col_nhwc_kernel(const float4* __restrict__ x, float4* __restrict__ cols, int N, int H, int W, int C4,
                                                           int KH, int KW, int stride_h, int stride_w, int pad_h, int pad_w, int Ho, int Wo,
                                                           long long total) {
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % C4);
        long long r = i / C4;
        int const kw = static_cast<int>(r % KW); r /= KW;
        int const kh = static_cast<int>(r % KH); r /= KH;
        int const wo = static_cast<int>(r % Wo); r /= Wo;
        int const ho = static_cast<int>(r % Ho);
        int const n = static_cast<int>(r / Ho);
        int const h = ho * stride_h - pad_h + kh, w = wo * stride_w - pad_w + kw;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (h >= 0 && h < H && w >= 0 && w < W) v = x[(static_cast<long long>(n) * H + h) * W * C4 + static_cast<long long>(w) * C4 + c];
        cols[i] = v;
    }
}

}  // namespace
}  // namespace gb

extern "C" GB_API int gb_im2col_nhwc(float const* x, float* cols, int N, int H, int W, int C, int KH, int KW, int stride_h, int stride_w, int pad_h,
                                     int pad_w, void* stream) {
    using namespace gb;
    if (N < 1 || H < 1 || W < 1 || C < 4 || C % 4 || KH < 1 || KW < 1 || stride_h < 1 || stride_w < 1 || pad_h < 0 || pad_w < 0)
        return cuda_fail("im2col_nhwc: bad geometry (C must be a positive multiple of 4)");
    int const Ho = (H + 2 * pad_h - KH) / stride_h + 1, Wo = (W + 2 * pad_w - KW) / stride_w + 1;
    if (Ho < 1 || Wo < 1) return cuda_fail("im2col_nhwc: empty output");
    long long const total = static_cast<long long>(N) * Ho * Wo * KH * KW * (C / 4);
    long long blocks = (total + 255) / 256;
    long long const cap = 32LL * device_sm_count();
    if (blocks > cap) blocks = cap;
    im2col_nhwc_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(cols), N, H, W, C / 4, KH, KW, stride_h, stride_w, pad_h, pad_w, Ho, Wo, total);
    return cuda_check(cudaGetLastError(), "im2col_nhwc");
}
