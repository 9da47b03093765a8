// Fused training-mode BatchNorm for channels-last activations: y = [ReLU]( BN(x) [+ residual] ), and its
// backward, one launch each. The flagship ResNet-50 step spends ≈30 % of one worker's forward+backward chain
// in 106 cuDNN batch-norm launches plus 98 separate ReLU / ReLU-backward launches and 32 residual adds
// (profiles/README.md); on the 32×32-input shapes these tensors are 50–800 KB, so a kernel that keeps its
// channel slab in L1/L2 between the statistics pass and the normalisation pass is launch-latency bound.
//
// Layout: x is [rows, C] (rows = N·H·W, C contiguous = NHWC). A CTA owns 32 channels: thread = (channel
// quad, row lane); 8 lanes cover the 32 channels of a row with float4 loads (128 B coalesced), 32 row lanes
// stride over the rows. Two-pass statistics (mean, then Σ(x−mean)²) like the reference's ATen/cuDNN
// semantics; fixed-order reductions (deterministic).
// Reference counterpart: cuDNN/ATen BatchNorm2d + ReLU + residual add (torchvision resnet.py via
// garfieldpp/tools.py:69-75).

#include <cmath>

#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

constexpr int kBnThreads = 256;          // 8 channel quads × 32 row lanes
constexpr int kBnChannels = 32;          // channels per CTA

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

// Sum a float4 over the 32 row lanes of the CTA (lanes with equal `quad`); result valid in every thread.
__device__ __forceinline__ float4 reduce_rows(float4 v, float4 (*scratch)[8]) {
    // within a warp: row lanes differ in lane bits 3,4
    for (int o = 8; o <= 16; o <<= 1) {
        v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
        v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
        v.z += __shfl_xor_sync(0xffffffffu, v.z, o);
        v.w += __shfl_xor_sync(0xffffffffu, v.w, o);
    }
    int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31, quad = lane & 7;
    __syncthreads();
    if (lane < 8) scratch[warp][quad] = v;
    __syncthreads();
    float4 total = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int w = 0; w < kBnThreads / 32; ++w) {  // fixed order
        float4 const p = scratch[w][quad];
        total.x += p.x; total.y += p.y; total.z += p.z; total.w += p.w;
    }
    return total;
}

__global__ void __launch_bounds__(kBnThreads) bn_nhwc_fwd_kernel(int rows, int C, const float* __restrict__ x, const float* __restrict__ residual,
                                                                 const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                                                 float momentum, int relu, float* __restrict__ y, float* __restrict__ save_mean,
                                                                 float* __restrict__ save_invstd, float* __restrict__ running_mean,
                                                                 float* __restrict__ running_var) {
    __shared__ float4 scratch[kBnThreads / 32][8];
    int const lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int const quad = lane & 7, rlane = (lane >> 3) + 4 * warp;       // 32 row lanes
    int const c0 = blockIdx.x * kBnChannels + quad * 4;
    bool const live = c0 < C;                                          // C is a multiple of 4
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (live)
        for (int r = rlane; r < rows; r += 32) {
            float4 const v = ld4(x + static_cast<long long>(r) * C + c0);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
    s = reduce_rows(s, scratch);
    float const inv_rows = 1.0f / static_cast<float>(rows);
    float4 const mean = make_float4(s.x * inv_rows, s.y * inv_rows, s.z * inv_rows, s.w * inv_rows);
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    if (live)
        for (int r = rlane; r < rows; r += 32) {
            float4 const v = ld4(x + static_cast<long long>(r) * C + c0);
            float const a = v.x - mean.x, b = v.y - mean.y, c = v.z - mean.z, d = v.w - mean.w;
            q.x = fmaf(a, a, q.x); q.y = fmaf(b, b, q.y); q.z = fmaf(c, c, q.z); q.w = fmaf(d, d, q.w);
        }
    q = reduce_rows(q, scratch);
    if (!live) return;
    float4 const var = make_float4(q.x * inv_rows, q.y * inv_rows, q.z * inv_rows, q.w * inv_rows);
    float4 const invstd = make_float4(rsqrtf(var.x + eps), rsqrtf(var.y + eps), rsqrtf(var.z + eps), rsqrtf(var.w + eps));
    float4 const g = ld4(gamma + c0), b = ld4(beta + c0);
    float4 const scale = make_float4(g.x * invstd.x, g.y * invstd.y, g.z * invstd.z, g.w * invstd.w);
    float4 const shift = make_float4(b.x - mean.x * scale.x, b.y - mean.y * scale.y, b.z - mean.z * scale.z, b.w - mean.w * scale.w);
    for (int r = rlane; r < rows; r += 32) {
        long long const at = static_cast<long long>(r) * C + c0;
        float4 const v = ld4(x + at);
        float4 o = make_float4(fmaf(v.x, scale.x, shift.x), fmaf(v.y, scale.y, shift.y), fmaf(v.z, scale.z, shift.z), fmaf(v.w, scale.w, shift.w));
        if (residual != nullptr) {
            float4 const e = ld4(residual + at);
            o.x += e.x; o.y += e.y; o.z += e.z; o.w += e.w;
        }
        if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
        st4(y + at, o);
    }
    if (rlane == 0) {
        st4(save_mean + c0, mean);
        st4(save_invstd + c0, invstd);
        if (running_mean != nullptr) {
            float const unbias = rows > 1 ? static_cast<float>(rows) / static_cast<float>(rows - 1) : 1.0f;
            float4 rm = ld4(running_mean + c0), rv = ld4(running_var + c0);
            float const k = 1.0f - momentum;
            rm = make_float4(k * rm.x + momentum * mean.x, k * rm.y + momentum * mean.y, k * rm.z + momentum * mean.z, k * rm.w + momentum * mean.w);
            rv = make_float4(k * rv.x + momentum * var.x * unbias, k * rv.y + momentum * var.y * unbias, k * rv.z + momentum * var.z * unbias,
                             k * rv.w + momentum * var.w * unbias);
            st4(running_mean + c0, rm);
            st4(running_var + c0, rv);
        }
    }
}

// dy is the gradient w.r.t. the block output; with relu the upstream mask is (y > 0). Produces dx (through
// the normalisation), dgamma, dbeta, and — when d_residual != nullptr — the masked gradient for the skip path.
__global__ void __launch_bounds__(kBnThreads) bn_nhwc_bwd_kernel(int rows, int C, const float* __restrict__ x, const float* __restrict__ y,
                                                                 const float* __restrict__ dy, const float* __restrict__ gamma,
                                                                 const float* __restrict__ save_mean, const float* __restrict__ save_invstd, int relu,
                                                                 float* __restrict__ dx, float* __restrict__ d_residual, float* __restrict__ dgamma,
                                                                 float* __restrict__ dbeta) {
    __shared__ float4 scratch[kBnThreads / 32][8];
    int const lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int const quad = lane & 7, rlane = (lane >> 3) + 4 * warp;
    int const c0 = blockIdx.x * kBnChannels + quad * 4;
    bool const live = c0 < C;
    float4 mean = make_float4(0.f, 0.f, 0.f, 0.f), invstd = mean, g = mean;
    if (live) { mean = ld4(save_mean + c0); invstd = ld4(save_invstd + c0); g = ld4(gamma + c0); }
    float4 s1 = make_float4(0.f, 0.f, 0.f, 0.f), s2 = s1;
    if (live)
        for (int r = rlane; r < rows; r += 32) {
            long long const at = static_cast<long long>(r) * C + c0;
            float4 d = ld4(dy + at);
            if (relu) {
                float4 const o = ld4(y + at);
                d.x = o.x > 0.f ? d.x : 0.f; d.y = o.y > 0.f ? d.y : 0.f; d.z = o.z > 0.f ? d.z : 0.f; d.w = o.w > 0.f ? d.w : 0.f;
            }
            float4 const v = ld4(x + at);
            s1.x += d.x; s1.y += d.y; s1.z += d.z; s1.w += d.w;
            s2.x = fmaf(d.x, (v.x - mean.x) * invstd.x, s2.x); s2.y = fmaf(d.y, (v.y - mean.y) * invstd.y, s2.y);
            s2.z = fmaf(d.z, (v.z - mean.z) * invstd.z, s2.z); s2.w = fmaf(d.w, (v.w - mean.w) * invstd.w, s2.w);
        }
    s1 = reduce_rows(s1, scratch);
    s2 = reduce_rows(s2, scratch);
    if (!live) return;
    float const inv_rows = 1.0f / static_cast<float>(rows);
    for (int r = rlane; r < rows; r += 32) {
        long long const at = static_cast<long long>(r) * C + c0;
        float4 d = ld4(dy + at);
        if (relu) {
            float4 const o = ld4(y + at);
            d.x = o.x > 0.f ? d.x : 0.f; d.y = o.y > 0.f ? d.y : 0.f; d.z = o.z > 0.f ? d.z : 0.f; d.w = o.w > 0.f ? d.w : 0.f;
        }
        if (d_residual != nullptr) st4(d_residual + at, d);
        float4 const v = ld4(x + at);
        float4 o;
        o.x = g.x * invstd.x * (d.x - inv_rows * (s1.x + (v.x - mean.x) * invstd.x * s2.x));
        o.y = g.y * invstd.y * (d.y - inv_rows * (s1.y + (v.y - mean.y) * invstd.y * s2.y));
        o.z = g.z * invstd.z * (d.z - inv_rows * (s1.z + (v.z - mean.z) * invstd.z * s2.z));
        o.w = g.w * invstd.w * (d.w - inv_rows * (s1.w + (v.w - mean.w) * invstd.w * s2.w));
        st4(dx + at, o);
    }
    if (rlane == 0) {
        st4(dgamma + c0, s2);
        st4(dbeta + c0, s1);
    }
}

}  // namespace
}  // namespace gb

using namespace gb;

extern "C" {

// x, y, residual: [rows, C] channels-last activations (C % 4 == 0, 16-byte aligned). relu / residual optional.
GB_API int gb_bn_nhwc_forward(float const* x, float const* residual, float const* gamma, float const* beta, float* y, float* save_mean,
                              float* save_invstd, float* running_mean, float* running_var, int rows, int C, float eps, float momentum, int relu,
                              void* stream) {
    if (rows < 1 || C < 4 || C % 4) return cuda_fail("bn_nhwc: need rows >= 1 and C a positive multiple of 4");
    bn_nhwc_fwd_kernel<<<(C + kBnChannels - 1) / kBnChannels, kBnThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        rows, C, x, residual, gamma, beta, eps, momentum, relu, y, save_mean, save_invstd, running_mean, running_var);
    return cuda_check(cudaGetLastError(), "bn_nhwc forward");
}

GB_API int gb_bn_nhwc_backward(float const* x, float const* y, float const* dy, float const* gamma, float const* save_mean,
                               float const* save_invstd, float* dx, float* d_residual, float* dgamma, float* dbeta, int rows, int C, int relu,
                               void* stream) {
    if (rows < 1 || C < 4 || C % 4) return cuda_fail("bn_nhwc: need rows >= 1 and C a positive multiple of 4");
    bn_nhwc_bwd_kernel<<<(C + kBnChannels - 1) / kBnChannels, kBnThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        rows, C, x, y, dy, gamma, save_mean, save_invstd, relu, dx, d_residual, dgamma, dbeta);
    return cuda_check(cudaGetLastError(), "bn_nhwc backward");
}

}  // extern "C"
