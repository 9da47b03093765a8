// Fused training-mode BatchNorm for channels-last activations: y = [ReLU]( BN(x) [+ residual] ), and its
// backward, one launch each. The flagship ResNet-50 step spends ≈30 % of one worker's forward+backward chain
// in 106 cuDNN batch-norm launches plus 98 separate ReLU / ReLU-backward launches and 32 residual adds
// (profiles/README.md); on the 32×32-input shapes these tensors are 50–800 KB, so a kernel that keeps its
// channel slab in L1/L2 between the statistics pass and the normalisation pass is launch-latency bound.
//
// Layout: x is [rows, C] (rows = N·H·W, C contiguous = NHWC). A CTA owns 4·QUADS channels (QUADS ∈ {1,2,4,8}
// chosen so that at least 64 CTAs are in flight): thread = (channel quad, row lane), float4 loads, 256/QUADS
// row lanes stride over the rows. One-pass pivoted statistics; fixed-order reductions (deterministic).
// Reference counterpart: cuDNN/ATen BatchNorm2d + ReLU + residual add (torchvision resnet.py via
// garfieldpp/tools.py:69-75).

#include <cmath>

#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

constexpr int kBnThreads = 256;

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 add4(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

// A CTA owns QUADS channel quads (4·QUADS channels) and spreads its 256 threads over 256/QUADS row lanes:
// narrow layers (C = 64) get 16 CTAs × 256 row lanes instead of 2 CTAs, wide ones (C = 2048) keep 128-byte
// coalesced rows. Sum a float4 over all row lanes of the CTA; result valid in every thread of that quad.
template <int QUADS>
__device__ __forceinline__ float4 reduce_rows(float4 v, float4 (*scratch)[QUADS]) {
#pragma unroll
    for (int o = QUADS; o < 32; o <<= 1) {
        v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
        v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
        v.z += __shfl_xor_sync(0xffffffffu, v.z, o);
        v.w += __shfl_xor_sync(0xffffffffu, v.w, o);
    }
    int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31, quad = lane % QUADS;
    __syncthreads();
    if (lane < QUADS) scratch[warp][quad] = v;
    __syncthreads();
    float4 total = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int w = 0; w < kBnThreads / 32; ++w) total = add4(total, scratch[w][quad]);  // fixed order
    return total;
}

template <int QUADS>
__global__ void __launch_bounds__(kBnThreads) bn_nhwc_fwd_kernel(int rows, int C, const float* __restrict__ x, const float* __restrict__ residual,
                                                                 const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                                                 float momentum, int relu, float* __restrict__ y, float* __restrict__ save_mean,
                                                                 float* __restrict__ save_invstd, float* __restrict__ running_mean,
                                                                 float* __restrict__ running_var) {
    __shared__ float4 scratch[kBnThreads / 32][QUADS];
    constexpr int kLanes = kBnThreads / QUADS;                          // row lanes per CTA
    int const lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int const quad = lane % QUADS, rlane = lane / QUADS + (32 / QUADS) * warp;
    int const c0 = (blockIdx.x * QUADS + quad) * 4;
    bool const live = c0 < C;                                            // C is a multiple of 4
    {   // blockIdx.y = statistics group: a contiguous block of `rows` rows with its own mean/variance/running buffers
        long long const goff = static_cast<long long>(blockIdx.y) * rows * C;
        int const soff = blockIdx.y * C;
        x += goff; y += goff;
        if (residual != nullptr) residual += goff;
        save_mean += soff; save_invstd += soff;
        if (running_mean != nullptr) { running_mean += soff; running_var += soff; }
    }
    // one-pass statistics around a per-channel pivot K (the first row): Σ(x−K) and Σ(x−K)² do not suffer the
    // cancellation of E[x²] − E[x]² because K is a sample of the same distribution
    float4 const K = live ? ld4(x + c0) : make_float4(0.f, 0.f, 0.f, 0.f);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = s;
    if (live) {
#pragma unroll 4
        for (int r = rlane; r < rows; r += kLanes) {
            float4 const v = ld4(x + static_cast<long long>(r) * C + c0);
            float const a = v.x - K.x, b = v.y - K.y, c = v.z - K.z, d = v.w - K.w;
            s.x += a; s.y += b; s.z += c; s.w += d;
            q.x = fmaf(a, a, q.x); q.y = fmaf(b, b, q.y); q.z = fmaf(c, c, q.z); q.w = fmaf(d, d, q.w);
        }
    }
    s = reduce_rows<QUADS>(s, scratch);
    q = reduce_rows<QUADS>(q, scratch);
    if (!live) return;
    float const inv_rows = 1.0f / static_cast<float>(rows);
    float4 const ds = make_float4(s.x * inv_rows, s.y * inv_rows, s.z * inv_rows, s.w * inv_rows);       // mean − K
    float4 const mean = add4(K, ds);
    float4 const var = make_float4(fmaxf(q.x * inv_rows - ds.x * ds.x, 0.f), fmaxf(q.y * inv_rows - ds.y * ds.y, 0.f),
                                   fmaxf(q.z * inv_rows - ds.z * ds.z, 0.f), fmaxf(q.w * inv_rows - ds.w * ds.w, 0.f));
    float4 const invstd = make_float4(rsqrtf(var.x + eps), rsqrtf(var.y + eps), rsqrtf(var.z + eps), rsqrtf(var.w + eps));
    float4 const g = ld4(gamma + c0), b = ld4(beta + c0);
    float4 const scale = make_float4(g.x * invstd.x, g.y * invstd.y, g.z * invstd.z, g.w * invstd.w);
    float4 const shift = make_float4(b.x - mean.x * scale.x, b.y - mean.y * scale.y, b.z - mean.z * scale.z, b.w - mean.w * scale.w);
#pragma unroll 4
    for (int r = rlane; r < rows; r += kLanes) {
        long long const at = static_cast<long long>(r) * C + c0;
        float4 const v = ld4(x + at);
        float4 o = make_float4(fmaf(v.x, scale.x, shift.x), fmaf(v.y, scale.y, shift.y), fmaf(v.z, scale.z, shift.z), fmaf(v.w, scale.w, shift.w));
        if (residual != nullptr) o = add4(o, ld4(residual + at));
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

__device__ __forceinline__ float4 relu_mask(float4 d, float4 o) {
    return make_float4(o.x > 0.f ? d.x : 0.f, o.y > 0.f ? d.y : 0.f, o.z > 0.f ? d.z : 0.f, o.w > 0.f ? d.w : 0.f);
}

// dy is the gradient w.r.t. the block output; with relu the upstream mask is (y > 0). Produces dx (through
// the normalisation), dgamma, dbeta, and — when d_residual != nullptr — the masked gradient for the skip path.
template <int QUADS>
__global__ void __launch_bounds__(kBnThreads) bn_nhwc_bwd_kernel(int rows, int C, const float* __restrict__ x, const float* __restrict__ y,
                                                                 const float* __restrict__ dy, const float* __restrict__ gamma,
                                                                 const float* __restrict__ save_mean, const float* __restrict__ save_invstd, int relu,
                                                                 float* __restrict__ dx, float* __restrict__ d_residual, float* __restrict__ dgamma,
                                                                 float* __restrict__ dbeta) {
    __shared__ float4 scratch[kBnThreads / 32][QUADS];
    constexpr int kLanes = kBnThreads / QUADS;
    int const lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int const quad = lane % QUADS, rlane = lane / QUADS + (32 / QUADS) * warp;
    int const c0 = (blockIdx.x * QUADS + quad) * 4;
    bool const live = c0 < C;
    {
        long long const goff = static_cast<long long>(blockIdx.y) * rows * C;
        int const soff = blockIdx.y * C;
        x += goff; dy += goff; dx += goff;
        if (y != nullptr) y += goff;
        if (d_residual != nullptr) d_residual += goff;
        save_mean += soff; save_invstd += soff; dgamma += soff; dbeta += soff;
    }
    float4 mean = make_float4(0.f, 0.f, 0.f, 0.f), invstd = mean, g = mean;
    if (live) { mean = ld4(save_mean + c0); invstd = ld4(save_invstd + c0); g = ld4(gamma + c0); }
    float4 s1 = make_float4(0.f, 0.f, 0.f, 0.f), s2 = s1;
    if (live) {
#pragma unroll 4
        for (int r = rlane; r < rows; r += kLanes) {
            long long const at = static_cast<long long>(r) * C + c0;
            float4 d = ld4(dy + at);
            if (relu) d = relu_mask(d, ld4(y + at));
            float4 const v = ld4(x + at);
            s1 = add4(s1, d);
            s2.x = fmaf(d.x, (v.x - mean.x) * invstd.x, s2.x); s2.y = fmaf(d.y, (v.y - mean.y) * invstd.y, s2.y);
            s2.z = fmaf(d.z, (v.z - mean.z) * invstd.z, s2.z); s2.w = fmaf(d.w, (v.w - mean.w) * invstd.w, s2.w);
        }
    }
    s1 = reduce_rows<QUADS>(s1, scratch);
    s2 = reduce_rows<QUADS>(s2, scratch);
    if (!live) return;
    float const inv_rows = 1.0f / static_cast<float>(rows);
#pragma unroll 4
    for (int r = rlane; r < rows; r += kLanes) {
        long long const at = static_cast<long long>(r) * C + c0;
        float4 d = ld4(dy + at);
        if (relu) d = relu_mask(d, ld4(y + at));
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

// Channel quads per CTA: as many as keep >= 64 CTAs in flight (128-byte rows for wide layers, more CTAs for narrow ones).
inline int pick_quads(int C, int groups = 1) {
    int const quads_total = C / 4;
    for (int q = 8; q > 1; q >>= 1)
        if (quads_total / q * groups >= 64) return q;
    return 1;
}

}  // namespace
}  // namespace gb

using namespace gb;

extern "C" {

// x, y, residual: [groups, rows, C] channels-last activations (C % 4 == 0, 16-byte aligned); every group of `rows`
// rows is normalised with its own statistics (save_mean / save_invstd / running_*: [groups, C]; gamma / beta: [C],
// shared). groups = 1 is plain BatchNorm2d; groups = G is G co-located workers' batches in one launch.
GB_API int gb_bn_nhwc_forward_grouped(float const* x, float const* residual, float const* gamma, float const* beta, float* y, float* save_mean,
                                      float* save_invstd, float* running_mean, float* running_var, int groups, int rows, int C, float eps,
                                      float momentum, int relu, void* stream) {
    if (groups < 1 || groups > 65535 || rows < 1 || C < 4 || C % 4) return cuda_fail("bn_nhwc: need 1 <= groups <= 65535, rows >= 1 and C a positive multiple of 4");
    int const q = pick_quads(C, groups);
    dim3 const grid((C / 4 + q - 1) / q, groups);
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
#define GB_BN_FWD(Q) bn_nhwc_fwd_kernel<Q><<<grid, kBnThreads, 0, st>>>(rows, C, x, residual, gamma, beta, eps, momentum, relu, y, save_mean, save_invstd, running_mean, running_var)
    if (q == 8) GB_BN_FWD(8); else if (q == 4) GB_BN_FWD(4); else if (q == 2) GB_BN_FWD(2); else GB_BN_FWD(1);
#undef GB_BN_FWD
    return cuda_check(cudaGetLastError(), "bn_nhwc forward");
}

GB_API int gb_bn_nhwc_forward(float const* x, float const* residual, float const* gamma, float const* beta, float* y, float* save_mean,
                              float* save_invstd, float* running_mean, float* running_var, int rows, int C, float eps, float momentum, int relu,
                              void* stream) {
    return gb_bn_nhwc_forward_grouped(x, residual, gamma, beta, y, save_mean, save_invstd, running_mean, running_var, 1, rows, C, eps, momentum,
                                      relu, stream);
}

// dgamma / dbeta: [groups, C] (per-group parameter gradients — one row per co-located worker).
GB_API int gb_bn_nhwc_backward_grouped(float const* x, float const* y, float const* dy, float const* gamma, float const* save_mean,
                                       float const* save_invstd, float* dx, float* d_residual, float* dgamma, float* dbeta, int groups, int rows,
                                       int C, int relu, void* stream) {
    if (groups < 1 || groups > 65535 || rows < 1 || C < 4 || C % 4) return cuda_fail("bn_nhwc: need 1 <= groups <= 65535, rows >= 1 and C a positive multiple of 4");
    int const q = pick_quads(C, groups);
    dim3 const grid((C / 4 + q - 1) / q, groups);
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
#define GB_BN_BWD(Q) bn_nhwc_bwd_kernel<Q><<<grid, kBnThreads, 0, st>>>(rows, C, x, y, dy, gamma, save_mean, save_invstd, relu, dx, d_residual, dgamma, dbeta)
    if (q == 8) GB_BN_BWD(8); else if (q == 4) GB_BN_BWD(4); else if (q == 2) GB_BN_BWD(2); else GB_BN_BWD(1);
#undef GB_BN_BWD
    return cuda_check(cudaGetLastError(), "bn_nhwc backward");
}

GB_API int gb_bn_nhwc_backward(float const* x, float const* y, float const* dy, float const* gamma, float const* save_mean,
                               float const* save_invstd, float* dx, float* d_residual, float* dgamma, float* dbeta, int rows, int C, int relu,
                               void* stream) {
    return gb_bn_nhwc_backward_grouped(x, y, dy, gamma, save_mean, save_invstd, dx, d_residual, dgamma, dbeta, 1, rows, C, relu, stream);
}

}  // extern "C"
