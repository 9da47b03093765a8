// Hand-written layer kernels for the small reference networks (MNIST `Net`, `Cifarnet`, `CNN`):
// direct NCHW convolution (forward / data-gradient / weight+bias-gradient), max-pooling, ReLU,
// inverted dropout, training-mode batch normalisation, and fused softmax losses
// (cross-entropy = log-softmax + NLL) with their gradients. fp32, any stride/padding.
//
// Reference counterpart: the cuDNN / ATen calls behind nn.Conv2d, MaxPool2d, ReLU, Dropout(2d),
// BatchNorm2d, NLLLoss, CrossEntropyLoss (garfieldpp/models/nets.py:40-144, tools.py:47-57;
// SURVEY.md §2.3 L1, L3-L5). The layer shapes of these networks are tiny (≤ 16 output channels on
// ≤ 32×32 images), so a direct SIMT formulation with coalesced output indexing is the right tool;
// the GEMM-shaped layers (nn.Linear) go to the tcgen05 kernel in gemm_tc.cu.

#include <cfloat>
#include <cmath>

#include "gar_common.cuh"
#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

constexpr int kNnThreads = 256;

inline int nn_grid(long long items) {
    long long blocks = (items + kNnThreads - 1) / kNnThreads;
    long long const cap = static_cast<long long>(device_sm_count()) * 16;
    if (blocks > cap) blocks = cap;
    return static_cast<int>(blocks < 1 ? 1 : blocks);
}

struct ConvShape {
    int n, cin, h, w, cout, kh, kw, stride, pad, ho, wo;
};

__global__ void __launch_bounds__(kNnThreads) conv_fwd_kernel(ConvShape s, const float* __restrict__ x, const float* __restrict__ wgt,
                                                              const float* __restrict__ bias, float* __restrict__ y, int relu) {
    long long const total = static_cast<long long>(s.n) * s.cout * s.ho * s.wo;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const wo = idx % s.wo, ho = (idx / s.wo) % s.ho, co = (idx / (static_cast<long long>(s.wo) * s.ho)) % s.cout;
        int const n = idx / (static_cast<long long>(s.wo) * s.ho * s.cout);
        float acc = bias != nullptr ? bias[co] : 0.0f;
        for (int ci = 0; ci < s.cin; ++ci) {
            const float* xp = x + (static_cast<long long>(n) * s.cin + ci) * s.h * s.w;
            const float* wp = wgt + (static_cast<long long>(co) * s.cin + ci) * s.kh * s.kw;
            for (int r = 0; r < s.kh; ++r) {
                int const hi = ho * s.stride - s.pad + r;
                if (hi < 0 || hi >= s.h) continue;
                for (int c = 0; c < s.kw; ++c) {
                    int const wi = wo * s.stride - s.pad + c;
                    if (wi < 0 || wi >= s.w) continue;
                    acc = fmaf(xp[hi * s.w + wi], __ldg(wp + r * s.kw + c), acc);
                }
            }
        }
        y[idx] = relu ? fmaxf(acc, 0.0f) : acc;
    }
}

__global__ void __launch_bounds__(kNnThreads) conv_dgrad_kernel(ConvShape s, const float* __restrict__ dy, const float* __restrict__ wgt,
                                                                float* __restrict__ dx) {
    long long const total = static_cast<long long>(s.n) * s.cin * s.h * s.w;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const wi = idx % s.w, hi = (idx / s.w) % s.h, ci = (idx / (static_cast<long long>(s.w) * s.h)) % s.cin;
        int const n = idx / (static_cast<long long>(s.w) * s.h * s.cin);
        float acc = 0.0f;
        for (int co = 0; co < s.cout; ++co) {
            const float* dyp = dy + (static_cast<long long>(n) * s.cout + co) * s.ho * s.wo;
            const float* wp = wgt + (static_cast<long long>(co) * s.cin + ci) * s.kh * s.kw;
            for (int r = 0; r < s.kh; ++r) {
                int const t = hi + s.pad - r;
                if (t < 0 || t % s.stride) continue;
                int const ho = t / s.stride;
                if (ho >= s.ho) continue;
                for (int c = 0; c < s.kw; ++c) {
                    int const u = wi + s.pad - c;
                    if (u < 0 || u % s.stride) continue;
                    int const wo = u / s.stride;
                    if (wo >= s.wo) continue;
                    acc = fmaf(dyp[ho * s.wo + wo], __ldg(wp + r * s.kw + c), acc);
                }
            }
        }
        dx[idx] = acc;
    }
}

__device__ __forceinline__ float block_sum(float v, float* scratch) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    float total = 0.0f;
    for (int i = 0; i < (blockDim.x >> 5); ++i) total += scratch[i];  // fixed order → deterministic
    return total;
}

// one block per weight element (co, ci, r, c): reduce over (n, ho, wo); blocks [nweights, nweights+cout) do the bias
__global__ void __launch_bounds__(kNnThreads) conv_wgrad_kernel(ConvShape s, const float* __restrict__ dy, const float* __restrict__ x,
                                                                float* __restrict__ dw, float* __restrict__ db) {
    __shared__ float scratch[kNnThreads / 32];
    int const nweights = s.cout * s.cin * s.kh * s.kw;
    long long const spatial = static_cast<long long>(s.ho) * s.wo;
    long long const terms = static_cast<long long>(s.n) * spatial;
    int const b = blockIdx.x;
    float acc = 0.0f;
    if (b < nweights) {
        int const c = b % s.kw, r = (b / s.kw) % s.kh, ci = (b / (s.kw * s.kh)) % s.cin, co = b / (s.kw * s.kh * s.cin);
        for (long long t = threadIdx.x; t < terms; t += blockDim.x) {
            int const wo = t % s.wo, ho = (t / s.wo) % s.ho, n = t / spatial;
            int const hi = ho * s.stride - s.pad + r, wi = wo * s.stride - s.pad + c;
            if (hi < 0 || hi >= s.h || wi < 0 || wi >= s.w) continue;
            acc = fmaf(dy[(static_cast<long long>(n) * s.cout + co) * spatial + ho * s.wo + wo],
                       x[((static_cast<long long>(n) * s.cin + ci) * s.h + hi) * s.w + wi], acc);
        }
        float const total = block_sum(acc, scratch);
        if (threadIdx.x == 0) dw[b] = total;
    } else if (db != nullptr) {
        int const co = b - nweights;
        for (long long t = threadIdx.x; t < terms; t += blockDim.x) {
            int const n = t / spatial;
            acc += dy[(static_cast<long long>(n) * s.cout + co) * spatial + (t % spatial)];
        }
        float const total = block_sum(acc, scratch);
        if (threadIdx.x == 0) db[co] = total;
    }
}

__global__ void __launch_bounds__(kNnThreads) maxpool_fwd_kernel(int planes, int h, int w, int k, int stride, int ho, int wo,
                                                                 const float* __restrict__ x, float* __restrict__ y, int* __restrict__ arg) {
    long long const total = static_cast<long long>(planes) * ho * wo;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const ow = idx % wo, oh = (idx / wo) % ho;
        long long const p = idx / (static_cast<long long>(wo) * ho);
        const float* xp = x + p * h * w;
        float best = -FLT_MAX;
        int where = 0;
        for (int r = 0; r < k; ++r)
            for (int c = 0; c < k; ++c) {
                int const hi = oh * stride + r, wi = ow * stride + c;
                if (hi < h && wi < w) {
                    float const v = xp[hi * w + wi];
                    if (v > best || v != v) {  // NaN propagates like torch
                        best = v;
                        where = hi * w + wi;
                    }
                }
            }
        y[idx] = best;
        arg[idx] = where;
    }
}

__global__ void __launch_bounds__(kNnThreads) maxpool_bwd_kernel(long long total_out, int hw_in, int hw_out, const float* __restrict__ dy,
                                                                 const int* __restrict__ arg, float* __restrict__ dx) {
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total_out;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        long long const p = idx / hw_out;
        atomicAdd(dx + p * hw_in + arg[idx], dy[idx]);  // overlapping windows (stride < k) may collide
    }
}

// rows × classes logits; mode 0: input already log-probabilities (NLLLoss), 1: raw logits (CrossEntropyLoss).
// Writes per-row losses and, when dlogits != nullptr, d(mean loss)/d(input).
__global__ void __launch_bounds__(128) softmax_loss_kernel(int rows, int classes, int mode, const float* __restrict__ in,
                                                           const long long* __restrict__ target, float* __restrict__ row_loss,
                                                           float* __restrict__ dlogits) {
    int const row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float* p = in + static_cast<long long>(row) * classes;
    long long const t = target[row];
    float lse = 0.0f;
    if (mode == 1) {
        float mx = -FLT_MAX;
        for (int c = lane; c < classes; c += 32) mx = fmaxf(mx, p[c]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.0f;
        for (int c = lane; c < classes; c += 32) sum += expf(p[c] - mx);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        lse = mx + logf(sum);
    }
    if (lane == 0) row_loss[row] = lse - p[t];
    if (dlogits != nullptr) {
        float const scale = 1.0f / static_cast<float>(rows);
        float* g = dlogits + static_cast<long long>(row) * classes;
        for (int c = lane; c < classes; c += 32) {
            float const soft = mode == 1 ? expf(p[c] - lse) : 0.0f;
            g[c] = (soft - (c == t ? 1.0f : 0.0f)) * scale;
        }
    }
}

// Training-mode BatchNorm2d over NCHW: one block per channel.
__global__ void __launch_bounds__(kNnThreads) bn_fwd_kernel(int n, int c, int hw, const float* __restrict__ x, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float eps, float momentum, float* __restrict__ y,
                                                            float* __restrict__ save_mean, float* __restrict__ save_invstd,
                                                            float* __restrict__ running_mean, float* __restrict__ running_var, int relu) {
    __shared__ float scratch[kNnThreads / 32];
    int const ch = blockIdx.x;
    long long const count = static_cast<long long>(n) * hw;
    float sum = 0.0f;
    for (long long t = threadIdx.x; t < count; t += blockDim.x) sum += x[(t / hw * c + ch) * hw + t % hw];
    float const mean = block_sum(sum, scratch) / static_cast<float>(count);
    float sq = 0.0f;
    for (long long t = threadIdx.x; t < count; t += blockDim.x) {
        float const dv = x[(t / hw * c + ch) * hw + t % hw] - mean;
        sq = fmaf(dv, dv, sq);
    }
    float const var = block_sum(sq, scratch) / static_cast<float>(count);
    float const invstd = rsqrtf(var + eps);
    float const g = gamma != nullptr ? gamma[ch] : 1.0f, b = beta != nullptr ? beta[ch] : 0.0f;
    for (long long t = threadIdx.x; t < count; t += blockDim.x) {
        long long const at = (t / hw * c + ch) * hw + t % hw;
        float const v = fmaf((x[at] - mean) * invstd, g, b);
        y[at] = relu ? fmaxf(v, 0.0f) : v;
    }
    if (threadIdx.x == 0) {
        save_mean[ch] = mean;
        save_invstd[ch] = invstd;
        if (running_mean != nullptr) {
            float const unbiased = count > 1 ? var * static_cast<float>(count) / static_cast<float>(count - 1) : var;
            running_mean[ch] = (1.0f - momentum) * running_mean[ch] + momentum * mean;
            running_var[ch] = (1.0f - momentum) * running_var[ch] + momentum * unbiased;
        }
    }
}

__global__ void __launch_bounds__(kNnThreads) bn_bwd_kernel(int n, int c, int hw, const float* __restrict__ x, const float* __restrict__ dy,
                                                            const float* __restrict__ gamma, const float* __restrict__ save_mean,
                                                            const float* __restrict__ save_invstd, float* __restrict__ dx,
                                                            float* __restrict__ dgamma, float* __restrict__ dbeta) {
    __shared__ float scratch[kNnThreads / 32];
    int const ch = blockIdx.x;
    long long const count = static_cast<long long>(n) * hw;
    float const mean = save_mean[ch], invstd = save_invstd[ch], g = gamma != nullptr ? gamma[ch] : 1.0f;
    float s1 = 0.0f, s2 = 0.0f;
    for (long long t = threadIdx.x; t < count; t += blockDim.x) {
        long long const at = (t / hw * c + ch) * hw + t % hw;
        float const d = dy[at];
        s1 += d;
        s2 = fmaf(d, (x[at] - mean) * invstd, s2);
    }
    float const sum_dy = block_sum(s1, scratch);
    float const sum_dy_xhat = block_sum(s2, scratch);
    float const inv_count = 1.0f / static_cast<float>(count);
    for (long long t = threadIdx.x; t < count; t += blockDim.x) {
        long long const at = (t / hw * c + ch) * hw + t % hw;
        float const xhat = (x[at] - mean) * invstd;
        dx[at] = g * invstd * (dy[at] - inv_count * (sum_dy + xhat * sum_dy_xhat));
    }
    if (threadIdx.x == 0) {
        if (dgamma != nullptr) dgamma[ch] = sum_dy_xhat;
        if (dbeta != nullptr) dbeta[ch] = sum_dy;
    }
}


// Global average pooling over the spatial dims of NCHW planes: y[p] = mean(x[p, :]); backward broadcasts dy/hw.
__global__ void __launch_bounds__(kNnThreads) avgpool_fwd_kernel(long long planes, int hw, const float* __restrict__ x, float* __restrict__ y) {
    long long const p = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (p >= planes) return;
    int const lane = threadIdx.x & 31;
    float s = 0.0f;
    for (int i = lane; i < hw; i += 32) s += x[p * hw + i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) y[p] = s / static_cast<float>(hw);
}

__global__ void __launch_bounds__(kNnThreads) avgpool_bwd_kernel(long long total, int hw, const float* __restrict__ dy, float* __restrict__ dx) {
    float const inv = 1.0f / static_cast<float>(hw);
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x)
        dx[i] = dy[i / hw] * inv;
}

// Inverted dropout with a counter-based Philox mask: y = keep ? x / (1 - p) : 0. The same (seed, index) pair
// regenerates the mask in the backward pass, so no mask tensor is stored.
__global__ void __launch_bounds__(kNnThreads) dropout_kernel(long long total, float p, unsigned long long seed, const float* __restrict__ x,
                                                             float* __restrict__ y) {
    float const scale = 1.0f / (1.0f - p);
    long long const quads = (total + 3) / 4;
    for (long long q = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; q < quads; q += static_cast<long long>(gridDim.x) * blockDim.x) {
        uint4 const r = philox4x32(seed, static_cast<unsigned long long>(q), 3u);
        uint32_t const bits[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            long long const i = q * 4 + k;
            if (i < total) y[i] = (u01(bits[k]) >= p) ? x[i] * scale : 0.0f;
        }
    }
}

// Binary cross-entropy on probabilities (nn.BCELoss, mean reduction, log clamped at -100 like torch):
// per-element loss and d(mean loss)/d(prob).
__global__ void __launch_bounds__(kNnThreads) bce_kernel(long long total, const float* __restrict__ prob, const float* __restrict__ target,
                                                         float* __restrict__ loss, float* __restrict__ dprob) {
    float const inv = 1.0f / static_cast<float>(total);
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float const q = prob[i], t = target[i];
        float const lq = fmaxf(logf(q), -100.0f), l1q = fmaxf(logf(1.0f - q), -100.0f);
        loss[i] = -(t * lq + (1.0f - t) * l1q);
        if (dprob != nullptr) dprob[i] = (q - t) / fmaxf(q * (1.0f - q), 1e-12f) * inv;
    }
}

}  // namespace
}  // namespace gb

using namespace gb;

extern "C" {

GB_API int gb_conv2d_forward(float const* x, float const* w, float const* bias, float* y, int n, int cin, int h, int wd, int cout, int kh, int kw,
                             int stride, int pad, int relu, void* stream) {
    ConvShape s{n, cin, h, wd, cout, kh, kw, stride, pad, (h + 2 * pad - kh) / stride + 1, (wd + 2 * pad - kw) / stride + 1};
    if (s.ho < 1 || s.wo < 1 || stride < 1) return cuda_fail("conv2d: empty output");
    conv_fwd_kernel<<<nn_grid(static_cast<long long>(n) * cout * s.ho * s.wo), kNnThreads, 0, static_cast<cudaStream_t>(stream)>>>(s, x, w, bias, y, relu);
    return cuda_check(cudaGetLastError(), "conv2d forward");
}

GB_API int gb_conv2d_backward(float const* x, float const* w, float const* dy, float* dx, float* dw, float* db, int n, int cin, int h, int wd,
                              int cout, int kh, int kw, int stride, int pad, void* stream) {
    ConvShape s{n, cin, h, wd, cout, kh, kw, stride, pad, (h + 2 * pad - kh) / stride + 1, (wd + 2 * pad - kw) / stride + 1};
    if (s.ho < 1 || s.wo < 1 || stride < 1) return cuda_fail("conv2d: empty output");
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    if (dx != nullptr) conv_dgrad_kernel<<<nn_grid(static_cast<long long>(n) * cin * h * wd), kNnThreads, 0, st>>>(s, dy, w, dx);
    if (dw != nullptr) conv_wgrad_kernel<<<cout * cin * kh * kw + (db != nullptr ? cout : 0), kNnThreads, 0, st>>>(s, dy, x, dw, db);
    return cuda_check(cudaGetLastError(), "conv2d backward");
}

GB_API int gb_maxpool2d_forward(float const* x, float* y, int* arg, int planes, int h, int w, int k, int stride, void* stream) {
    int const ho = (h - k) / stride + 1, wo = (w - k) / stride + 1;
    if (ho < 1 || wo < 1) return cuda_fail("maxpool: empty output");
    maxpool_fwd_kernel<<<nn_grid(static_cast<long long>(planes) * ho * wo), kNnThreads, 0, static_cast<cudaStream_t>(stream)>>>(planes, h, w, k, stride, ho, wo, x, y, arg);
    return cuda_check(cudaGetLastError(), "maxpool forward");
}

// dx must be zero-filled by the caller.
GB_API int gb_maxpool2d_backward(float const* dy, int const* arg, float* dx, int planes, int hw_in, int hw_out, void* stream) {
    long long const total = static_cast<long long>(planes) * hw_out;
    maxpool_bwd_kernel<<<nn_grid(total), kNnThreads, 0, static_cast<cudaStream_t>(stream)>>>(total, hw_in, hw_out, dy, arg, dx);
    return cuda_check(cudaGetLastError(), "maxpool backward");
}

GB_API int gb_softmax_loss(float const* in, int64_t const* target, float* row_loss, float* dlogits, int rows, int classes, int mode, void* stream) {
    if (rows < 1 || classes < 1) return cuda_fail("softmax loss: empty input");
    softmax_loss_kernel<<<(rows + 3) / 4, 128, 0, static_cast<cudaStream_t>(stream)>>>(rows, classes, mode, in, reinterpret_cast<long long const*>(target), row_loss, dlogits);
    return cuda_check(cudaGetLastError(), "softmax loss");
}

GB_API int gb_batchnorm_forward(float const* x, float const* gamma, float const* beta, float* y, float* save_mean, float* save_invstd,
                                float* running_mean, float* running_var, int n, int c, int hw, float eps, float momentum, int relu, void* stream) {
    bn_fwd_kernel<<<c, kNnThreads, 0, static_cast<cudaStream_t>(stream)>>>(n, c, hw, x, gamma, beta, eps, momentum, y, save_mean, save_invstd, running_mean, running_var, relu);
    return cuda_check(cudaGetLastError(), "batchnorm forward");
}

GB_API int gb_batchnorm_backward(float const* x, float const* dy, float const* gamma, float const* save_mean, float const* save_invstd, float* dx,
                                 float* dgamma, float* dbeta, int n, int c, int hw, void* stream) {
    bn_bwd_kernel<<<c, kNnThreads, 0, static_cast<cudaStream_t>(stream)>>>(n, c, hw, x, dy, gamma, save_mean, save_invstd, dx, dgamma, dbeta);
    return cuda_check(cudaGetLastError(), "batchnorm backward");
}


GB_API int gb_global_avgpool_forward(float const* x, float* y, int64_t planes, int hw, void* stream) {
    if (planes < 1 || hw < 1) return cuda_fail("avgpool: empty input");
    avgpool_fwd_kernel<<<static_cast<int>((planes + 7) / 8), kNnThreads, 0, static_cast<cudaStream_t>(stream)>>>(planes, hw, x, y);
    return cuda_check(cudaGetLastError(), "avgpool forward");
}

GB_API int gb_global_avgpool_backward(float const* dy, float* dx, int64_t planes, int hw, void* stream) {
    long long const total = static_cast<long long>(planes) * hw;
    avgpool_bwd_kernel<<<nn_grid(total), kNnThreads, 0, static_cast<cudaStream_t>(stream)>>>(total, hw, dy, dx);
    return cuda_check(cudaGetLastError(), "avgpool backward");
}

// y = dropout(x) with keep probability 1-p; calling it again on dy with the same seed gives dx.
GB_API int gb_dropout(float const* x, float* y, int64_t total, float p, uint64_t seed, void* stream) {
    if (p < 0.0f || p >= 1.0f) return cuda_fail("dropout: need 0 <= p < 1");
    dropout_kernel<<<nn_grid((total + 3) / 4), kNnThreads, 0, static_cast<cudaStream_t>(stream)>>>(total, p, seed, x, y);
    return cuda_check(cudaGetLastError(), "dropout");
}

GB_API int gb_bce_loss(float const* prob, float const* target, float* loss, float* dprob, int64_t total, void* stream) {
    if (total < 1) return cuda_fail("bce: empty input");
    bce_kernel<<<nn_grid(total), kNnThreads, 0, static_cast<cudaStream_t>(stream)>>>(total, prob, target, loss, dprob);
    return cuda_check(cudaGetLastError(), "bce loss");
}

}  // extern "C"
