// Channels-last helper kernels around the tensor-core convolutions (csrc/cuda/conv_tc.cu): the few layers of a
// CNN that are not GEMMs, written so that a whole forward/backward pass runs on this library's kernels only.
//
//   im2col_pad     patch matrix for layers the implicit-GEMM loads cannot address (Cin not a multiple of 4, e.g. the
//                  RGB stem): cols[pixel][(kh, kw, c)] zero-padded to a row length that is a multiple of 4 floats
//   pad_rows       [rows, k] → [rows, kpad] zero-padded copy (weights of such layers)
//   col2im         scatter-free inverse of im2col (data gradient of such layers): every input pixel sums its patches
//   upsample_zero  out[n, h·s, w·s, :] = in[n, h, w, :], zero elsewhere (data gradient of strided convolutions)
//   maxpool        NHWC max pooling with the arg-max tap kept in one byte per output; deterministic gather backward
//   softmax_ce     grouped cross-entropy: per-worker mean losses and d(Σ_g mean_g)/d(logits) in one launch
//   colsum         per-worker column sums (bias gradients)
//   avgpool        NHWC global average pooling forward / backward
//
// Reference counterpart: ATen/cuDNN kernels behind nn.MaxPool2d, nn.CrossEntropyLoss, nn.AdaptiveAvgPool2d and the
// strided-convolution backward (garfieldpp/tools.py:47-57,69-75).

#include <cfloat>

#include "garfield_b200.h"
#include "launch.cuh"
#include "pdl.cuh"

namespace gb {
namespace {

inline unsigned grid_for(long long total, int block = 256) {
    long long blocks = (total + block - 1) / block;
    long long const cap = 32LL * device_sm_count();
    if (blocks > cap) blocks = cap;
    return static_cast<unsigned>(blocks < 1 ? 1 : blocks);
}

// One block = kIm2colPixels consecutive output pixels × all kpad columns. The pixel → (image, row, column) decomposition is done
// once per pixel (first warp, shared memory), the column → (kh, kw, c) decomposition once per thread and column; the inner loop
// is one predicated load and one coalesced store per element (the first version redid both decompositions per element and
// ran at 0.46 TB/s on the stem's 30 MB patch matrix).
constexpr int kIm2colPixels = 32;
__global__ void __launch_bounds__(256) im2col_pad_kernel(const float* __restrict__ x, float* __restrict__ cols, int H, int W, int C, int pitch, int KH,
                                                          int KW, int sh, int sw, int ph, int pw, int Ho, int Wo, int kpad, long long pixels) {
    pdl_enter();
    __shared__ long long base[kIm2colPixels];     // offset of pixel (n, 0, 0) ... in floats: n·H·W·pitch
    __shared__ int hb[kIm2colPixels], wb[kIm2colPixels];
    long long const pix0 = static_cast<long long>(blockIdx.x) * kIm2colPixels;
    if (threadIdx.x < kIm2colPixels) {
        long long const r = pix0 + threadIdx.x;
        if (r < pixels) {
            int const wo = static_cast<int>(r % Wo);
            long long const t = r / Wo;
            int const ho = static_cast<int>(t % Ho);
            base[threadIdx.x] = (t / Ho) * H * W * pitch;
            hb[threadIdx.x] = ho * sh - ph;
            wb[threadIdx.x] = wo * sw - pw;
        }
    }
    __syncthreads();
    int const k_real = KH * KW * C;
    int const count = static_cast<int>(min(static_cast<long long>(kIm2colPixels), pixels - pix0));
    for (int k = threadIdx.x; k < kpad; k += blockDim.x) {
        bool const real = k < k_real;
        int const c = k % C, kw = (k / C) % KW, kh = k / (C * KW);
        for (int p = 0; p < count; ++p) {
            int const h = hb[p] + kh, w = wb[p] + kw;
            float v = 0.f;
            if (real && h >= 0 && h < H && w >= 0 && w < W) v = x[base[p] + (static_cast<long long>(h) * W + w) * pitch + c];
            cols[(pix0 + p) * kpad + k] = v;
        }
    }
}

__global__ void __launch_bounds__(256) pad_rows_kernel(const float* __restrict__ src, float* __restrict__ dst, int k, long long ld, int kpad,
                                                        long long total) {
    pdl_enter();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % kpad);
        long long const r = i / kpad;
        dst[i] = c < k ? src[r * ld + c] : 0.f;
    }
}

// dx[n, h, w, c] = Σ_{(kh, kw): (h + ph − kh) % sh == 0, …} dcols[(n, ho, wo)][(kh, kw, c)]
__global__ void __launch_bounds__(256) col2im_kernel(const float* __restrict__ dcols, float* __restrict__ dx, int H, int W, int C, int pitch, int KH,
                                                      int KW, int sh, int sw, int ph, int pw, int Ho, int Wo, int kpad, long long total) {
    pdl_enter();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % C);
        long long r = i / C;
        int const w = static_cast<int>(r % W); r /= W;
        int const h = static_cast<int>(r % H);
        long long const n = r / H;
        float acc = 0.f;
        for (int kh = 0; kh < KH; ++kh) {
            int const hn = h + ph - kh;
            if (hn < 0 || hn % sh) continue;
            int const ho = hn / sh;
            if (ho >= Ho) continue;
            for (int kw = 0; kw < KW; ++kw) {
                int const wn = w + pw - kw;
                if (wn < 0 || wn % sw) continue;
                int const wo = wn / sw;
                if (wo >= Wo) continue;
                acc += dcols[((n * Ho + ho) * Wo + wo) * kpad + (kh * KW + kw) * C + c];
            }
        }
        dx[((n * H + h) * W + w) * pitch + c] = acc;
    }
}

__global__ void __launch_bounds__(256) upsample_zero_kernel(const float4* __restrict__ in, float4* __restrict__ out, int Ho, int Wo, int C4, int sh,
                                                             int sw, int H, int W, long long total) {
    pdl_enter();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % C4);
        long long r = i / C4;
        int const w = static_cast<int>(r % W); r /= W;
        int const h = static_cast<int>(r % H);
        long long const n = r / H;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (h % sh == 0 && w % sw == 0 && h / sh < Ho && w / sw < Wo) v = in[((n * Ho + h / sh) * Wo + w / sw) * C4 + c];
        out[i] = v;
    }
}

__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const float4* __restrict__ x, float4* __restrict__ y, uchar4* __restrict__ arg, int H, int W,
                                                           int C4, int K, int s, int p, int Ho, int Wo, long long total) {
    pdl_enter();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % C4);
        long long r = i / C4;
        int const wo = static_cast<int>(r % Wo); r /= Wo;
        int const ho = static_cast<int>(r % Ho);
        long long const n = r / Ho;
        float4 best = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX);
        uchar4 at = make_uchar4(255, 255, 255, 255);
        for (int kh = 0; kh < K; ++kh) {
            int const h = ho * s - p + kh;
            if (h < 0 || h >= H) continue;
            for (int kw = 0; kw < K; ++kw) {
                int const w = wo * s - p + kw;
                if (w < 0 || w >= W) continue;
                float4 const v = x[((n * H + h) * W + w) * C4 + c];
                unsigned char const tap = static_cast<unsigned char>(kh * K + kw);
                // first maximum in scan order wins and NaN propagates, like ATen
                if (v.x > best.x || v.x != v.x || at.x == 255) { best.x = v.x; at.x = tap; }
                if (v.y > best.y || v.y != v.y || at.y == 255) { best.y = v.y; at.y = tap; }
                if (v.z > best.z || v.z != v.z || at.z == 255) { best.z = v.z; at.z = tap; }
                if (v.w > best.w || v.w != v.w || at.w == 255) { best.w = v.w; at.w = tap; }
            }
        }
        y[i] = best;
        arg[i] = at;
    }
}

__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const float4* __restrict__ dy, const uchar4* __restrict__ arg, float4* __restrict__ dx, int H,
                                                           int W, int C4, int K, int s, int p, int Ho, int Wo, long long total) {
    pdl_enter();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % C4);
        long long r = i / C4;
        int const w = static_cast<int>(r % W); r /= W;
        int const h = static_cast<int>(r % H);
        long long const n = r / H;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int kh = 0; kh < K; ++kh) {
            int const hn = h + p - kh;
            if (hn < 0 || hn % s) continue;
            int const ho = hn / s;
            if (ho >= Ho) continue;
            for (int kw = 0; kw < K; ++kw) {
                int const wn = w + p - kw;
                if (wn < 0 || wn % s) continue;
                int const wo = wn / s;
                if (wo >= Wo) continue;
                long long const o = ((n * Ho + ho) * Wo + wo) * C4 + c;
                uchar4 const a = arg[o];
                float4 const g = dy[o];
                unsigned char const tap = static_cast<unsigned char>(kh * K + kw);
                if (a.x == tap) acc.x += g.x;
                if (a.y == tap) acc.y += g.y;
                if (a.z == tap) acc.z += g.z;
                if (a.w == tap) acc.w += g.w;
            }
        }
        dx[i] = acc;
    }
}

// One warp per row: losses[g] += row loss / rows_per_group (pre-zeroed by the host wrapper through `zero_first`),
// dlogits[row][c] = (softmax − onehot) / rows_per_group, padded columns zeroed.
__global__ void __launch_bounds__(256) softmax_ce_grouped_kernel(const float* __restrict__ logits, int ld, const long long* __restrict__ target,
                                                                  float* __restrict__ row_loss, float* __restrict__ dlogits, int ldd, int rows,
                                                                  int classes, int rows_per_group) {
    pdl_enter();
    int const warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= rows) return;
    const float* in = logits + static_cast<long long>(warp) * ld;
    float mx = -FLT_MAX;
    for (int c = lane; c < classes; c += 32) mx = fmaxf(mx, in[c]);
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    for (int c = lane; c < classes; c += 32) sum += expf(in[c] - mx);
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    int const t = static_cast<int>(target[warp]);
    float const inv = 1.f / static_cast<float>(rows_per_group), lse = logf(sum) + mx;
    if (lane == 0) row_loss[warp] = lse - in[t];
    float* out = dlogits + static_cast<long long>(warp) * ldd;
    for (int c = lane; c < ldd; c += 32) out[c] = c < classes ? (expf(in[c] - lse) - (c == t ? 1.f : 0.f)) * inv : 0.f;
}

// out[g][c] = scale · Σ_{rows of group g} in[row][c]; one block per (group, 32-column slab), fixed-order reduction.
__global__ void __launch_bounds__(256) colsum_grouped_kernel(const float* __restrict__ in, long long ld, float* __restrict__ out, int rows_per_group,
                                                              int cols, float scale) {
    pdl_enter();
    __shared__ float part[8][33];
    int const g = blockIdx.y, c = blockIdx.x * 32 + (threadIdx.x & 31), sub = threadIdx.x >> 5;
    float acc = 0.f;
    if (c < cols)
        for (int r = sub; r < rows_per_group; r += 8) acc += in[(static_cast<long long>(g) * rows_per_group + r) * ld + c];
    part[sub][threadIdx.x & 31] = acc;
    __syncthreads();
    if (sub == 0 && c < cols) {
        float s = 0.f;
        for (int k = 0; k < 8; ++k) s += part[k][threadIdx.x];
        out[static_cast<long long>(g) * cols + c] = s * scale;
    }
}

// ReLU on a flat tensor (forward) and its gradient mask from the OUTPUT (backward): y > 0 ⇔ x > 0.
// out = a + b over `total` floats (float4 body, scalar tail): the sum of the two gradients that meet where a residual block's
// input forks — autograd's own accumulation is a library kernel that takes no part in the programmatic launch chain.
__global__ void __launch_bounds__(256) add_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, long long total,
                                                  int vec) {
    pdl_enter();
    long long const step = static_cast<long long>(gridDim.x) * blockDim.x;
    long long const quads = vec ? total / 4 : 0;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < quads; i += step) {
        float4 const x = reinterpret_cast<const float4*>(a)[i], y = reinterpret_cast<const float4*>(b)[i];
        reinterpret_cast<float4*>(out)[i] = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
    }
    for (long long i = quads * 4 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += step) out[i] = a[i] + b[i];
}

__global__ void __launch_bounds__(256) relu_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long total) {
    pdl_enter();
    long long const n4 = total / 4;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float4 v = reinterpret_cast<const float4*>(x)[i];
        v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
        reinterpret_cast<float4*>(y)[i] = v;
    }
    for (long long i = n4 * 4 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x)
        y[i] = fmaxf(x[i], 0.f);
}
__global__ void __launch_bounds__(256) relu_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ dx, long long total) {
    pdl_enter();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x)
        dx[i] = y[i] > 0.f ? dy[i] : 0.f;
}

// Max pooling for any channel count (scalar lanes): the float4 kernels above need C % 4 == 0.
__global__ void __launch_bounds__(256) maxpool_fwd_scalar_kernel(const float* __restrict__ x, float* __restrict__ y, unsigned char* __restrict__ arg, int H,
                                                                  int W, int C, int K, int s, int p, int Ho, int Wo, long long total) {
    pdl_enter();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % C);
        long long r = i / C;
        int const wo = static_cast<int>(r % Wo); r /= Wo;
        int const ho = static_cast<int>(r % Ho);
        long long const n = r / Ho;
        float best = -FLT_MAX;
        unsigned char at = 255;
        for (int kh = 0; kh < K; ++kh) {
            int const h = ho * s - p + kh;
            if (h < 0 || h >= H) continue;
            for (int kw = 0; kw < K; ++kw) {
                int const w = wo * s - p + kw;
                if (w < 0 || w >= W) continue;
                float const v = x[((n * H + h) * W + w) * C + c];
                if (v > best || v != v || at == 255) { best = v; at = static_cast<unsigned char>(kh * K + kw); }
            }
        }
        y[i] = best;
        arg[i] = at;
    }
}
__global__ void __launch_bounds__(256) maxpool_bwd_scalar_kernel(const float* __restrict__ dy, const unsigned char* __restrict__ arg, float* __restrict__ dx,
                                                                  int H, int W, int C, int K, int s, int p, int Ho, int Wo, long long total) {
    pdl_enter();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % C);
        long long r = i / C;
        int const w = static_cast<int>(r % W); r /= W;
        int const h = static_cast<int>(r % H);
        long long const n = r / H;
        float acc = 0.f;
        for (int kh = 0; kh < K; ++kh) {
            int const hn = h + p - kh;
            if (hn < 0 || hn % s) continue;
            int const ho = hn / s;
            if (ho >= Ho) continue;
            for (int kw = 0; kw < K; ++kw) {
                int const wn = w + p - kw;
                if (wn < 0 || wn % s) continue;
                int const wo = wn / s;
                if (wo >= Wo) continue;
                long long const o = ((n * Ho + ho) * Wo + wo) * C + c;
                if (arg[o] == static_cast<unsigned char>(kh * K + kw)) acc += dy[o];
            }
        }
        dx[i] = acc;
    }
}

__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int hw, int C, long long total) {
    pdl_enter();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % C);
        long long const n = i / C;
        float acc = 0.f;
        for (int p = 0; p < hw; ++p) acc += x[(n * hw + p) * C + c];
        y[i] = acc / static_cast<float>(hw);
    }
}
__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int hw, int C, long long total) {
    pdl_enter();
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % C);
        long long const n = i / C / hw;
        dx[i] = dy[n * C + c] / static_cast<float>(hw);
    }
}

}  // namespace
}  // namespace gb

using namespace gb;

extern "C" GB_API int gb_im2col_pad_nhwc(float const* x, float* cols, int N, int H, int W, int C, int pitch, int KH, int KW, int stride_h,
                                         int stride_w, int pad_h, int pad_w, int kpad, void* stream) {
    if (N < 1 || H < 1 || W < 1 || C < 1 || pitch < C || KH < 1 || KW < 1 || stride_h < 1 || stride_w < 1 || pad_h < 0 || pad_w < 0 || kpad < KH * KW * C)
        return cuda_fail("im2col_pad_nhwc: bad geometry");
    int const Ho = (H + 2 * pad_h - KH) / stride_h + 1, Wo = (W + 2 * pad_w - KW) / stride_w + 1;
    if (Ho < 1 || Wo < 1) return cuda_fail("im2col_pad_nhwc: empty output");
    long long const pixels = static_cast<long long>(N) * Ho * Wo;
    long long const blocks = (pixels + kIm2colPixels - 1) / kIm2colPixels;
    if (blocks > 2147483647LL) return cuda_fail("im2col_pad_nhwc: too many output pixels");
    launch_pdl(im2col_pad_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), static_cast<cudaStream_t>(stream), x, cols, H, W, C, pitch, KH, KW, stride_h, stride_w,
                                                                                                   pad_h, pad_w, Ho, Wo, kpad, pixels);
    return cuda_check(cudaGetLastError(), "im2col_pad_nhwc");
}

extern "C" GB_API int gb_pad_rows(float const* src, float* dst, int64_t rows, int k, int64_t ld, int kpad, void* stream) {
    if (rows < 1 || k < 1 || kpad < k || ld < k) return cuda_fail("pad_rows: bad shape");
    long long const total = rows * kpad;
    launch_pdl(pad_rows_kernel, dim3(grid_for(total)), dim3(256), static_cast<cudaStream_t>(stream), src, dst, k, ld, kpad, total);
    return cuda_check(cudaGetLastError(), "pad_rows");
}

extern "C" GB_API int gb_col2im_nhwc(float const* dcols, float* dx, int N, int H, int W, int C, int pitch, int KH, int KW, int stride_h, int stride_w,
                                     int pad_h, int pad_w, int kpad, void* stream) {
    if (N < 1 || H < 1 || W < 1 || C < 1 || pitch < C || KH < 1 || KW < 1 || stride_h < 1 || stride_w < 1 || kpad < KH * KW * C)
        return cuda_fail("col2im_nhwc: bad geometry");
    int const Ho = (H + 2 * pad_h - KH) / stride_h + 1, Wo = (W + 2 * pad_w - KW) / stride_w + 1;
    long long const total = static_cast<long long>(N) * H * W * C;
    launch_pdl(col2im_kernel, dim3(grid_for(total)), dim3(256), static_cast<cudaStream_t>(stream), dcols, dx, H, W, C, pitch, KH, KW, stride_h, stride_w, pad_h, pad_w, Ho,
                                                                                  Wo, kpad, total);
    return cuda_check(cudaGetLastError(), "col2im_nhwc");
}

extern "C" GB_API int gb_upsample_zero_nhwc(float const* in, float* out, int N, int Ho, int Wo, int C, int stride_h, int stride_w, int H, int W,
                                            void* stream) {
    if (N < 1 || Ho < 1 || Wo < 1 || C < 4 || C % 4 || stride_h < 1 || stride_w < 1 || H < 1 || W < 1) return cuda_fail("upsample_zero_nhwc: bad geometry");
    long long const total = static_cast<long long>(N) * H * W * (C / 4);
    launch_pdl(upsample_zero_kernel, dim3(grid_for(total)), dim3(256), static_cast<cudaStream_t>(stream), reinterpret_cast<const float4*>(in), reinterpret_cast<float4*>(out),
                                                                                         Ho, Wo, C / 4, stride_h, stride_w, H, W, total);
    return cuda_check(cudaGetLastError(), "upsample_zero_nhwc");
}

extern "C" GB_API int gb_maxpool_nhwc_forward(float const* x, float* y, unsigned char* arg, int N, int H, int W, int C, int k, int stride, int pad,
                                              void* stream) {
    if (N < 1 || H < 1 || W < 1 || C < 1 || k < 1 || k > 15 || stride < 1 || pad < 0 || 2 * pad > k) return cuda_fail("maxpool_nhwc: bad geometry");
    int const Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
    if (C % 4) {
        long long const all = static_cast<long long>(N) * Ho * Wo * C;
        launch_pdl(maxpool_fwd_scalar_kernel, dim3(grid_for(all)), dim3(256), static_cast<cudaStream_t>(stream), x, y, arg, H, W, C, k, stride, pad, Ho, Wo, all);
        return cuda_check(cudaGetLastError(), "maxpool_nhwc_forward");
    }
    long long const total = static_cast<long long>(N) * Ho * Wo * (C / 4);
    launch_pdl(maxpool_fwd_kernel, dim3(grid_for(total)), dim3(256), static_cast<cudaStream_t>(stream), reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(y),
                                                                                       reinterpret_cast<uchar4*>(arg), H, W, C / 4, k, stride, pad, Ho, Wo,
                                                                                       total);
    return cuda_check(cudaGetLastError(), "maxpool_nhwc_forward");
}

extern "C" GB_API int gb_maxpool_nhwc_backward(float const* dy, unsigned char const* arg, float* dx, int N, int H, int W, int C, int k, int stride,
                                               int pad, void* stream) {
    if (N < 1 || H < 1 || W < 1 || C < 1 || k < 1 || k > 15 || stride < 1 || pad < 0) return cuda_fail("maxpool_nhwc: bad geometry");
    int const Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
    if (C % 4) {
        long long const all = static_cast<long long>(N) * H * W * C;
        launch_pdl(maxpool_bwd_scalar_kernel, dim3(grid_for(all)), dim3(256), static_cast<cudaStream_t>(stream), dy, arg, dx, H, W, C, k, stride, pad, Ho, Wo, all);
        return cuda_check(cudaGetLastError(), "maxpool_nhwc_backward");
    }
    long long const total = static_cast<long long>(N) * H * W * (C / 4);
    launch_pdl(maxpool_bwd_kernel, dim3(grid_for(total)), dim3(256), static_cast<cudaStream_t>(stream), reinterpret_cast<const float4*>(dy),
                                                                                       reinterpret_cast<const uchar4*>(arg), reinterpret_cast<float4*>(dx),
                                                                                       H, W, C / 4, k, stride, pad, Ho, Wo, total);
    return cuda_check(cudaGetLastError(), "maxpool_nhwc_backward");
}

extern "C" GB_API int gb_softmax_ce_grouped(float const* logits, int ld, int64_t const* target, float* row_loss, float* dlogits, int ldd, int rows,
                                            int classes, int rows_per_group, void* stream) {
    if (rows < 1 || classes < 1 || ld < classes || ldd < classes || rows_per_group < 1) return cuda_fail("softmax_ce_grouped: bad shape");
    int const blocks = (rows * 32 + 255) / 256;
    launch_pdl(softmax_ce_grouped_kernel, dim3(blocks), dim3(256), static_cast<cudaStream_t>(stream), logits, ld, reinterpret_cast<const long long*>(target), row_loss,
                                                                                     dlogits, ldd, rows, classes, rows_per_group);
    return cuda_check(cudaGetLastError(), "softmax_ce_grouped");
}

extern "C" GB_API int gb_colsum_grouped(float const* in, int64_t ld, float* out, int groups, int rows_per_group, int cols, float scale, void* stream) {
    if (groups < 1 || rows_per_group < 1 || cols < 1 || ld < cols) return cuda_fail("colsum_grouped: bad shape");
    dim3 const grid(static_cast<unsigned>((cols + 31) / 32), static_cast<unsigned>(groups));
    launch_pdl(colsum_grouped_kernel, dim3(grid), dim3(256), static_cast<cudaStream_t>(stream), in, ld, out, rows_per_group, cols, scale);
    return cuda_check(cudaGetLastError(), "colsum_grouped");
}

extern "C" GB_API int gb_add(float const* a, float const* b, float* out, int64_t total, void* stream) {
    using namespace gb;
    if (total < 1) return 0;
    int const vec = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0 ? 1 : 0;
    launch_pdl(add_kernel, dim3(grid_for(total / 4 + 1)), dim3(256), static_cast<cudaStream_t>(stream), a, b, out, static_cast<long long>(total), vec);
    return cuda_check(cudaGetLastError(), "add");
}

extern "C" GB_API int gb_relu_forward(float const* x, float* y, int64_t total, void* stream) {
    if (total < 1) return cuda_fail("relu: empty tensor");
    if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15u) return cuda_fail("relu: tensors must be 16-byte aligned");
    launch_pdl(relu_fwd_kernel, dim3(grid_for(total / 4 + 1)), dim3(256), static_cast<cudaStream_t>(stream), x, y, total);
    return cuda_check(cudaGetLastError(), "relu_forward");
}
extern "C" GB_API int gb_relu_backward(float const* dy, float const* y, float* dx, int64_t total, void* stream) {
    if (total < 1) return cuda_fail("relu: empty tensor");
    launch_pdl(relu_bwd_kernel, dim3(grid_for(total)), dim3(256), static_cast<cudaStream_t>(stream), dy, y, dx, total);
    return cuda_check(cudaGetLastError(), "relu_backward");
}

extern "C" GB_API int gb_avgpool_nhwc_forward(float const* x, float* y, int64_t n, int hw, int C, void* stream) {
    if (n < 1 || hw < 1 || C < 1) return cuda_fail("avgpool_nhwc: bad shape");
    long long const total = n * C;
    launch_pdl(avgpool_fwd_kernel, dim3(grid_for(total)), dim3(256), static_cast<cudaStream_t>(stream), x, y, hw, C, total);
    return cuda_check(cudaGetLastError(), "avgpool_nhwc_forward");
}
extern "C" GB_API int gb_avgpool_nhwc_backward(float const* dy, float* dx, int64_t n, int hw, int C, void* stream) {
    if (n < 1 || hw < 1 || C < 1) return cuda_fail("avgpool_nhwc: bad shape");
    long long const total = n * hw * C;
    launch_pdl(avgpool_bwd_kernel, dim3(grid_for(total)), dim3(256), static_cast<cudaStream_t>(stream), dy, dx, hw, C, total);
    return cuda_check(cudaGetLastError(), "avgpool_nhwc_backward");
}
