// Fused training-mode BatchNorm for channels-last activations: y = [ReLU]( BN(x) [+ residual] ), and its
// backward, one launch each. The flagship ResNet-50 step spends ≈30 % of one worker's forward+backward chain
// in 106 cuDNN batch-norm launches plus 98 separate ReLU / ReLU-backward launches and 32 residual adds
// (profiles/README.md); on the 32×32-input shapes these tensors are 50 KB – 13 MB, i.e. L2-resident.
//
// Layout: x is [groups, rows, C] (rows = N·H·W of one statistics group, C contiguous = NHWC). The work is cut
// three ways so that even a 64-channel layer fills the GPU:
//   * channel slabs   — a CTA owns 4·QUADS channels (QUADS ∈ {1,2,4,8}); thread = (channel quad, row lane), float4 loads;
//   * row splits      — the rows of one slab are divided over the CTAs of a THREAD-BLOCK CLUSTER (≤ 8); each CTA
//                       reduces its rows, the partial sums are exchanged through distributed shared memory
//                       (fixed order → deterministic), and each CTA then normalises its own rows;
//   * groups          — gridDim.z independent statistics groups (the batches of co-located workers).
// When a thread's share of rows is ≤ 8 float4 the values stay in REGISTERS between the statistics pass and the
// normalisation pass (x is read exactly once); larger shares are re-read (L1/L2 hits).
// One-pass statistics around a per-channel pivot (the group's first row), so Σ(x−K)² has no E[x²]−E[x]² cancellation.
// Reference counterpart: cuDNN/ATen BatchNorm2d + ReLU + residual add (torchvision resnet.py via
// garfieldpp/tools.py:69-75).

#include <cooperative_groups.h>

#include <cmath>
#include <cstdlib>

#include "garfield_b200.h"
#include "launch.cuh"

namespace cg = cooperative_groups;

namespace gb {
namespace {

constexpr int kBnThreads = 256;
constexpr int kKeep = 8;          // float4 values per thread that may stay in registers between the two passes
constexpr int kMaxSplits = 8;     // portable cluster size

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 add4(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 zero4() { return make_float4(0.f, 0.f, 0.f, 0.f); }

// Sum two float4 accumulators over all row lanes of the CTA (one barrier pair for both); results valid in every
// thread of that quad.
template <int QUADS>
__device__ __forceinline__ void reduce_rows(float4& a, float4& b, float4 (*scratch)[2][QUADS]) {
#pragma unroll
    for (int o = QUADS; o < 32; o <<= 1) {
        a.x += __shfl_xor_sync(0xffffffffu, a.x, o); a.y += __shfl_xor_sync(0xffffffffu, a.y, o);
        a.z += __shfl_xor_sync(0xffffffffu, a.z, o); a.w += __shfl_xor_sync(0xffffffffu, a.w, o);
        b.x += __shfl_xor_sync(0xffffffffu, b.x, o); b.y += __shfl_xor_sync(0xffffffffu, b.y, o);
        b.z += __shfl_xor_sync(0xffffffffu, b.z, o); b.w += __shfl_xor_sync(0xffffffffu, b.w, o);
    }
    int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31, quad = lane % QUADS;
    if (lane < QUADS) { scratch[warp][0][quad] = a; scratch[warp][1][quad] = b; }
    __syncthreads();
    float4 ta = zero4(), tb = zero4();
#pragma unroll
    for (int w = 0; w < kBnThreads / 32; ++w) {  // fixed order
        ta = add4(ta, scratch[w][0][quad]);
        tb = add4(tb, scratch[w][1][quad]);
    }
    a = ta; b = tb;
}

// Combine the two per-CTA partial sums over the row splits of the cluster (rank order → every CTA gets the same bits).
template <int QUADS>
__device__ __forceinline__ void reduce_cluster(float4& a, float4& b, float4 (*part)[QUADS], int splits, int quad) {
    if (splits == 1) return;
    cg::cluster_group cluster = cg::this_cluster();
    if (threadIdx.x < QUADS) { part[0][threadIdx.x] = a; part[1][threadIdx.x] = b; }   // threads 0..QUADS-1 hold quad = threadIdx.x
    cluster.sync();
    float4 ta = zero4(), tb = zero4();
    for (int r = 0; r < splits; ++r) {
        float4 const* peer = reinterpret_cast<float4 const*>(cluster.map_shared_rank(&part[0][0], r));
        ta = add4(ta, peer[quad]);
        tb = add4(tb, peer[QUADS + quad]);
    }
    cluster.sync();   // nobody leaves (or overwrites part) while a peer still reads it
    a = ta; b = tb;
}

struct BnGeom {
    int r0, r1;          // this CTA's rows
    int c0;              // first channel of this thread's quad
    int quad, rlane;
    bool live;
};

template <int QUADS>
__device__ __forceinline__ BnGeom bn_geometry(int rows, int C, int splits) {
    BnGeom g;
    int const lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    g.quad = lane % QUADS;
    g.rlane = lane / QUADS + (32 / QUADS) * warp;
    g.c0 = (blockIdx.y * QUADS + g.quad) * 4;
    g.live = g.c0 < C;                                                   // C is a multiple of 4
    int const chunk = (rows + splits - 1) / splits;
    g.r0 = min(rows, static_cast<int>(blockIdx.x) * chunk);
    g.r1 = min(rows, g.r0 + chunk);
    return g;
}

template <int QUADS, bool KEEP>
__global__ void __launch_bounds__(kBnThreads) bn_nhwc_fwd_kernel(int rows, int C, int splits, const float* __restrict__ x,
                                                                 const float* __restrict__ residual, const float* __restrict__ gamma,
                                                                 const float* __restrict__ beta, float eps, float momentum, int relu,
                                                                 float* __restrict__ y, float* __restrict__ save_mean, float* __restrict__ save_invstd,
                                                                 float* __restrict__ running_mean, float* __restrict__ running_var) {
    __shared__ float4 scratch[kBnThreads / 32][2][QUADS];
    __shared__ float4 part[2][QUADS];
    constexpr int kLanes = kBnThreads / QUADS;                          // row lanes per CTA
    BnGeom const t = bn_geometry<QUADS>(rows, C, splits);
    cudaGridDependencySynchronize();   // no-op unless launched with programmatic stream serialization
    cudaTriggerProgrammaticLaunchCompletion();   // the next kernel may start its prologue
    {   // blockIdx.z = statistics group: a contiguous block of `rows` rows with its own mean/variance/running buffers
        long long const goff = static_cast<long long>(blockIdx.z) * rows * C;
        int const soff = blockIdx.z * C;
        x += goff; y += goff;
        if (residual != nullptr) residual += goff;
        save_mean += soff; save_invstd += soff;
        if (running_mean != nullptr) { running_mean += soff; running_var += soff; }
    }
    float4 const K = t.live ? ld4(x + t.c0) : zero4();                   // pivot: the group's first row (same in every split)
    // issued now, consumed after the reduction: their latency hides behind the statistics pass
    bool const writes_stats = t.live && t.rlane == 0 && blockIdx.x == 0;
    float4 const g = t.live ? ld4(gamma + t.c0) : zero4(), b = t.live ? ld4(beta + t.c0) : zero4();
    float4 rm = zero4(), rv = zero4();
    if (writes_stats && running_mean != nullptr) { rm = ld4(running_mean + t.c0); rv = ld4(running_var + t.c0); }
    float4 s = zero4(), q = zero4();
    float4 keep[KEEP ? kKeep : 1];
    if (t.live) {
        if (KEEP) {
#pragma unroll
            for (int i = 0; i < kKeep; ++i) {
                int const r = t.r0 + t.rlane + i * kLanes;
                keep[i] = r < t.r1 ? ld4(x + static_cast<long long>(r) * C + t.c0) : K;   // K contributes zero
            }
#pragma unroll
            for (int i = 0; i < kKeep; ++i) {
                float const a = keep[i].x - K.x, b = keep[i].y - K.y, c = keep[i].z - K.z, d = keep[i].w - K.w;
                s.x += a; s.y += b; s.z += c; s.w += d;
                q.x = fmaf(a, a, q.x); q.y = fmaf(b, b, q.y); q.z = fmaf(c, c, q.z); q.w = fmaf(d, d, q.w);
            }
        } else {
#pragma unroll 8
            for (int r = t.r0 + t.rlane; r < t.r1; r += kLanes) {
                float4 const v = ld4(x + static_cast<long long>(r) * C + t.c0);
                float const a = v.x - K.x, b = v.y - K.y, c = v.z - K.z, d = v.w - K.w;
                s.x += a; s.y += b; s.z += c; s.w += d;
                q.x = fmaf(a, a, q.x); q.y = fmaf(b, b, q.y); q.z = fmaf(c, c, q.z); q.w = fmaf(d, d, q.w);
            }
        }
    }
    reduce_rows<QUADS>(s, q, scratch);
    reduce_cluster<QUADS>(s, q, part, splits, t.quad);
    if (!t.live) return;
    float const inv_rows = 1.0f / static_cast<float>(rows);
    float4 const ds = make_float4(s.x * inv_rows, s.y * inv_rows, s.z * inv_rows, s.w * inv_rows);       // mean − K
    float4 const mean = add4(K, ds);
    float4 const var = make_float4(fmaxf(q.x * inv_rows - ds.x * ds.x, 0.f), fmaxf(q.y * inv_rows - ds.y * ds.y, 0.f),
                                   fmaxf(q.z * inv_rows - ds.z * ds.z, 0.f), fmaxf(q.w * inv_rows - ds.w * ds.w, 0.f));
    float4 const invstd = make_float4(rsqrtf(var.x + eps), rsqrtf(var.y + eps), rsqrtf(var.z + eps), rsqrtf(var.w + eps));
    float4 const scale = make_float4(g.x * invstd.x, g.y * invstd.y, g.z * invstd.z, g.w * invstd.w);
    float4 const shift = make_float4(b.x - mean.x * scale.x, b.y - mean.y * scale.y, b.z - mean.z * scale.z, b.w - mean.w * scale.w);
    auto emit = [&](int r, float4 v) {
        long long const at = static_cast<long long>(r) * C + t.c0;
        float4 o = make_float4(fmaf(v.x, scale.x, shift.x), fmaf(v.y, scale.y, shift.y), fmaf(v.z, scale.z, shift.z), fmaf(v.w, scale.w, shift.w));
        if (residual != nullptr) o = add4(o, ld4(residual + at));
        if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
        st4(y + at, o);
    };
    if (KEEP) {
#pragma unroll
        for (int i = 0; i < kKeep; ++i) {
            int const r = t.r0 + t.rlane + i * kLanes;
            if (r < t.r1) emit(r, keep[i]);
        }
    } else {
#pragma unroll 8
        for (int r = t.r0 + t.rlane; r < t.r1; r += kLanes) emit(r, ld4(x + static_cast<long long>(r) * C + t.c0));
    }
    if (writes_stats) {
        st4(save_mean + t.c0, mean);
        st4(save_invstd + t.c0, invstd);
        if (running_mean != nullptr) {
            float const unbias = rows > 1 ? static_cast<float>(rows) / static_cast<float>(rows - 1) : 1.0f;
            float const k = 1.0f - momentum;
            rm = make_float4(k * rm.x + momentum * mean.x, k * rm.y + momentum * mean.y, k * rm.z + momentum * mean.z, k * rm.w + momentum * mean.w);
            rv = make_float4(k * rv.x + momentum * var.x * unbias, k * rv.y + momentum * var.y * unbias, k * rv.z + momentum * var.z * unbias,
                             k * rv.w + momentum * var.w * unbias);
            st4(running_mean + t.c0, rm);
            st4(running_var + t.c0, rv);
        }
    }
}

__device__ __forceinline__ float4 relu_mask(float4 d, float4 o) {
    return make_float4(o.x > 0.f ? d.x : 0.f, o.y > 0.f ? d.y : 0.f, o.z > 0.f ? d.z : 0.f, o.w > 0.f ? d.w : 0.f);
}

// dy is the gradient w.r.t. the block output; with relu the upstream mask is (y > 0). Produces dx (through
// the normalisation), dgamma, dbeta, and — when d_residual != nullptr — the masked gradient for the skip path.
template <int QUADS, bool KEEP>
__global__ void __launch_bounds__(kBnThreads) bn_nhwc_bwd_kernel(int rows, int C, int splits, const float* __restrict__ x, const float* __restrict__ y,
                                                                 const float* __restrict__ dy, const float* __restrict__ gamma,
                                                                 const float* __restrict__ save_mean, const float* __restrict__ save_invstd, int relu,
                                                                 float* __restrict__ dx, float* __restrict__ d_residual, float* __restrict__ dgamma,
                                                                 float* __restrict__ dbeta) {
    __shared__ float4 scratch[kBnThreads / 32][2][QUADS];
    __shared__ float4 part[2][QUADS];
    constexpr int kLanes = kBnThreads / QUADS;
    BnGeom const t = bn_geometry<QUADS>(rows, C, splits);
    cudaGridDependencySynchronize();
    cudaTriggerProgrammaticLaunchCompletion();
    {
        long long const goff = static_cast<long long>(blockIdx.z) * rows * C;
        int const soff = blockIdx.z * C;
        x += goff; dy += goff; dx += goff;
        if (y != nullptr) y += goff;
        if (d_residual != nullptr) d_residual += goff;
        save_mean += soff; save_invstd += soff; dgamma += soff; dbeta += soff;
    }
    float4 mean = zero4(), invstd = zero4(), g = zero4();
    if (t.live) { mean = ld4(save_mean + t.c0); invstd = ld4(save_invstd + t.c0); g = ld4(gamma + t.c0); }
    float4 s1 = zero4(), s2 = zero4();
    float4 keep_d[KEEP ? kKeep : 1], keep_v[KEEP ? kKeep : 1];
    auto load = [&](int r, float4& d, float4& v) {
        long long const at = static_cast<long long>(r) * C + t.c0;
        d = ld4(dy + at);
        if (relu) d = relu_mask(d, ld4(y + at));
        v = ld4(x + at);
        v = make_float4((v.x - mean.x) * invstd.x, (v.y - mean.y) * invstd.y, (v.z - mean.z) * invstd.z, (v.w - mean.w) * invstd.w);   // x̂
    };
    if (t.live) {
        if (KEEP) {
#pragma unroll
            for (int i = 0; i < kKeep; ++i) {
                int const r = t.r0 + t.rlane + i * kLanes;
                keep_d[i] = zero4(); keep_v[i] = zero4();
                if (r < t.r1) load(r, keep_d[i], keep_v[i]);
            }
#pragma unroll
            for (int i = 0; i < kKeep; ++i) {
                s1 = add4(s1, keep_d[i]);
                s2.x = fmaf(keep_d[i].x, keep_v[i].x, s2.x); s2.y = fmaf(keep_d[i].y, keep_v[i].y, s2.y);
                s2.z = fmaf(keep_d[i].z, keep_v[i].z, s2.z); s2.w = fmaf(keep_d[i].w, keep_v[i].w, s2.w);
            }
        } else {
#pragma unroll 4
            for (int r = t.r0 + t.rlane; r < t.r1; r += kLanes) {
                float4 d, v;
                load(r, d, v);
                s1 = add4(s1, d);
                s2.x = fmaf(d.x, v.x, s2.x); s2.y = fmaf(d.y, v.y, s2.y); s2.z = fmaf(d.z, v.z, s2.z); s2.w = fmaf(d.w, v.w, s2.w);
            }
        }
    }
    reduce_rows<QUADS>(s1, s2, scratch);
    reduce_cluster<QUADS>(s1, s2, part, splits, t.quad);
    if (!t.live) return;
    float const inv_rows = 1.0f / static_cast<float>(rows);
    float4 const gi = make_float4(g.x * invstd.x, g.y * invstd.y, g.z * invstd.z, g.w * invstd.w);
    auto emit = [&](int r, float4 d, float4 v) {
        long long const at = static_cast<long long>(r) * C + t.c0;
        if (d_residual != nullptr) st4(d_residual + at, d);
        float4 o;
        o.x = gi.x * (d.x - inv_rows * (s1.x + v.x * s2.x));
        o.y = gi.y * (d.y - inv_rows * (s1.y + v.y * s2.y));
        o.z = gi.z * (d.z - inv_rows * (s1.z + v.z * s2.z));
        o.w = gi.w * (d.w - inv_rows * (s1.w + v.w * s2.w));
        st4(dx + at, o);
    };
    if (KEEP) {
#pragma unroll
        for (int i = 0; i < kKeep; ++i) {
            int const r = t.r0 + t.rlane + i * kLanes;
            if (r < t.r1) emit(r, keep_d[i], keep_v[i]);
        }
    } else {
#pragma unroll 4
        for (int r = t.r0 + t.rlane; r < t.r1; r += kLanes) {
            float4 d, v;
            load(r, d, v);
            emit(r, d, v);
        }
    }
    if (t.rlane == 0 && blockIdx.x == 0) {
        st4(dgamma + t.c0, s2);
        st4(dbeta + t.c0, s1);
    }
}

struct BnPlan {
    int quads, splits;
    bool keep;
    dim3 grid;
};

// One wave of fat CTAs: take the widest channel slab (128-byte rows) that still yields >= one CTA per SM from
// channels x groups alone, going down to 32-byte rows (one sector) before giving up; only when even that leaves SMs
// idle split the rows over a cluster (whose DSMEM exchange costs two cluster barriers), until every SM has a CTA or a
// CTA would have less than one row per lane (a split is only worth its barriers while half the SMs would otherwise idle).
inline BnPlan plan(int groups, int rows, int C) {
    BnPlan p;
    int const quads_total = C / 4;
    int const sms = device_sm_count();
    p.quads = quads_total >= 2 ? 2 : 1;
    for (int q = 8; q > 2; q >>= 1)
        if (static_cast<long long>((quads_total + q - 1) / q) * groups >= sms && quads_total % q == 0) { p.quads = q; break; }
    int const slabs = (quads_total + p.quads - 1) / p.quads;
    int const lanes = kBnThreads / p.quads;
    p.splits = 1;
    while (p.splits < kMaxSplits && static_cast<long long>(slabs) * groups * p.splits * 2 <= sms && rows / (p.splits * 2) >= lanes) p.splits *= 2;
    int const chunk = (rows + p.splits - 1) / p.splits;
    p.keep = (chunk + lanes - 1) / lanes <= kKeep;
    p.grid = dim3(p.splits, slabs, groups);
    return p;
}

inline bool pdl_enabled() {
    static int const on = [] { char const* v = std::getenv("GARFIELD_PDL"); return (v != nullptr && v[0] == '0') ? 0 : 1; }();
    return on != 0;
}

template <typename Kernel, typename... Args>
cudaError_t launch_bn(Kernel kernel, BnPlan const& p, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = p.grid;
    cfg.blockDim = dim3(kBnThreads);
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int count = 0;
    if (p.splits > 1) {
        attr[count].id = cudaLaunchAttributeClusterDimension;
        attr[count].val.clusterDim.x = p.splits;
        attr[count].val.clusterDim.y = 1;
        attr[count].val.clusterDim.z = 1;
        ++count;
    }
    if (pdl_enabled()) {   // programmatic dependent launch: CTAs are scheduled while the producer drains; the kernel
        attr[count].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // waits in cudaGridDependencySynchronize()
        attr[count].val.programmaticStreamSerializationAllowed = 1;
        ++count;
    }
    cfg.attrs = attr;
    cfg.numAttrs = count;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
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
    BnPlan const p = plan(groups, rows, C);
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    cudaError_t err;
#define GB_BN_FWD(Q, KP) err = launch_bn(bn_nhwc_fwd_kernel<Q, KP>, p, st, rows, C, p.splits, x, residual, gamma, beta, eps, momentum, relu, y, save_mean, save_invstd, running_mean, running_var)
#define GB_BN_FWD_Q(Q) do { if (p.keep) GB_BN_FWD(Q, true); else GB_BN_FWD(Q, false); } while (0)
    if (p.quads == 8) GB_BN_FWD_Q(8); else if (p.quads == 4) GB_BN_FWD_Q(4); else if (p.quads == 2) GB_BN_FWD_Q(2); else GB_BN_FWD_Q(1);
#undef GB_BN_FWD_Q
#undef GB_BN_FWD
    return cuda_check(err, "bn_nhwc forward");
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
    if (relu && y == nullptr) return cuda_fail("bn_nhwc backward: relu needs the forward output");
    BnPlan const p = plan(groups, rows, C);
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    cudaError_t err;
#define GB_BN_BWD(Q, KP) err = launch_bn(bn_nhwc_bwd_kernel<Q, KP>, p, st, rows, C, p.splits, x, y, dy, gamma, save_mean, save_invstd, relu, dx, d_residual, dgamma, dbeta)
#define GB_BN_BWD_Q(Q) do { if (p.keep) GB_BN_BWD(Q, true); else GB_BN_BWD(Q, false); } while (0)
    if (p.quads == 8) GB_BN_BWD_Q(8); else if (p.quads == 4) GB_BN_BWD_Q(4); else if (p.quads == 2) GB_BN_BWD_Q(2); else GB_BN_BWD_Q(1);
#undef GB_BN_BWD_Q
#undef GB_BN_BWD
    return cuda_check(err, "bn_nhwc backward");
}

GB_API int gb_bn_nhwc_backward(float const* x, float const* y, float const* dy, float const* gamma, float const* save_mean,
                               float const* save_invstd, float* dx, float* d_residual, float* dgamma, float* dbeta, int rows, int C, int relu,
                               void* stream) {
    return gb_bn_nhwc_backward_grouped(x, y, dy, gamma, save_mean, save_invstd, dx, d_residual, dgamma, dbeta, 1, rows, C, relu, stream);
}

}  // extern "C"
