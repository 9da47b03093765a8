// Flat-arena elementwise kernels: optimizers (torch.optim semantics on one d-length vector),
// Byzantine attack generation, and small utilities (scale / axpy / sum of squares).
//
// Counterparts in the reference are ATen ops executed on the HOST by the parameter server:
//   optimizer step   /root/reference/pytorch_impl/libs/garfieldpp/server.py:277-287 (+ tools.py:107-123)
//   attacks          /root/reference/pytorch_impl/libs/garfieldpp/byzWorker.py:78-143, byzServer.py:86-107
// Here they run on the flat device arenas with 128-bit accesses; the SGD step additionally
// exists fused into the aggregation epilogue (gar_coordwise.cu).

#include <cmath>

#include "gar_common.cuh"
#include "ptx.cuh"
#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

constexpr int kEwThreads = 256;

inline int ew_grid(long long work_items) {
    long long blocks = (work_items + kEwThreads - 1) / kEwThreads;
    long long const cap = static_cast<long long>(device_sm_count()) * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

// Generic grid-stride map over d elements, 4 at a time when the base pointers allow it.
template <class F>
__global__ void __launch_bounds__(kEwThreads) map_kernel(long long d, F f) {
    long long const stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < d; i += stride) f(i);
}

template <class F>
int launch_map(long long d, cudaStream_t stream, F f, char const* what) {
    map_kernel<<<ew_grid(d), kEwThreads, 0, stream>>>(d, f);
    return cuda_check(cudaGetLastError(), what);
}

struct Mirrors {
    float* p[kMaxMirrors];
    int n;
};

}  // namespace
}  // namespace gb

using namespace gb;

extern "C" {

// ---------------------------------------------------------------------------- optimizers ---

GB_API int gb_sgd_step(float* param, float const* grad, float* mom, int64_t d, float lr, float momentum, float dampening,
                       float weight_decay, int nesterov, int first_step, float* const* mirrors, int n_mirrors, void* stream) {
    if (n_mirrors < 0 || n_mirrors > kMaxMirrors) return cuda_fail("sgd: too many mirrors");
    if (momentum != 0.0f && mom == nullptr) return cuda_fail("sgd: momentum buffer missing");
    Mirrors mir{};
    mir.n = n_mirrors;
    for (int k = 0; k < n_mirrors; ++k) mir.p[k] = mirrors[k];
    return launch_map(d, static_cast<cudaStream_t>(stream), [=] __device__(long long i) {
        float p = param[i];
        float g = fmaf(weight_decay, p, grad[i]);
        if (momentum != 0.0f) {
            float const buf = first_step ? g : fmaf(momentum, mom[i], (1.0f - dampening) * g);
            mom[i] = buf;
            g = nesterov ? fmaf(momentum, buf, g) : buf;
        }
        p = fmaf(-lr, g, p);
        param[i] = p;
#pragma unroll
        for (int k = 0; k < kMaxMirrors; ++k)
            if (k < mir.n) mir.p[k][i] = p;
    }, "sgd step");
}

GB_API int gb_adam_step(float* param, float const* grad, float* exp_avg, float* exp_avg_sq, int64_t d, float lr, float beta1, float beta2,
                        float eps, float weight_decay, int step, int decoupled, void* stream) {
    float const bc1 = 1.0f - powf(beta1, static_cast<float>(step));
    float const bc2_sqrt = sqrtf(1.0f - powf(beta2, static_cast<float>(step)));
    float const step_size = lr / bc1;
    return launch_map(d, static_cast<cudaStream_t>(stream), [=] __device__(long long i) {
        float p = param[i];
        float g = grad[i];
        if (decoupled) p *= 1.0f - lr * weight_decay;
        else g = fmaf(weight_decay, p, g);
        float const m = fmaf(beta1, exp_avg[i], (1.0f - beta1) * g);
        float const v = fmaf(beta2, exp_avg_sq[i], (1.0f - beta2) * g * g);
        exp_avg[i] = m;
        exp_avg_sq[i] = v;
        float const denom = sqrtf(v) / bc2_sqrt + eps;
        param[i] = p - step_size * (m / denom);
    }, "adam step");
}

GB_API int gb_rmsprop_step(float* param, float const* grad, float* square_avg, float* mom, int64_t d, float lr, float alpha, float eps,
                           float weight_decay, float momentum, void* stream) {
    if (momentum != 0.0f && mom == nullptr) return cuda_fail("rmsprop: momentum buffer missing");
    return launch_map(d, static_cast<cudaStream_t>(stream), [=] __device__(long long i) {
        float p = param[i];
        float const g = fmaf(weight_decay, p, grad[i]);
        float const sq = fmaf(alpha, square_avg[i], (1.0f - alpha) * g * g);
        square_avg[i] = sq;
        float const avg = sqrtf(sq) + eps;
        if (momentum != 0.0f) {
            float const buf = fmaf(momentum, mom[i], g / avg);
            mom[i] = buf;
            p -= lr * buf;
        } else {
            p -= lr * (g / avg);
        }
        param[i] = p;
    }, "rmsprop step");
}

GB_API int gb_adagrad_step(float* param, float const* grad, float* state_sum, int64_t d, float lr, float lr_decay, float eps,
                           float weight_decay, int step, void* stream) {
    float const clr = lr / (1.0f + static_cast<float>(step - 1) * lr_decay);
    return launch_map(d, static_cast<cudaStream_t>(stream), [=] __device__(long long i) {
        float const p = param[i];
        float const g = fmaf(weight_decay, p, grad[i]);
        float const s = fmaf(g, g, state_sum[i]);
        state_sum[i] = s;
        param[i] = p - clr * (g / (sqrtf(s) + eps));
    }, "adagrad step");
}

// ------------------------------------------------------------------------------- attacks ---

// out[i] ~ U[0,1)   (ref: byzWorker.py:78-85 torch.rand)
GB_API int gb_attack_random(float* out, int64_t d, uint64_t seed, void* stream) {
    long long const quads = (d + 3) / 4;
    return launch_map(quads, static_cast<cudaStream_t>(stream), [=] __device__(long long q) {
        uint4 const r = philox4x32(seed, static_cast<unsigned long long>(q), 1u);
        uint32_t const bits[4] = {r.x, r.y, r.z, r.w};
        for (int k = 0; k < 4; ++k) {
            long long const i = q * 4 + k;
            if (i < d) out[i] = u01(bits[k]);
        }
    }, "attack random");
}

// out = factor * in   (reverse attack: factor = -100; ref: byzWorker.py:87-94)
GB_API int gb_scale(float const* in, float* out, int64_t d, float factor, void* stream) {
    return launch_map(d, static_cast<cudaStream_t>(stream), [=] __device__(long long i) { out[i] = factor * in[i]; }, "scale");
}

// out = in with each coordinate zeroed with probability p   (intended drop attack; ref: byzWorker.py:96-106, Q5)
GB_API int gb_attack_drop(float const* in, float* out, int64_t d, float p, uint64_t seed, void* stream) {
    long long const quads = (d + 3) / 4;
    return launch_map(quads, static_cast<cudaStream_t>(stream), [=] __device__(long long q) {
        uint4 const r = philox4x32(seed, static_cast<unsigned long long>(q), 2u);
        uint32_t const bits[4] = {r.x, r.y, r.z, r.w};
        for (int k = 0; k < 4; ++k) {
            long long const i = q * 4 + k;
            if (i < d) out[i] = (u01(bits[k]) < p) ? 0.0f : in[i];
        }
    }, "attack drop");
}

// Colluding attacks over the fw "honest" estimates in rows[0..fw):
//   kind 0 "lie":    out = μ + z·σ   (σ unbiased; 0 when fw == 1 unless nan_single)   ref: byzWorker.py:108-125
//   kind 1 "empire": out = −eps·μ                                                    ref: byzWorker.py:127-143
GB_API int gb_attack_collude(float const* const* rows, int fw, int64_t d, int kind, float coeff, int nan_single, float* out, void* stream) {
    if (fw < 1 || fw > kMaxRows) return cuda_fail("attack: need 1 <= fw <= 64 estimate rows");
    RowTable table{};
    for (int i = 0; i < fw; ++i) table.p[i] = rows[i];
    return launch_map(d, static_cast<cudaStream_t>(stream), [=] __device__(long long i) {
        float mean = 0.0f;
        for (int r = 0; r < fw; ++r) mean += table.p[r][i];
        mean /= static_cast<float>(fw);
        if (kind == 1) {
            out[i] = -coeff * mean;
            return;
        }
        float var = 0.0f;
        for (int r = 0; r < fw; ++r) {
            float const dv = table.p[r][i] - mean;
            var = fmaf(dv, dv, var);
        }
        float sigma;
        if (fw > 1) sigma = sqrtf(var / static_cast<float>(fw - 1));
        else sigma = nan_single ? NAN : 0.0f;
        out[i] = fmaf(coeff, sigma, mean);
    }, "attack collude");
}

// ------------------------------------------------------------------------------ utilities ---

// y = a·x + b·y
GB_API int gb_axpby(float const* x, float* y, int64_t d, float a, float b, void* stream) {
    return launch_map(d, static_cast<cudaStream_t>(stream), [=] __device__(long long i) { y[i] = fmaf(a, x[i], b * y[i]); }, "axpby");
}

// Copy one source vector into up to 8 destinations (model publication to peer arenas).
GB_API int gb_broadcast_copy(float const* src, float* const* dsts, int n_dst, int64_t d, void* stream) {
    if (n_dst < 1 || n_dst > kMaxMirrors) return cuda_fail("broadcast_copy: need 1..8 destinations");
    Mirrors mir{};
    mir.n = n_dst;
    for (int k = 0; k < n_dst; ++k) mir.p[k] = dsts[k];
    return launch_map(d, static_cast<cudaStream_t>(stream), [=] __device__(long long i) {
        float const v = src[i];
#pragma unroll
        for (int k = 0; k < kMaxMirrors; ++k)
            if (k < mir.n) mir.p[k][i] = v;
    }, "broadcast copy");
}

}  // extern "C"

// --------------------------------------------------------------- cross-GPU readiness flags ---
//
// Device-side replacement of the reference's RPC completion / quorum wait (garfieldpp/server.py:134-155):
// a producer publishes "my buffer for epoch e is complete" with a system-scope release store into a
// peer-mapped flag word; a consumer kernel spins (system-scope acquire loads) on a table of such
// flags until all reached the epoch or a deadline passed. Late peers are counted, not waited for:
// the robust rules tolerate f stale rows by construction.
//
// Only meaningful between kernels running on DIFFERENT GPUs (one process per GPU): two kernels that
// wait on each other on one GPU may never be co-scheduled.

namespace gb {
namespace {
__global__ void flag_signal_kernel(uint32_t* flag, uint32_t value) {
    __threadfence_system();  // everything this stream wrote before is visible before the flag flips
    ptx::st_release_sys(flag, value);
}

struct FlagTable {
    uint32_t* p[kMaxRows];
};

__global__ void flag_wait_kernel(FlagTable flags, int count, uint32_t value, unsigned long long timeout_ns, uint32_t* late) {
    int const i = threadIdx.x;
    if (i >= count) return;
    unsigned long long const deadline = ptx::globaltimer_ns() + timeout_ns;
    bool ready = false;
    while (!(ready = (static_cast<int32_t>(ptx::ld_acquire_sys(flags.p[i]) - value) >= 0))) {
        if (ptx::globaltimer_ns() > deadline) break;
        __nanosleep(200);
    }
    if (!ready && late != nullptr) atomicAdd(late, 1u);
}
}  // namespace
}  // namespace gb

namespace gb {
namespace {
// Device-resident epoch: the step counter lives in device memory so that a CUDA graph of the whole iteration, replayed every
// step, signals and waits for the right epoch without any kernel argument changing.
__global__ void epoch_bump_kernel(uint32_t* epoch) { *epoch += 1u; }

__global__ void flag_signal_ep_kernel(uint32_t* flag, const uint32_t* epoch) {
    __threadfence_system();
    ptx::st_release_sys(flag, *epoch);
}

// Wait until every flag reached *epoch − lag, or `quorum` of them did AND the deadline passed... precisely:
//   * all flags there                      → done
//   * deadline passed                      → done, the missing ones are late
// arrived[i] (optional) = 1 when flag i made it. Thread i polls flag i; the block agrees through shared memory.
__global__ void flag_wait_ep_kernel(FlagTable flags, int count, const uint32_t* epoch, uint32_t lag, unsigned long long timeout_ns, uint32_t* late,
                                    uint32_t* arrived) {
    int const i = threadIdx.x;
    if (i >= count) return;
    uint32_t const value = *epoch - lag;
    unsigned long long const deadline = ptx::globaltimer_ns() + timeout_ns;
    bool ready = false;
    while (!(ready = (static_cast<int32_t>(ptx::ld_acquire_sys(flags.p[i]) - value) >= 0))) {
        if (ptx::globaltimer_ns() > deadline) break;
        __nanosleep(100);
    }
    if (!ready && late != nullptr) atomicAdd(late, 1u);
    if (arrived != nullptr) arrived[i] = ready ? 1u : 0u;
}

// Quorum decision by the lead server (device-side form of the reference's "hand the rule the replies that made it",
// garfieldpp/server.py:134-155): poll the peers' "gradients ready" flags until all arrived, or the deadline passed with at
// least `quorum` of them there (below the quorum it keeps waiting, up to 16 deadlines). The resulting bit mask (bit i = flag
// i arrived) is stored into every rank's decision word, then every rank's decision flag is released with the epoch.
struct DecisionTable {
    uint32_t* word[kMaxRows];
    uint32_t* flag[kMaxRows];
};
__global__ void quorum_decide_kernel(FlagTable flags, int count, const uint32_t* epoch, int quorum, unsigned long long timeout_ns,
                                     DecisionTable out, int n_out, uint32_t* late) {
    __shared__ unsigned long long mask_sh;
    int const i = threadIdx.x;
    uint32_t const value = *epoch;
    if (i == 0) mask_sh = 0ull;
    __syncthreads();
    unsigned long long const start = ptx::globaltimer_ns();
    bool ready = false;
    if (i < count) {
        for (;;) {
            if (!ready && static_cast<int32_t>(ptx::ld_acquire_sys(flags.p[i]) - value) >= 0) {
                ready = true;
                atomicOr(&mask_sh, 1ull << i);
            }
            unsigned long long const seen = *reinterpret_cast<volatile unsigned long long*>(&mask_sh);
            int const have = __popcll(seen);
            unsigned long long const waited = ptx::globaltimer_ns() - start;
            if (have == count) break;
            if (waited > timeout_ns && have >= quorum) break;
            if (waited > 16ull * timeout_ns) break;
            __nanosleep(100);
        }
    }
    __syncthreads();
    unsigned long long const mask = mask_sh;
    if (i < count && !((mask >> i) & 1ull) && late != nullptr) atomicAdd(late, 1u);
    if (i < n_out) {
        out.word[i][0] = static_cast<uint32_t>(mask & 0xffffffffull);
        out.word[i][1] = static_cast<uint32_t>(mask >> 32);
        __threadfence_system();
        ptx::st_release_sys(out.flag[i], value);
    }
}

struct SlotTable {
    const float* p[kMaxRows];
};
// out[i] = Σ_r slots[r][i] in rank order (deterministic; every rank computes bitwise the same sum). Slots may be peer-mapped.
__global__ void __launch_bounds__(256) reduce_slots_kernel(SlotTable slots, int nslots, int count, float* __restrict__ out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        float acc = 0.f;
        for (int r = 0; r < nslots; ++r) acc += slots.p[r][i];
        out[i] = acc;
    }
}
}  // namespace
}  // namespace gb

extern "C" {

GB_API int gb_flag_signal(uint32_t* flag, uint32_t value, void* stream) {
    flag_signal_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(flag, value);
    return cuda_check(cudaGetLastError(), "flag signal");
}

// Wait until every flags[i] >= value (wrap-safe) or `timeout_ns` elapsed; *late += number of flags that
// did not make it. count <= 64.
GB_API int gb_flag_wait(uint32_t* const* flags, int count, uint32_t value, uint64_t timeout_ns, uint32_t* late, void* stream) {
    if (count < 1 || count > kMaxRows) return cuda_fail("flag wait: need 1..64 flags");
    FlagTable table{};
    for (int i = 0; i < count; ++i) table.p[i] = flags[i];
    flag_wait_kernel<<<1, 64, 0, static_cast<cudaStream_t>(stream)>>>(table, count, value, timeout_ns, late);
    return cuda_check(cudaGetLastError(), "flag wait");
}

GB_API int gb_epoch_bump(uint32_t* epoch, void* stream) {
    epoch_bump_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(epoch);
    return cuda_check(cudaGetLastError(), "epoch bump");
}

GB_API int gb_flag_signal_ep(uint32_t* flag, uint32_t const* epoch, void* stream) {
    flag_signal_ep_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(flag, epoch);
    return cuda_check(cudaGetLastError(), "flag signal (device epoch)");
}

GB_API int gb_flag_wait_ep(uint32_t* const* flags, int count, uint32_t const* epoch, uint32_t lag, uint64_t timeout_ns, uint32_t* late,
                           uint32_t* arrived, void* stream) {
    if (count < 1 || count > kMaxRows) return cuda_fail("flag wait: need 1..64 flags");
    FlagTable table{};
    for (int i = 0; i < count; ++i) table.p[i] = flags[i];
    flag_wait_ep_kernel<<<1, 64, 0, static_cast<cudaStream_t>(stream)>>>(table, count, epoch, lag, timeout_ns, late, arrived);
    return cuda_check(cudaGetLastError(), "flag wait (device epoch)");
}

GB_API int gb_quorum_decide(uint32_t* const* flags, int count, uint32_t const* epoch, int quorum, uint64_t timeout_ns, uint32_t* const* words,
                            uint32_t* const* decision_flags, int n_out, uint32_t* late, void* stream) {
    if (count < 1 || count > kMaxRows || n_out < 1 || n_out > kMaxRows || quorum < 1 || quorum > count)
        return cuda_fail("quorum decide: need 1..64 flags, 1..64 destinations and 1 <= quorum <= count");
    FlagTable table{};
    DecisionTable out{};
    for (int i = 0; i < count; ++i) table.p[i] = flags[i];
    for (int i = 0; i < n_out; ++i) { out.word[i] = words[i]; out.flag[i] = decision_flags[i]; }
    quorum_decide_kernel<<<1, 64, 0, static_cast<cudaStream_t>(stream)>>>(table, count, epoch, quorum, timeout_ns, out, n_out, late);
    return cuda_check(cudaGetLastError(), "quorum decide");
}

GB_API int gb_reduce_slots(float const* const* slots, int nslots, int count, float* out, void* stream) {
    if (nslots < 1 || nslots > kMaxRows || count < 1) return cuda_fail("reduce_slots: need 1..64 slots");
    SlotTable table{};
    for (int r = 0; r < nslots; ++r) table.p[r] = slots[r];
    int const blocks = (count + 255) / 256 < 16 ? (count + 255) / 256 : 16;
    reduce_slots_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(table, nslots, count, out);
    return cuda_check(cudaGetLastError(), "reduce_slots");
}

}  // extern "C"

// ------------------------------------------------------------------- multi-tensor gather ---
//
// Copy up to kMultiCopyMax separate device tensors into their slots of one flat vector in ONE launch
// (the per-parameter gradients autograd produced → the worker's gradient row). Replaces the 161
// accumulate-into-.grad add kernels per ResNet-50 backward pass (≈10 % of the step chain) and the
// reference's torch.cat of the gradients (garfieldpp/worker.py:93-94). The table travels in the
// kernel parameters, so the launch is capturable in a CUDA graph without any host buffer.

namespace gb {
namespace {
constexpr int kMultiCopyMax = 112;
constexpr int kMultiCopyChunk = 4096;  // floats per block
struct MultiCopyTable {
    const float* src[kMultiCopyMax];
    unsigned long long dst_off[kMultiCopyMax];
    unsigned int n[kMultiCopyMax];
};
__global__ void __launch_bounds__(256) multi_copy_kernel(MultiCopyTable t, float* __restrict__ dst) {
    int const which = blockIdx.y;
    unsigned int const n = t.n[which];
    unsigned int const begin = blockIdx.x * kMultiCopyChunk;
    if (begin >= n) return;
    unsigned int const end = min(n, begin + kMultiCopyChunk);
    const float* __restrict__ src = t.src[which];
    float* __restrict__ out = dst + t.dst_off[which];
    bool const vec = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0;
    if (vec) {
        unsigned int const end4 = begin + ((end - begin) & ~3u);
        for (unsigned int i = begin + threadIdx.x * 4; i < end4; i += blockDim.x * 4)
            *reinterpret_cast<float4*>(out + i) = *reinterpret_cast<const float4*>(src + i);
        for (unsigned int i = end4 + threadIdx.x; i < end; i += blockDim.x) out[i] = src[i];
    } else {
        for (unsigned int i = begin + threadIdx.x; i < end; i += blockDim.x) out[i] = src[i];
    }
}
}  // namespace
}  // namespace gb

extern "C" GB_API int gb_multi_copy(float const* const* srcs, uint64_t const* dst_offsets, uint32_t const* sizes, int count, float* dst,
                                    void* stream) {
    using namespace gb;
    if (count < 0) return cuda_fail("multi_copy: negative count");
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    for (int base = 0; base < count; base += kMultiCopyMax) {
        int const m = count - base < kMultiCopyMax ? count - base : kMultiCopyMax;
        MultiCopyTable t{};
        unsigned int longest = 0;
        for (int i = 0; i < m; ++i) {
            t.src[i] = srcs[base + i];
            t.dst_off[i] = dst_offsets[base + i];
            t.n[i] = sizes[base + i];
            if (sizes[base + i] > longest) longest = sizes[base + i];
        }
        if (longest == 0) continue;
        dim3 grid((longest + kMultiCopyChunk - 1) / kMultiCopyChunk, m);
        multi_copy_kernel<<<grid, 256, 0, st>>>(t, dst);
        if (cuda_check(cudaGetLastError(), "multi_copy launch")) return 1;
    }
    return 0;
}
