// pull_median_operand: the head of the MSMW model pull (north-star path 3) — the coordinate range of the FIRST layer's
// weight is pulled from all P server replicas (local or peer-mapped rows), reduced with the coordinate-wise lower median and
// written twice in the same pass: into the worker's flat model and, zero-padded to a 16-byte row pitch, as the K-major
// [Cout][kpad] operand of the first forward GEMM (what the tensor-core convolution of the stem consumes). The first GEMM
// can therefore start as soon as this (tiny) kernel is done, while the remaining coordinate chunks of the pull — issued in
// forward-layer order on their own stream, each gating the layers that need it — are still crossing NVLink.
//
// Reference: workers fetch the models of all servers over RPC, aggregate them on the host, load the result into the module
// and only then start the forward pass (applications/ByzSGD/trainer.py:237-244, Garfield_CC/trainer.py:170-196).

#include <cmath>

#include "gar_common.cuh"
#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

constexpr int kMaxReplicas = 16;

__global__ void __launch_bounds__(256) pull_median_operand_kernel(RowTable rows, int P, int rows_out, int k, int kpad, float* __restrict__ flat,
                                                                   float* __restrict__ operand) {
    long long const total = static_cast<long long>(rows_out) * kpad;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        int const c = static_cast<int>(i % kpad);
        long long const r = i / kpad;
        float med = 0.f;
        if (c < k) {
            long long const x = r * k + c;
            float v[kMaxReplicas];
            bool nan = false;
            int cnt = 0;
            for (int p = 0; p < P; ++p) {            // insertion sort of the P replicas' values (P <= 16)
                float const val = rows.p[p][x];
                nan |= val != val;
                int pos = cnt++;
                while (pos > 0 && v[pos - 1] > val) { v[pos] = v[pos - 1]; --pos; }
                v[pos] = val;
            }
            med = nan ? NAN : v[(P - 1) >> 1];        // torch.median semantics: lower median, NaN propagates
            flat[x] = med;
        }
        operand[i] = med;
    }
}

// out[x] (every rank, every slot) = scale · Σ_ranks Σ_slots in[slot·pitch + x]: one multimem.ld_reduce per slot and 16 bytes (the
// switch adds the ranks' copies), one multimem.st per slot for the result.
__global__ void __launch_bounds__(256) multimem_average_kernel(const float* __restrict__ mc_in, float* __restrict__ mc_out, int slots, long long pitch,
                                                                long long lo, long long hi, float scale) {
    long long const n4 = (hi - lo) / 4;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += static_cast<long long>(gridDim.x) * blockDim.x) {
        long long const x = lo + 4 * i;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int s = 0; s < slots; ++s) {
            float4 const v = ptx::multimem_ld_reduce_add(mc_in + s * pitch + x);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
        for (int s = 0; s < slots; ++s) ptx::multimem_st(mc_out + s * pitch + x, acc);
    }
}

}  // namespace
}  // namespace gb

extern "C" GB_API int gb_multimem_average(float const* mc_in, float* mc_out, int slots, int64_t pitch, int64_t lo, int64_t hi, float scale,
                                          void* stream) {
    using namespace gb;
    if (mc_in == nullptr || mc_out == nullptr || slots < 1 || pitch < 4 || lo < 0 || hi < lo || (lo % 4) || (hi % 4) || (pitch % 4))
        return cuda_fail("multimem_average: need multicast addresses and 16-byte aligned slices");
    if ((reinterpret_cast<uintptr_t>(mc_in) | reinterpret_cast<uintptr_t>(mc_out)) & 15u) return cuda_fail("multimem_average: unaligned multicast address");
    if (hi == lo) return 0;
    long long blocks = ((hi - lo) / 4 + 255) / 256;
    if (blocks > 8LL * device_sm_count()) blocks = 8LL * device_sm_count();
    multimem_average_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(mc_in, mc_out, slots, pitch, lo, hi, scale);
    return cuda_check(cudaGetLastError(), "multimem_average");
}

extern "C" GB_API int gb_pull_median_operand(float const* const* rows, int replicas, int rows_out, int k, int kpad, float* flat, float* operand,
                                             void* stream) {
    using namespace gb;
    if (replicas < 1 || replicas > kMaxReplicas || rows_out < 1 || k < 1 || kpad < k) return cuda_fail("pull_median_operand: bad arguments (1..16 replicas)");
    RowTable table{};
    for (int p = 0; p < replicas; ++p) table.p[p] = rows[p];
    long long const total = static_cast<long long>(rows_out) * kpad;
    long long blocks = (total + 255) / 256;
    if (blocks > 4LL * device_sm_count()) blocks = 4LL * device_sm_count();
    pull_median_operand_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(table, replicas, rows_out, k, kpad, flat,
                                                                                                          operand);
    return cuda_check(cudaGetLastError(), "pull_median_operand");
}
