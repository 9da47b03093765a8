// Double-precision aggregation kernels: the fp64 instantiation of the two-pass scheme (pairwise distances + selection,
// then a per-coordinate rule on the rows or on W·rows) for CUDA float64 inputs.
//
// The reference's native modules dispatch on dtype ∈ {float, double} × {cpu, gpu}
// (/root/reference/pytorch_impl/libs/native/include/aggregator.hpp:123-133, aggregate_gpu_double; kernels templated on T in
// py_krum/krum.cu, py_bulyan/bulyan.cu, py_median/median.cu). The fp32 kernels of this library are TMA-streamed and tuned to
// the HBM roofline; tcgen05 has no fp64 path and fp64 gradients are a rarity (verification runs), so this variant is plain
// SIMT: coalesced grid-stride loads, shared-memory tiles, the SAME selection code (gar_select.cuh) and the same semantics.
// Still one launch per pass, no host round trip, deterministic (fixed-order reductions).

#include <cfloat>
#include <cmath>
#include <cstdio>

#include "gar_common.cuh"
#include "gar_select.cuh"
#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

constexpr int kF64Threads = 256;
constexpr int kF64Tile = 64;           // coordinates per shared-memory tile

struct RowTable64 {
    const double* p[kMaxRows];
};

struct Pairwise64Args {
    RowTable64 rows;
    int n;
    long long d;
    double* partials;        // gridDim.x × n × n
    unsigned int* counter;
    double* dist;            // n × n squared distances
    float* W;
    int* info;
    SelectArgs sel;
};

__global__ void __launch_bounds__(kF64Threads) pairwise_f64_kernel(const __grid_constant__ Pairwise64Args a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ unsigned int is_last;
    double* tile = reinterpret_cast<double*>(smem_raw);            // n × kF64Tile
    int const n = a.n, nn = n * n, pairs = n * (n - 1) / 2;
    constexpr int kMaxOwn = (kMaxRows * (kMaxRows - 1) / 2 + kF64Threads - 1) / kF64Threads;   // pairs owned by one thread
    double acc[kMaxOwn];
#pragma unroll
    for (int k = 0; k < kMaxOwn; ++k) acc[k] = 0.0;
    long long const tiles = (a.d + kF64Tile - 1) / kF64Tile;
    for (long long t = blockIdx.x; t < tiles; t += gridDim.x) {
        long long const x0 = t * kF64Tile;
        for (int e = threadIdx.x; e < n * kF64Tile; e += kF64Threads) {
            int const i = e / kF64Tile, c = e % kF64Tile;
            tile[e] = (x0 + c < a.d) ? a.rows.p[i][x0 + c] : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kMaxOwn; ++k) {
            int const p = threadIdx.x + k * kF64Threads;
            if (p < pairs) {
                // unrank pair p -> (i < j), row-major over the upper triangle
                int i = 0, rem = p, len = n - 1;
                while (rem >= len) { rem -= len; ++i; --len; }
                int const j = i + 1 + rem;
                double s = 0.0;
                for (int c = 0; c < kF64Tile; ++c) {
                    double const diff = tile[i * kF64Tile + c] - tile[j * kF64Tile + c];
                    s = fma(diff, diff, s);
                }
                acc[k] += s;
            }
        }
        __syncthreads();
    }
    double* mine = a.partials + static_cast<size_t>(blockIdx.x) * nn;
#pragma unroll
    for (int k = 0; k < kMaxOwn; ++k) {
        int const p = threadIdx.x + k * kF64Threads;
        if (p < pairs) {
            int i = 0, rem = p, len = n - 1;
            while (rem >= len) { rem -= len; ++i; --len; }
            int const j = i + 1 + rem;
            mine[i * n + j] = acc[k];
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = (atomicAdd(a.counter, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    for (int e = threadIdx.x; e < nn; e += blockDim.x) {
        int const i = e / n, j = e % n;
        double out = INFINITY;
        if (i != j) {
            int const lo = i < j ? i : j, hi = i < j ? j : i;
            double total = 0.0;
            for (unsigned b = 0; b < gridDim.x; ++b) total += __ldcg(a.partials + static_cast<size_t>(b) * nn + lo * n + hi);
            out = (fabs(total) <= DBL_MAX) ? total : INFINITY;      // non-finite -> +inf
        }
        a.dist[e] = out;
    }
    if (threadIdx.x == 0) *a.counter = 0u;
    __threadfence_block();
    __syncthreads();
    run_selection(a, *reinterpret_cast<SelectScratch*>(smem_raw));
}

struct Coord64Args {
    RowTable64 rows;
    int n, rule, rank, upper, theta;
    long long d;
    const float* W;
    double* out;
};

// One thread per coordinate; the n (or theta) values of the coordinate live in a local array and are insertion-sorted.
__global__ void __launch_bounds__(kF64Threads) coordwise_f64_kernel(const __grid_constant__ Coord64Args a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // theta × n weights. The selection emits uniform averaging weights (1/m over the chosen rows) as floats; in double
    // precision they are re-derived exactly: 1.0 / (number of non-zero entries of the row).
    double* Ws = reinterpret_cast<double*>(smem_raw);
    for (int k = threadIdx.x; k < a.theta; k += blockDim.x) {
        int chosen = 0;
        for (int i = 0; i < a.n; ++i) chosen += a.W[k * a.n + i] != 0.0f ? 1 : 0;
        for (int i = 0; i < a.n; ++i) Ws[k * a.n + i] = a.W[k * a.n + i] != 0.0f ? 1.0 / static_cast<double>(chosen) : 0.0;
    }
    __syncthreads();
    for (long long x = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; x < a.d; x += static_cast<long long>(gridDim.x) * blockDim.x) {
        double v[kMaxRows];
        int cnt;
        if (a.theta > 0) {
            cnt = a.theta;
            for (int k = 0; k < cnt; ++k) v[k] = 0.0;
            for (int i = 0; i < a.n; ++i) {
                bool used = false;
                for (int k = 0; k < cnt; ++k) used |= Ws[k * a.n + i] != 0.0;
                if (!used) continue;                     // rows with zero weight may hold NaN: never touch them
                double const xi = a.rows.p[i][x];
                for (int k = 0; k < cnt; ++k) v[k] = fma(Ws[k * a.n + i], xi, v[k]);
            }
        } else {
            cnt = a.n;
            for (int i = 0; i < cnt; ++i) v[i] = a.rows.p[i][x];
        }
        double result;
        if (a.rule == kRuleLinear) {
            result = v[0];
        } else {
            bool has_nan = false;
            int finite = 0;
            for (int i = 0; i < cnt; ++i) {
                has_nan |= v[i] != v[i];
                bool const fin = fabs(v[i]) <= DBL_MAX;
                finite += fin ? 1 : 0;
                if (a.rule == kRuleMedianFiniteUpper) { if (!fin) v[i] = INFINITY; }
                else if (v[i] != v[i]) v[i] = INFINITY;
            }
            for (int i = 1; i < cnt; ++i) {             // insertion sort (cnt <= 64)
                double const key = v[i];
                int j = i - 1;
                while (j >= 0 && v[j] > key) { v[j + 1] = v[j]; --j; }
                v[j + 1] = key;
            }
            if (a.rule == kRuleMedianLower) {
                result = has_nan ? NAN : v[(cnt - 1) >> 1];
            } else if (a.rule == kRuleMedianFiniteUpper) {
                result = finite == 0 ? 0.0 : v[finite >> 1];
            } else if (a.rule == kRuleTrimmedMean) {
                double s = 0.0;
                for (int i = a.rank; i < cnt - a.rank; ++i) s += v[i];
                result = has_nan ? NAN : s / static_cast<double>(cnt - 2 * a.rank);
            } else {                                     // averaged median: mean of the `rank` values closest to the median
                double const med = v[a.upper ? (cnt >> 1) : ((cnt - 1) >> 1)];
                double key[kMaxRows];
                for (int i = 0; i < cnt; ++i) key[i] = fabs(v[i] - med);
                for (int i = 1; i < cnt; ++i) {         // sort values by closeness (stable)
                    double const kk = key[i], vv = v[i];
                    int j = i - 1;
                    while (j >= 0 && key[j] > kk) { key[j + 1] = key[j]; v[j + 1] = v[j]; --j; }
                    key[j + 1] = kk; v[j + 1] = vv;
                }
                double s = 0.0;
                for (int i = 0; i < a.rank; ++i) s += v[i];
                result = has_nan ? NAN : s / static_cast<double>(a.rank);
            }
        }
        a.out[x] = result;
    }
}

}  // namespace
}  // namespace gb

using namespace gb;

extern "C" GB_API int64_t gb_pairwise_scratch_bytes_f64(int n) {
    return 256 + 2LL * device_sm_count() * n * n * static_cast<int64_t>(sizeof(double));
}

// fp64 counterpart of gb_pairwise_select: dist is n×n doubles; W / info as in the fp32 entry point.
extern "C" GB_API int gb_pairwise_select_f64(double const* const* rows, int n, int64_t d, int mode, int f, int m, int neighbors, int squared,
                                             void* scratch, int64_t scratch_bytes, double* dist, float* W, int* info, void* stream) {
    if (n < 1 || n > kMaxRows) return cuda_fail("pairwise_f64: need 1 <= n <= 64 rows");
    if (d < 1) return cuda_fail("pairwise_f64: empty rows");
    if (dist == nullptr || scratch == nullptr) return cuda_fail("pairwise_f64: missing output/scratch buffer");
    if (mode != kSelectNone && (W == nullptr || info == nullptr)) return cuda_fail("pairwise_f64: selection needs W and info buffers");
    if (char const* bad = selection_error(mode, n, f, m, neighbors)) {
        static thread_local char msg[160];
        std::snprintf(msg, sizeof(msg), "pairwise_f64: %s", bad);
        return cuda_fail(msg);
    }
    Pairwise64Args a{};
    for (int i = 0; i < n; ++i) a.rows.p[i] = rows[i];
    a.n = n; a.d = d; a.dist = dist; a.W = W; a.info = info;
    a.sel.mode = mode; a.sel.f = f; a.sel.m = m; a.sel.neighbors = neighbors; a.sel.squared = squared;
    long long const tiles = (d + kF64Tile - 1) / kF64Tile;
    long long grid = 2LL * device_sm_count();
    if (grid > tiles) grid = tiles;
    if (scratch_bytes < 256 + grid * n * n * static_cast<int64_t>(sizeof(double))) return cuda_fail("pairwise_f64: scratch buffer too small");
    a.counter = static_cast<unsigned int*>(scratch);
    a.partials = reinterpret_cast<double*>(static_cast<unsigned char*>(scratch) + 256);
    size_t smem = static_cast<size_t>(n) * kF64Tile * sizeof(double);
    if (smem < sizeof(SelectScratch)) smem = sizeof(SelectScratch);
    auto kern = pairwise_f64_kernel;
    if (cuda_check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)), "pairwise_f64 smem attribute")) return 1;
    if (cuda_check(cudaMemsetAsync(a.partials, 0, static_cast<size_t>(grid) * n * n * sizeof(double), static_cast<cudaStream_t>(stream)), "pairwise_f64 memset"))
        return 1;
    kern<<<static_cast<unsigned>(grid), kF64Threads, smem, static_cast<cudaStream_t>(stream)>>>(a);
    return cuda_check(cudaGetLastError(), "pairwise_f64 launch");
}

// fp64 counterpart of gb_coordwise (rules 0-4; theta > 0: the rule is applied to W·rows, W float theta×n).
extern "C" GB_API int gb_coordwise_f64(double const* const* rows, int n, int64_t d, int rule, int rank, int upper, float const* W, int theta,
                                       double* out, void* stream) {
    if (n < 1 || n > kMaxRows) return cuda_fail("coordwise_f64: need 1 <= n <= 64 rows");
    if (d < 1 || out == nullptr) return cuda_fail("coordwise_f64: empty rows or missing output");
    if (theta < 0 || theta > kMaxRows || (theta > 0 && W == nullptr)) return cuda_fail("coordwise_f64: bad weight matrix");
    if (rule < kRuleLinear || rule > kRuleAveragedMedian) return cuda_fail("coordwise_f64: unsupported rule");
    if (rule == kRuleLinear && theta != 1) return cuda_fail("coordwise_f64: the linear rule needs a 1×n weight row");
    int const cnt = theta > 0 ? theta : n;
    if (rule == kRuleTrimmedMean && (rank < 0 || 2 * rank >= cnt)) return cuda_fail("coordwise_f64: trimmed mean needs 0 <= 2f < n");
    if (rule == kRuleAveragedMedian && (rank < 1 || rank > cnt)) return cuda_fail("coordwise_f64: averaged median needs 1 <= beta <= n");
    Coord64Args a{};
    for (int i = 0; i < n; ++i) a.rows.p[i] = rows[i];
    a.n = n; a.rule = rule; a.rank = rank; a.upper = upper; a.theta = theta; a.d = d; a.W = W; a.out = out;
    long long blocks = (d + kF64Threads - 1) / kF64Threads;
    if (blocks > 16LL * device_sm_count()) blocks = 16LL * device_sm_count();
    size_t const smem = static_cast<size_t>(theta > 0 ? theta * n : 1) * sizeof(double);
    coordwise_f64_kernel<<<static_cast<unsigned>(blocks), kF64Threads, smem, static_cast<cudaStream_t>(stream)>>>(a);
    return cuda_check(cudaGetLastError(), "coordwise_f64 launch");
}
