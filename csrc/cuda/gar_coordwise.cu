// coordwise_select: one streaming pass over n gradient rows that produces, per coordinate, a
// rank statistic of the n values (or of θ weighted mixes of them), optionally followed by a
// fused SGD update of the parameter arena and stores into peer arenas.
//
// Counterpart of the reference kernels K3 `selection_average`, K8 `averaged_median`,
// K9 `simple_median`, K10 `median` (/root/reference/pytorch_impl/libs/native/include/
// operations.cu.hpp:57-139, py_bulyan/bulyan.cu:215-243, py_median/median.cu:60-83) and of
// TF `median_nan` / `averaged_median` (tensorflow_impl/rsrcs/native/op_median/gpu.cu.cpp:66-83),
// re-designed for sm_100a:
//   * rows are pulled tile by tile with cp.async.bulk (TMA 1-D) into a 3-stage mbarrier ring
//     by a dedicated producer warp — works unchanged on peer-mapped rows (fused gather+GAR);
//   * selection is a fully unrolled register sorting network (no quick-select, no per-thread
//     shared-memory slices, no n ≤ 96 limit from the 48 KiB dynamic smem of K10);
//   * Bulyan's θ×d intermediate never exists: the θ round averages of a coordinate are formed
//     in registers from the weight matrix W (θ×n) right before the network;
//   * the optimizer update and the model publication are fused in the epilogue.

#include <cfloat>
#include <cmath>

#include "gar_common.cuh"
#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

constexpr float kInf = INFINITY;

__device__ __forceinline__ void apply_update(FusedUpdate const& u, long long x, float g, float* out) {
    if (out != nullptr) out[x] = g;
    if (u.mode == 0) return;
    float p = u.param[x];
    g = fmaf(u.weight_decay, p, g);
    if (u.momentum != 0.0f) {
        float buf = u.first_step ? g : fmaf(u.momentum, u.mom[x], (1.0f - u.dampening) * g);
        u.mom[x] = buf;
        g = u.nesterov ? fmaf(u.momentum, buf, g) : buf;
    }
    p = fmaf(-u.lr, g, p);
    u.param[x] = p;
    for (int k = 0; k < u.n_mirrors; ++k) u.mirror[k][x] = p;
}

// Reduce the `cnt` values of one coordinate (already gathered in v[0..NP), +inf padded).
template <int NP, bool PAIRS>
__device__ __forceinline__ float reduce_column(float (&v)[NP], int cnt, CoordArgs const& a, long long x, float first_row_value) {
    int const rule = a.rule;
    if (rule == kRuleLinear) return v[0];

    bool has_nan = false;
    int finite = 0;
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        if (i < cnt) {
            float const val = v[i];
            has_nan |= (val != val);
            bool const fin = fabsf(val) <= FLT_MAX;
            finite += fin ? 1 : 0;
            if (rule == kRuleMedianFiniteUpper) v[i] = fin ? val : kInf;
            else if (val != val) v[i] = kInf;
        } else {
            v[i] = kInf;
        }
    }
    sort_regs<NP>(v);

    float result = 0.0f;
    if (rule == kRuleMedianLower || rule == kRuleCondense) {
        int const k = (cnt - 1) >> 1;
#pragma unroll
        for (int i = 0; i < NP; ++i) result = (i == k) ? v[i] : result;
        if (has_nan) result = NAN;
        if (rule == kRuleCondense) {
            uint4 const r = philox4x32(a.seed, static_cast<unsigned long long>(x), 0u);
            result = (u01(r.x) < a.p) ? result : first_row_value;
        }
    } else if (rule == kRuleMedianFiniteUpper) {
        int const k = finite >> 1;
#pragma unroll
        for (int i = 0; i < NP; ++i) result = (i == k) ? v[i] : result;
        if (finite == 0) result = 0.0f;
    } else if (rule == kRuleTrimmedMean) {
        int const lo = a.rank, hi = cnt - a.rank;
        float sum = 0.0f;
#pragma unroll
        for (int i = 0; i < NP; ++i) sum += (i >= lo && i < hi) ? v[i] : 0.0f;
        result = has_nan ? NAN : sum / static_cast<float>(hi - lo);
    } else if (PAIRS && rule == kRuleAveragedMedian) {
        int const k = a.upper ? (cnt >> 1) : ((cnt - 1) >> 1);
        float med = 0.0f;
#pragma unroll
        for (int i = 0; i < NP; ++i) med = (i == k) ? v[i] : med;
        float key[NP];
#pragma unroll
        for (int i = 0; i < NP; ++i) key[i] = (i < cnt) ? fabsf(v[i] - med) : kInf;
        sort_pairs<NP>(key, v);
        float sum = 0.0f;
#pragma unroll
        for (int i = 0; i < NP; ++i) sum += (i < a.rank) ? v[i] : 0.0f;
        result = has_nan ? NAN : sum / static_cast<float>(a.rank);
    }
    return result;
}

// Gather the column of coordinate x through `load(i)` and reduce it.
template <int NP, bool PAIRS, class Load>
__device__ __forceinline__ float column(CoordArgs const& a, float const* Ws, unsigned long long used_mask, long long x, Load load) {
    float v[NP];
    float first = 0.0f;
    int cnt;
    if (a.theta > 0) {  // mixes y_k = Σ_i W[k][i]·x_i (zero weights skipped: excluded rows may hold NaN)
        cnt = a.theta;
#pragma unroll
        for (int k = 0; k < NP; ++k) v[k] = 0.0f;
        for (int i = 0; i < a.n; ++i) {
            if (!((used_mask >> i) & 1ull)) continue;  // warp-uniform
            float const xi = load(i);
#pragma unroll
            for (int k = 0; k < NP; ++k)
                if (k < a.theta) v[k] = fmaf(Ws[k * a.n + i], xi, v[k]);
        }
    } else {
        cnt = a.n;
#pragma unroll
        for (int i = 0; i < NP; ++i) v[i] = (i < a.n) ? load(i) : kInf;
        first = v[0];
    }
    return reduce_column<NP, PAIRS>(v, cnt, a, x, first);
}

template <int NP, bool PAIRS>
__global__ void __launch_bounds__(kThreads) coordwise_kernel(const __grid_constant__ CoordArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ TileRing ring;
    float* Ws = reinterpret_cast<float*>(smem_raw);                       // theta × n weights
    int const w_floats = ((a.theta * a.n + 31) / 32) * 32;
    float* stages = Ws + w_floats;                                        // kStages × n × tile
    int const warp = threadIdx.x >> 5;
    int const tile = a.tile;
    long long const full_tiles = a.use_tma ? a.d / tile : 0;

    for (int i = threadIdx.x; i < a.theta * a.n; i += blockDim.x) Ws[i] = a.W[i];
    ring_init(ring);  // contains the __syncthreads that also publishes Ws
    unsigned long long used_mask = 0ull;  // rows with at least one non-zero weight
    for (int i = 0; i < a.n; ++i) {
        bool used = a.theta == 0;
        for (int k = 0; k < a.theta; ++k) used |= (Ws[k * a.n + i] != 0.0f);
        used_mask |= used ? (1ull << i) : 0ull;
    }

    if (warp == 0) {
        if (full_tiles > 0 && ptx::elect_one()) {
            ring_produce(ring, stages, a.rows, a.n, tile, full_tiles, [&](int i) { return ((used_mask >> i) & 1ull) != 0ull; });
        }
    } else {
        int const ctid = threadIdx.x - 32;
        constexpr int kConsumers = 32 * kConsumerWarps;
        size_t const stage_floats = static_cast<size_t>(a.n) * tile;
        long long it = 0;
        for (long long t = blockIdx.x; t < full_tiles; t += gridDim.x, ++it) {
            int const s = static_cast<int>(it % kStages);
            uint32_t const phase = static_cast<uint32_t>((it / kStages) & 1);
            ptx::mbar_wait(&ring.full[s], phase);
            float const* src = stages + s * stage_floats;
            for (int c = ctid; c < tile; c += kConsumers) {
                long long const x = t * tile + c;
                float const r = column<NP, PAIRS>(a, Ws, used_mask, x, [&](int i) { return src[static_cast<size_t>(i) * tile + c]; });
                apply_update(a.upd, x, r, a.out);
            }
            __syncwarp();
            if ((threadIdx.x & 31) == 0) ptx::mbar_arrive(&ring.empty[s]);
        }
        // remainder (and everything when rows are unaligned): direct coalesced global loads
        long long const done = full_tiles * tile;
        for (long long x = done + static_cast<long long>(blockIdx.x) * kConsumers + ctid; x < a.d;
             x += static_cast<long long>(gridDim.x) * kConsumers) {
            float const r = column<NP, PAIRS>(a, Ws, used_mask, x, [&](int i) { return a.rows.p[i][x]; });
            apply_update(a.upd, x, r, a.out);
        }
    }
}

template <int NP, bool PAIRS>
cudaError_t launch(CoordArgs const& a, int grid, size_t smem, cudaStream_t stream) {
    auto kern = coordwise_kernel<NP, PAIRS>;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (err != cudaSuccess) return err;
    kern<<<grid, kThreads, smem, stream>>>(a);
    return cudaGetLastError();
}

}  // namespace

int pick_tile(int n) {
    int tile = (32768 / (4 * n)) / 128 * 128;
    if (tile < 128) tile = 128;
    if (tile > 1024) tile = 1024;
    return tile;
}

bool rows_aligned(RowTable const& rows, int n) {
    for (int i = 0; i < n; ++i)
        if (reinterpret_cast<uintptr_t>(rows.p[i]) & 15u) return false;
    return true;
}

}  // namespace gb

extern "C" GB_API int gb_coordwise(float const* const* rows, int n, int64_t d, int rule, int rank, int upper, float p, uint64_t seed,
                                   float const* W, int theta, float* out, gb_fused_update const* upd, void* stream) {
    using namespace gb;
    if (n < 1 || n > kMaxRows) return cuda_fail("coordwise: need 1 <= n <= 64 rows");
    if (d < 1) return cuda_fail("coordwise: empty rows");
    if (theta < 0 || theta > kMaxRows || (theta > 0 && W == nullptr)) return cuda_fail("coordwise: bad weight matrix");
    if (rule == kRuleLinear && theta != 1) return cuda_fail("coordwise: the linear rule needs a 1×n weight row");
    int const cnt = theta > 0 ? theta : n;
    if (rule == kRuleTrimmedMean && (rank < 0 || 2 * rank >= cnt)) return cuda_fail("coordwise: trimmed mean needs 0 <= 2f < n");
    if (rule == kRuleAveragedMedian && (rank < 1 || rank > cnt)) return cuda_fail("coordwise: averaged median needs 1 <= beta <= n");
    if (out == nullptr && (upd == nullptr || upd->mode == 0)) return cuda_fail("coordwise: no output requested");

    CoordArgs a{};
    for (int i = 0; i < n; ++i) a.rows.p[i] = rows[i];
    a.n = n; a.rule = rule; a.rank = rank; a.upper = upper; a.theta = theta; a.p = p; a.seed = seed; a.d = d; a.W = W; a.out = out;
    if (upd != nullptr && upd->mode != 0) {
        if (upd->n_mirrors < 0 || upd->n_mirrors > kMaxMirrors) return cuda_fail("coordwise: too many mirrors");
        a.upd.mode = upd->mode; a.upd.nesterov = upd->nesterov; a.upd.first_step = upd->first_step; a.upd.n_mirrors = upd->n_mirrors;
        a.upd.lr = upd->lr; a.upd.momentum = upd->momentum; a.upd.dampening = upd->dampening; a.upd.weight_decay = upd->weight_decay;
        a.upd.param = upd->param; a.upd.mom = upd->mom;
        for (int k = 0; k < upd->n_mirrors; ++k) a.upd.mirror[k] = upd->mirror[k];
    }
    a.tile = pick_tile(n);
    a.use_tma = rows_aligned(a.rows, n) ? 1 : 0;

    int const w_floats = ((theta * n + 31) / 32) * 32;
    size_t const smem = sizeof(float) * (static_cast<size_t>(w_floats) + static_cast<size_t>(kStages) * n * a.tile);
    long long const tiles = (d + a.tile - 1) / a.tile;
    int const ctas_per_sm = smem <= 100 * 1024 ? 2 : 1;
    long long grid = static_cast<long long>(device_sm_count()) * ctas_per_sm;
    if (grid > tiles) grid = tiles;
    if (grid < 1) grid = 1;
    cudaStream_t const st = static_cast<cudaStream_t>(stream);

    cudaError_t err;
    bool const pairs = (rule == kRuleAveragedMedian);
    int const np = cnt <= 4 ? 4 : cnt <= 8 ? 8 : cnt <= 16 ? 16 : cnt <= 32 ? 32 : 64;
    if (rule == kRuleLinear) err = launch<4, false>(a, static_cast<int>(grid), smem, st);
    else if (pairs) {
        switch (np) {
            case 4: err = launch<4, true>(a, static_cast<int>(grid), smem, st); break;
            case 8: err = launch<8, true>(a, static_cast<int>(grid), smem, st); break;
            case 16: err = launch<16, true>(a, static_cast<int>(grid), smem, st); break;
            case 32: err = launch<32, true>(a, static_cast<int>(grid), smem, st); break;
            default: err = launch<64, true>(a, static_cast<int>(grid), smem, st); break;
        }
    } else {
        switch (np) {
            case 4: err = launch<4, false>(a, static_cast<int>(grid), smem, st); break;
            case 8: err = launch<8, false>(a, static_cast<int>(grid), smem, st); break;
            case 16: err = launch<16, false>(a, static_cast<int>(grid), smem, st); break;
            case 32: err = launch<32, false>(a, static_cast<int>(grid), smem, st); break;
            default: err = launch<64, false>(a, static_cast<int>(grid), smem, st); break;
        }
    }
    return cuda_check(err, "coordwise launch");
}
