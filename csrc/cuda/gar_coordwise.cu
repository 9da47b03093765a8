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
//   * selection is a fully unrolled register sorting network, 4 coordinates per thread
//     (128-bit shared loads / global stores), the rule is a compile-time parameter — no
//     quick-select, no per-thread shared-memory slices, no n ≤ 96 limit (K10's 48 KiB smem);
//   * Bulyan's θ×d intermediate never exists: the θ round averages of a coordinate are formed
//     in registers from the weight matrix W (θ×n) right before the network;
//   * the optimizer update and the model publication are fused in the epilogue.

#include <cstdlib>

#include "gar_coordwise.cuh"
#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {

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
    if (theta > 0 && rule != kRuleLinear && rule != kRuleAveragedMedian) return cuda_fail("coordwise: mixes support the linear and averaged-median rules");
    int const cnt = theta > 0 ? theta : n;
    if (rule == kRuleTrimmedMean && (rank < 0 || 2 * rank >= cnt)) return cuda_fail("coordwise: trimmed mean needs 0 <= 2f < n");
    if (rule == kRuleAveragedMedian && (rank < 1 || rank > cnt)) return cuda_fail("coordwise: averaged median needs 1 <= beta <= n");
    if (out == nullptr && (upd == nullptr || upd->mode == 0)) return cuda_fail("coordwise: no output requested");

    CoordArgs a{};
    for (int i = 0; i < n; ++i) a.rows.p[i] = rows[i];
    a.n = n; a.rule = rule; a.rank = rank; a.upper = upper; a.theta = theta; a.p = p; a.seed = seed; a.d = d; a.W = W; a.out = out;
    bool aligned = rows_aligned(a.rows, n) && (reinterpret_cast<uintptr_t>(out) & 15u) == 0;
    if (upd != nullptr && upd->mode != 0) {
        if (upd->n_mirrors < 0 || upd->n_mirrors > kMaxMirrors) return cuda_fail("coordwise: too many mirrors");
        if (upd->param == nullptr || (upd->momentum != 0.0f && upd->mom == nullptr)) return cuda_fail("coordwise: fused update needs param (and momentum) buffers");
        a.upd.mode = upd->mode; a.upd.nesterov = upd->nesterov; a.upd.first_step = upd->first_step; a.upd.n_mirrors = upd->n_mirrors;
        a.upd.lr = upd->lr; a.upd.momentum = upd->momentum; a.upd.dampening = upd->dampening; a.upd.weight_decay = upd->weight_decay;
        a.upd.param = upd->param; a.upd.mom = upd->mom; a.upd.multicast = upd->multicast;
        if (upd->multicast != nullptr && upd->n_mirrors != 0) return cuda_fail("coordwise: multicast publication excludes per-peer mirrors");
        aligned = aligned && (reinterpret_cast<uintptr_t>(upd->multicast) & 15u) == 0;
        aligned = aligned && (reinterpret_cast<uintptr_t>(upd->param) & 15u) == 0 && (reinterpret_cast<uintptr_t>(upd->mom) & 15u) == 0;
        for (int k = 0; k < upd->n_mirrors; ++k) {
            a.upd.mirror[k] = upd->mirror[k];
            aligned = aligned && (reinterpret_cast<uintptr_t>(upd->mirror[k]) & 15u) == 0;
        }
    }
    a.tile = pick_tile(n);
    a.use_tma = aligned ? 1 : 0;  // otherwise every coordinate takes the scalar direct-load path

    // register-array size of the sorting network: the next bucket >= cnt (pruned Batcher networks)
    int const np = rule == kRuleLinear ? 1 : cnt <= 4 ? 4 : cnt <= 32 ? (cnt + 3) / 4 * 4 : cnt <= 48 ? 48 : 64;
    // transposed weights, rows padded to 4 (gar_coordwise.cuh); the weighted sum keeps one weight per input
    int const w_floats = theta > 0 ? (((np == 1 ? 1 : (np + 3) / 4 * 4) * ((n + 3) / 4 * 4) + 31) / 32) * 32 : 0;
    // Weighted sums and W·x mixes of more than 16 rows: the 32 KB stages leave tiles of 256-384 coordinates, i.e. 64-96 of the 256
    // consumer threads busy at 4 coordinates each. One CTA per SM with ≈220 KB of stages instead (n = 23, d = 23.5 M: weighted sum
    // 0.61 -> 0.43 ms, Bulyan's second pass 2.54 -> 1.96 ms); the sorting rules without a mix keep two CTAs per SM (median 0.70
    // against 0.90 ms).
    if ((rule == kRuleLinear || theta > 0) && a.tile < 512) {
        int tile = (220 * 1024 - w_floats * 4) / (kStages * n * 4) / 128 * 128;
        if (tile > 1024) tile = 1024;
        if (tile > a.tile) a.tile = tile;
    }
    size_t const smem = sizeof(float) * (static_cast<size_t>(w_floats) + static_cast<size_t>(kStages) * n * a.tile);
    long long const tiles = (d + a.tile - 1) / a.tile;
    int const ctas_per_sm = smem <= 100 * 1024 ? 2 : 1;
    long long grid = static_cast<long long>(device_sm_count()) * ctas_per_sm;
    if (grid > tiles) grid = tiles;
    if (grid < 1) grid = 1;
    cudaStream_t const st = static_cast<cudaStream_t>(stream);

    cudaError_t err;
    int const g = static_cast<int>(grid);
    switch (np) {
        case 1: err = launch_coordwise<1, kRuleLinear, true, 4>(a, g, smem, st); break;
        case 4: err = dispatch_coordwise<4>(a, g, smem, st); break;
        case 8: err = dispatch_coordwise<8>(a, g, smem, st); break;
        case 12: err = dispatch_coordwise<12>(a, g, smem, st); break;
        case 16: err = dispatch_coordwise<16>(a, g, smem, st); break;
        case 20: err = dispatch_coordwise<20>(a, g, smem, st); break;
        case 24: err = dispatch_coordwise<24>(a, g, smem, st); break;
        case 28: err = dispatch_coordwise<28>(a, g, smem, st); break;
        case 32: err = dispatch_coordwise<32>(a, g, smem, st); break;
        case 48: err = dispatch_coordwise<48>(a, g, smem, st); break;
        default: err = dispatch_coordwise<64>(a, g, smem, st); break;
    }
    return cuda_check(err, "coordwise launch");
}
