// Shared declarations of the aggregation kernels: row tables, rule ids, the TMA-bulk tile
// ring used by every streaming kernel, register sorting networks and a counter-based RNG.
#pragma once

#include <cstdint>
#include <utility>
#include <cuda_runtime.h>

#include "ptx.cuh"

namespace gb {

constexpr int kMaxRows = 64;          // max number of gradients per aggregation call
constexpr int kMaxMirrors = 8;        // max peer copies written by the fused update epilogue
constexpr int kConsumerWarps = 8;
constexpr int kThreads = 32 * (1 + kConsumerWarps);  // warp 0 = TMA producer
constexpr int kStages = 3;

// n row pointers, passed by value in the kernel parameter space. Rows may live in the memory
// of another GPU (peer-mapped): the kernels only ever read them.
struct RowTable {
    const float* p[kMaxRows];
};

// Per-coordinate rules (the reference's K8/K9/K10 kernels plus the TF-only and new ones).
enum CoordRule : int {
    kRuleLinear = 0,             // out = Σ_i W[0][i]·x_i                (selection_average, K3)
    kRuleMedianLower = 1,        // torch.median semantics, NaN-propagating
    kRuleMedianFiniteUpper = 2,  // ⌊c/2⌋-th of the c finite values, 0 if none  (K10)
    kRuleTrimmedMean = 3,        // mean of sorted ranks [f, n-f)
    kRuleAveragedMedian = 4,     // mean of the β values closest to the (lower|upper) median (K8)
    kRuleCondense = 5,           // Bernoulli(p) ? lower median : x_0
};

// Optional optimizer step fused into the aggregation epilogue (torch.optim.SGD semantics),
// followed by stores of the updated parameter into up to kMaxMirrors peer arenas.
struct FusedUpdate {
    int mode;  // 0 = none, 1 = SGD(momentum, dampening, weight_decay, nesterov)
    int nesterov;
    int first_step;  // momentum buffer is initialised with the gradient on the first step
    int n_mirrors;
    float lr, momentum, dampening, weight_decay;
    float* param;
    float* mom;
    float* mirror[kMaxMirrors];
    float* multicast;  // NVLS multicast address of `param` (or null): publication by multimem.st
};

struct CoordArgs {
    RowTable rows;
    int n;            // number of input rows
    int rule;         // CoordRule
    int rank;         // f (trimmed mean) | β (averaged median)
    int upper;        // averaged median: use the upper median
    int theta;        // > 0: values fed to the rule are y = W·x  (W is theta × n, row-major)
    int tile;         // coordinates per ring stage
    int use_tma;      // 0: rows are not 16-byte aligned → direct global loads only
    float p;          // condense probability
    unsigned long long seed;
    long long d;
    const float* W;
    float* out;       // may be null when a fused update consumes the result
    FusedUpdate upd;
};

enum SelectMode : int {
    kSelectNone = 0,          // distances only
    kSelectKrum = 1,          // m best scores; score = Σ `neighbors` smallest (sqrt if !squared)
    kSelectBulyanPython = 2,  // static-score rounds (reference python semantics)
    kSelectBulyanNative = 3,  // pruned-distance score updates (reference native semantics)
    kSelectMda = 4,           // minimum-diameter subset of size n-f
    kSelectClosest = 5,       // the m rows (of the first n-1) closest to the LAST row (aksel)
};

struct SelectArgs {
    int mode, f, m, neighbors, squared;
};

struct PairwiseArgs {
    RowTable rows;
    int n;
    int tile;
    int use_tma;
    long long d;
    float* partials;        // scratch: gridDim.x × n × n floats
    unsigned int* counter;  // scratch: 1 word, must be 0 at launch (reset by the kernel)
    float* dist;            // out: n × n squared distances (+inf diagonal / non-finite)
    float* W;               // out: theta × n averaging weights (theta = 1 unless Bulyan)
    int* info;              // out: [0] = number of selection rounds, [1..] selected ids of round 0
    SelectArgs sel;
    unsigned char owner[36];   // pairwise_multi_kernel: consumer warp that owns block pair p (host plan, gb_pairwise_plan)
};

// ------------------------------------------------------------------- sorting networks ---

// Batcher odd-even merge sort on a register array of NP values (NP any size <= 64): the network of
// the next power of two, pruned of every comparator that touches a position >= NP (those positions
// would hold +inf padding, which such a comparator never moves): 19 / 42 / 63 / 103 / 132 / 162 / 191 /
// 384 / 543 compare-exchanges for NP = 8 / 12 / 16 / 20 / 24 / 28 / 32 / 48 / 64. The comparator list is a
// constexpr table and the network is expanded with a fold expression, so every register index is a
// constant expression (a partially unrolled loop nest would push the array to local memory).
// NaNs must have been replaced beforehand (fminf/fmaxf drop them).
constexpr int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

template <int NP>
struct OddEvenNetwork {
    static constexpr int kFull = next_pow2(NP);
    static constexpr int kMax = NP <= 2 ? 1 : NP * 9;  // >= comparator count for NP <= 64
    struct Table {
        int a[kMax];
        int b[kMax];
        int count;
    };
    static constexpr Table build() {
        Table t{};
        t.count = 0;
        for (int p = 1; p < kFull; p <<= 1)
            for (int k = p; k >= 1; k >>= 1)
                for (int j = k % p; j + k < kFull; j += 2 * k)
                    for (int i = 0; i < k; ++i)
                        if ((i + j) / (2 * p) == (i + j + k) / (2 * p) && i + j + k < NP) {
                            t.a[t.count] = i + j;
                            t.b[t.count] = i + j + k;
                            ++t.count;
                        }
        return t;
    }
    static constexpr Table table = build();
};

template <int NP, int C>
__device__ __forceinline__ void compare_exchange(float (&v)[NP]) {
    constexpr int a = OddEvenNetwork<NP>::table.a[C], b = OddEvenNetwork<NP>::table.b[C];
    float const lo = fminf(v[a], v[b]), hi = fmaxf(v[a], v[b]);
    v[a] = lo;
    v[b] = hi;
}

template <int NP, int C>
__device__ __forceinline__ void compare_exchange_pair(float (&key)[NP], float (&val)[NP]) {
    constexpr int a = OddEvenNetwork<NP>::table.a[C], b = OddEvenNetwork<NP>::table.b[C];
    bool const swap = key[a] > key[b];
    float const ka = key[a], kb = key[b], va = val[a], vb = val[b];
    key[a] = swap ? kb : ka;
    key[b] = swap ? ka : kb;
    val[a] = swap ? vb : va;
    val[b] = swap ? va : vb;
}

template <int NP, int... Cs>
__device__ __forceinline__ void sort_regs_seq(float (&v)[NP], std::integer_sequence<int, Cs...>) {
    (compare_exchange<NP, Cs>(v), ...);
}
template <int NP, int... Cs>
__device__ __forceinline__ void sort_pairs_seq(float (&key)[NP], float (&val)[NP], std::integer_sequence<int, Cs...>) {
    (compare_exchange_pair<NP, Cs>(key, val), ...);
}

template <int NP>
__device__ __forceinline__ void sort_regs(float (&v)[NP]) {
    sort_regs_seq<NP>(v, std::make_integer_sequence<int, OddEvenNetwork<NP>::table.count>{});
}

// Same network ordering (key, value) pairs by key.
template <int NP>
__device__ __forceinline__ void sort_pairs(float (&key)[NP], float (&val)[NP]) {
    sort_pairs_seq<NP>(key, val, std::make_integer_sequence<int, OddEvenNetwork<NP>::table.count>{});
}

// Bitonic merger: sorts a register array that is already BITONIC (here: |sorted values − their median|, which falls to zero at
// the median and rises after it; +inf padding beyond the count keeps that shape) with ⌈log2 NP⌉ half-cleaner stages — 52
// compare-exchanges at NP = 24 where the full network needs 132. Comparators reaching beyond NP are pruned like above.
template <int NP>
struct BitonicMerger {
    static constexpr int kFull = next_pow2(NP);
    static constexpr int kMax = NP * 6;
    struct Table {
        int a[kMax];
        int b[kMax];
        int count;
    };
    static constexpr Table build() {
        Table t{};
        t.count = 0;
        for (int k = kFull / 2; k >= 1; k >>= 1)
            for (int i = 0; i + k < NP; ++i)
                if ((i & k) == 0) {
                    t.a[t.count] = i;
                    t.b[t.count] = i + k;
                    ++t.count;
                }
        return t;
    }
    static constexpr Table table = build();
};
template <int NP, int C>
__device__ __forceinline__ void merge_exchange(float (&v)[NP]) {
    constexpr int a = BitonicMerger<NP>::table.a[C], b = BitonicMerger<NP>::table.b[C];
    float const lo = fminf(v[a], v[b]), hi = fmaxf(v[a], v[b]);
    v[a] = lo;
    v[b] = hi;
}
template <int NP, int... Cs>
__device__ __forceinline__ void merge_bitonic_seq(float (&v)[NP], std::integer_sequence<int, Cs...>) {
    (merge_exchange<NP, Cs>(v), ...);
}
template <int NP>
__device__ __forceinline__ void merge_bitonic(float (&v)[NP]) {
    merge_bitonic_seq<NP>(v, std::make_integer_sequence<int, BitonicMerger<NP>::table.count>{});
}

// v[k] for a runtime (warp-uniform) k without dynamic register indexing.
template <int NP>
__device__ __forceinline__ float pick_reg(float const (&v)[NP], int k) {
    float r = v[0];
#pragma unroll
    for (int i = 1; i < NP; ++i)
        if (i == k) r = v[i];
    return r;
}

// v sorted → its k-th value, for the ranks a median asks for. With the rank a compile-time constant only ONE output of the
// network is live, so the compiler drops every comparator (or comparator half) that cannot reach it: 193 instead of 264
// FMNMX at NP = 24, 91 instead of 126 at NP = 16. The count is warp-uniform, so the dispatch costs nothing; other ranks
// take the full network.
template <int NP, int K>
__device__ __forceinline__ float sorted_at(float (&v)[NP]) {
    sort_regs<NP>(v);
    return v[K];
}
template <int NP>
__device__ __forceinline__ float select_rank(float (&v)[NP], int k) {
    constexpr int kMid = (NP - 1) >> 1;
    if (k == kMid) return sorted_at<NP, kMid>(v);
    if constexpr (NP >= 4) {
        if (k == kMid - 1) return sorted_at<NP, kMid - 1>(v);
        if (k == kMid + 1) return sorted_at<NP, kMid + 1>(v);
    }
    sort_regs<NP>(v);
    return pick_reg<NP>(v, k);
}

// ------------------------------------------------------------------------ Philox RNG ---

// Philox4x32-10 keyed by `seed`, counter = (index, stream). Returns 4 uniform uint32.
__device__ __forceinline__ uint4 philox4x32(unsigned long long seed, unsigned long long index, unsigned int stream) {
    uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
    uint32_t c0 = static_cast<uint32_t>(index), c1 = static_cast<uint32_t>(index >> 32), c2 = stream, c3 = 0x9E3779B9u;
#pragma unroll
    for (int round = 0; round < 10; ++round) {
        uint32_t const hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t const hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t const n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ float u01(uint32_t bits) {  // [0, 1)
    return static_cast<float>(bits >> 8) * (1.0f / 16777216.0f);
}

// ---------------------------------------------------------------------- tile ring ---

// Shared-memory ring of `kStages` tiles, each n rows × `tile` floats, filled by one elected
// producer lane with cp.async.bulk (TMA 1-D) and drained by kConsumerWarps warps.
struct TileRing {
    uint64_t full[kStages];
    uint64_t empty[kStages];
};

__device__ __forceinline__ void ring_init(TileRing& ring) {
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            ptx::mbar_init(&ring.full[s], 1);
            ptx::mbar_init(&ring.empty[s], kConsumerWarps);
        }
        ptx::fence_mbar_init();
    }
    __syncthreads();
}

// Producer side: one lane streams this CTA's full tiles. `want(i)` tells whether row i is needed.
template <class Want>
__device__ __forceinline__ void ring_produce(TileRing& ring, float* stage_base, RowTable const& rows, int n, int tile, long long full_tiles,
                                             Want want) {
    size_t const stage_floats = static_cast<size_t>(n) * tile;
    uint32_t const row_bytes = static_cast<uint32_t>(tile) * 4u;
    int needed = 0;
    for (int i = 0; i < n; ++i) needed += want(i) ? 1 : 0;
    long long it = 0;
    for (long long t = blockIdx.x; t < full_tiles; t += gridDim.x, ++it) {
        int const s = static_cast<int>(it % kStages);
        uint32_t const phase = static_cast<uint32_t>((it / kStages) & 1);
        ptx::mbar_wait(&ring.empty[s], phase ^ 1u);
        ptx::mbar_arrive_expect_tx(&ring.full[s], row_bytes * static_cast<uint32_t>(needed));
        float* dst = stage_base + s * stage_floats;
        for (int i = 0; i < n; ++i)
            if (want(i)) ptx::bulk_g2s(dst + static_cast<size_t>(i) * tile, rows.p[i] + t * tile, row_bytes, &ring.full[s]);
    }
}

}  // namespace gb
