// pairwise_select: ONE streaming pass over n gradient rows that accumulates all n(n−1)/2
// squared distances by direct differences, followed — in the last CTA to finish — by a
// fixed-order reduction of the per-CTA partials and the whole on-device selection logic of
// Krum / Multi-Krum / Bulyan / MDA / Aksel, which emits an averaging-weight matrix W (θ×n).
// The second pass (`gb_coordwise` with W) produces the output; nothing returns to the host.
//
// Counterpart of the reference's K1 `squared_difference` + K2 `cub::DeviceReduce::Sum`
// launched n(n−1)/2 times, K4 `distances_finalize`, K5 `merge_sort_limited`, K6
// `compute_scores_prune_distances`, K7 `remove_smallest_scoring_gradient` and of the host
// sort / nth_element / combination search round-trips
// (/root/reference/pytorch_impl/libs/native/py_krum/krum.cu:68-158, py_bulyan/bulyan.cu:67-207,
//  256-322, py_brute/brute.cu:64-154, include/operations.cu.hpp:43-48):
//   * n·d words of HBM/NVLink traffic instead of ≈4·d·n(n−1)/2, 1 launch instead of n(n−1)+,
//     no cudaMalloc, no D2H/H2D, no n ≤ 23 limit;
//   * rows are streamed by TMA bulk copies into an mbarrier ring (peer-mapped rows work);
//   * warps own 4×4 row-block pairs: 8 LDS.128 feed 64 FADD+FFMA pairs per lane;
//   * per-warp shared accumulators + fixed-order CTA/grid reductions → bitwise run-to-run
//     deterministic distances.

#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <mutex>
#include <type_traits>

#include "../cpu/combinations.hpp"
#include "gar_common.cuh"
#include "gar_select.cuh"
#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

// Rows per register block: warps own KG×KG row-block pairs. Per lane and 128 coordinates a block pair costs 2·KG LDS.128 and
// 2·KG² FADD/FFMA pairs: at KG = 4 shared-memory reads and FP32 issue are balanced (both ≈50 % of the SM). (8-row blocks halve
// the shared-memory traffic per FMA but were slower than pairwise_multi_kernel with packed FP32 at every n — 0.33 against
// 0.26 ms at n = 16 — and are no longer instantiated.)

__device__ __forceinline__ void block_pair_of(int item, int B, int& bi, int& bj) {
    // enumerate (bi <= bj) row-block pairs in row-major order
    bi = 0;
    int rowlen = B;
    while (item >= rowlen) {
        item -= rowlen;
        ++bi;
        --rowlen;
    }
    bj = bi + item;
}

// acc += Σ over the 4 coordinates of (a − b)²   (4 FADD + 4 FFMA)
__device__ __forceinline__ float sq4_acc(float acc, float4 a, float4 b) {
    float const dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, dw = a.w - b.w;
    return fmaf(dw, dw, fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, acc))));
}

// Accumulate the 128-coordinate chunks first, first+stride, ... (lane = 4 coordinates) of the
// row-block pair whose rows start at `pa` / `pb` in shared memory (row pitch `tile` floats).
// FULL: all 4 rows of both blocks exist (no predicates). DIAG: bi == bj (upper triangle only).
template <int kG, bool FULL, bool DIAG>
__device__ __forceinline__ void accumulate_block(float (&acc)[kG][kG], float const* pa, float const* pb, int tile, int rows_a, int rows_b,
                                                 int first, int stride, int chunks) {
    for (int ch = first; ch < chunks; ch += stride) {
        float4 a[kG], b[kG];
#pragma unroll
        for (int r = 0; r < kG; ++r) {
            if (FULL || r < rows_a) a[r] = *reinterpret_cast<float4 const*>(pa + r * tile + ch * 128);
            else a[r] = make_float4(0.f, 0.f, 0.f, 0.f);
            if constexpr (DIAG) {
                b[r] = a[r];
            } else {
                if (FULL || r < rows_b) b[r] = *reinterpret_cast<float4 const*>(pb + r * tile + ch * 128);
                else b[r] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
#pragma unroll
        for (int r = 0; r < kG; ++r)
#pragma unroll
            for (int c = 0; c < kG; ++c)
                if (!DIAG || c > r) acc[r][c] = sq4_acc(acc[r][c], a[r], b[c]);
    }
}

// Packed FP32 (sm_100 FADD2 / FFMA2: two lanes of a 64-bit register pair per instruction) for the register-resident kernel: half
// the issue slots per squared difference, and the operands are register PAIRS — the scalar form subtracts .x from .x, .y from .y
// of two LDS.128 results, i.e. always two registers of equal index modulo 4. A sum is kept as (even coordinates, odd coordinates)
// and folded when it is flushed.
using pair2 = unsigned long long;
__device__ __forceinline__ pair2 pack2(float lo, float hi) {
    pair2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ pair2 sub2(pair2 a, pair2 b) {
    pair2 r;
    asm("sub.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ pair2 square_acc2(pair2 d, pair2 acc) {
    pair2 r;
    asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(r) : "l"(d), "l"(acc));
    return r;
}
__device__ __forceinline__ float fold2(pair2 v) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
    return lo + hi;
}
template <bool FULL, bool DIAG>
__device__ __forceinline__ void accumulate_block_packed(pair2 (&acc)[4][4], float const* pa, float const* pb, int tile, int rows_a, int rows_b,
                                                        int first, int stride, int chunks) {
    for (int ch = first; ch < chunks; ch += stride) {
        pair2 a01[4], a23[4], b01[4], b23[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (FULL || r < rows_a) v = *reinterpret_cast<float4 const*>(pa + r * tile + ch * 128);
            a01[r] = pack2(v.x, v.y);
            a23[r] = pack2(v.z, v.w);
            if constexpr (DIAG) {
                b01[r] = a01[r];
                b23[r] = a23[r];
            } else {
                float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
                if (FULL || r < rows_b) w = *reinterpret_cast<float4 const*>(pb + r * tile + ch * 128);
                b01[r] = pack2(w.x, w.y);
                b23[r] = pack2(w.z, w.w);
            }
        }
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (!DIAG || c > r) {
                    acc[r][c] = square_acc2(sub2(a01[r], b01[c]), acc[r][c]);
                    acc[r][c] = square_acc2(sub2(a23[r], b23[c]), acc[r][c]);
                }
    }
}
template <int kG>
__device__ __forceinline__ void flush(float (&acc)[kG][kG], int bi, int bj, int n, float* slot);
__device__ __forceinline__ void flush_packed(pair2 (&acc)[4][4], int bi, int bj, int n, float* slot) {
    float folded[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            folded[r][c] = fold2(acc[r][c]);
            acc[r][c] = 0ull;
        }
    flush<4>(folded, bi, bj, n, slot);
}

// Dispatch on (all rows present, diagonal block). `src` = stage base with the lane offset added.
template <int kG>
__device__ __forceinline__ void accumulate_pair(float (&acc)[kG][kG], float const* src, int tile, int n, int bi, int bj, int first,
                                                int stride, int chunks) {
    float const* pa = src + bi * kG * tile;
    float const* pb = src + bj * kG * tile;
    int const rows_a = min(kG, n - bi * kG), rows_b = min(kG, n - bj * kG);
    bool const full = rows_a == kG && rows_b == kG;
    if (bi == bj) {
        if (full) accumulate_block<kG, true, true>(acc, pa, pb, tile, rows_a, rows_b, first, stride, chunks);
        else accumulate_block<kG, false, true>(acc, pa, pb, tile, rows_a, rows_b, first, stride, chunks);
    } else {
        if (full) accumulate_block<kG, true, false>(acc, pa, pb, tile, rows_a, rows_b, first, stride, chunks);
        else accumulate_block<kG, false, false>(acc, pa, pb, tile, rows_a, rows_b, first, stride, chunks);
    }
}

__device__ __forceinline__ void accumulate_pair_packed(pair2 (&acc)[4][4], float const* src, int tile, int n, int bi, int bj, int first, int stride,
                                                       int chunks) {
    float const* pa = src + bi * 4 * tile;
    float const* pb = src + bj * 4 * tile;
    int const rows_a = min(4, n - bi * 4), rows_b = min(4, n - bj * 4);
    bool const full = rows_a == 4 && rows_b == 4;
    if (bi == bj) {
        if (full) accumulate_block_packed<true, true>(acc, pa, pb, tile, rows_a, rows_b, first, stride, chunks);
        else accumulate_block_packed<false, true>(acc, pa, pb, tile, rows_a, rows_b, first, stride, chunks);
    } else {
        if (full) accumulate_block_packed<true, false>(acc, pa, pb, tile, rows_a, rows_b, first, stride, chunks);
        else accumulate_block_packed<false, false>(acc, pa, pb, tile, rows_a, rows_b, first, stride, chunks);
    }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Add this warp's register accumulators of block pair (bi, bj) into its private n×n slot.
template <int kG>
__device__ __forceinline__ void flush(float (&acc)[kG][kG], int bi, int bj, int n, float* slot) {
    int const lane = threadIdx.x & 31;
#pragma unroll
    for (int r = 0; r < kG; ++r)
#pragma unroll
        for (int c = 0; c < kG; ++c) {
            float const total = warp_sum(acc[r][c]);
            int const i = bi * kG + r, j = bj * kG + c;
            if (lane == 0 && i < n && j < n && i < j) slot[i * n + j] += total;
            acc[r][c] = 0.0f;
        }
}

// Coordinates that do not fill a tile (and rows the bulk copies cannot take: unaligned) — direct global loads, one coordinate
// per lane, every block pair flushed into the warp's slot.
template <int kG>
__device__ __forceinline__ void remainder_pass(PairwiseArgs const& a, long long done, int cw, int lane, int B, int BP, float* slot) {
    long long const rest = a.d - done;
    if (rest <= 0) return;
    int const n = a.n;
    long long const per_cta = (rest + gridDim.x - 1) / gridDim.x;
    long long const lo = done + per_cta * blockIdx.x;
    long long const hi = min(a.d, lo + per_cta);
    float acc[kG][kG];
#pragma unroll
    for (int r = 0; r < kG; ++r)
#pragma unroll
        for (int c = 0; c < kG; ++c) acc[r][c] = 0.0f;
    for (int p = cw; p < BP; p += kConsumerWarps) {
        int bi, bj;
        block_pair_of(p, B, bi, bj);
        for (long long x = lo + lane; x < hi; x += 32) {
            float va[kG], vb[kG];
#pragma unroll
            for (int r = 0; r < kG; ++r) {
                int const ra = bi * kG + r, rb = bj * kG + r;
                va[r] = ra < n ? a.rows.p[ra][x] : 0.0f;
                vb[r] = rb < n ? a.rows.p[rb][x] : 0.0f;
            }
#pragma unroll
            for (int r = 0; r < kG; ++r)
#pragma unroll
                for (int c = 0; c < kG; ++c)
                    if (bi != bj || c > r) {
                        float const df = va[r] - vb[c];
                        acc[r][c] = fmaf(df, df, acc[r][c]);
                    }
        }
        flush<kG>(acc, bi, bj, n, slot);
    }
}

// After the streaming loop: CTA partial = Σ over warps in fixed order → global scratch; the last CTA to finish reduces the
// partials in CTA order (double), symmetrises, maps non-finite values to +inf and runs the selection.
__device__ __forceinline__ void finish_cta(PairwiseArgs const& a, float* slots, unsigned char* smem_raw, unsigned int& is_last) {
    int const n = a.n, nn = n * n;
    __syncthreads();
    float* mine = a.partials + static_cast<size_t>(blockIdx.x) * nn;
    for (int e = threadIdx.x; e < nn; e += blockDim.x) {
        float s = 0.0f;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) s += slots[w * nn + e];
        mine[e] = s;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = (atomicAdd(a.counter, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // Σ over the CTAs' partials, in a fixed order (double): the whole GPU waits for this one CTA, so the sum of each of the
    // n(n−1)/2 pairs is cut into S segments (b ≡ seg mod S) taken by different threads, each with 4 independent accumulators —
    // at n = 8 that is 37 loads behind one another per thread where one thread per matrix entry had 296.
    int const pairs = n * (n - 1) / 2;
    int const S = pairs == 0 ? 1 : max(1, min(8, static_cast<int>(blockDim.x) / pairs));
    double* const red = reinterpret_cast<double*>(smem_raw);       // pairs × S segment sums (slots and stages are idle by now)
    unsigned const G = gridDim.x;
    for (int w = threadIdx.x; w < pairs * S; w += blockDim.x) {
        int const pr = w / S, seg = w - pr * S;
        int i = 0, rest = pr;                                      // pair index → (i < j), row-major over the upper triangle
        while (rest >= n - 1 - i) { rest -= n - 1 - i; ++i; }
        int const src = i * n + (i + 1 + rest);
        double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
        unsigned b = static_cast<unsigned>(seg);
        for (; b + 3u * S < G; b += 4u * S) {
            acc0 += static_cast<double>(__ldcg(a.partials + static_cast<size_t>(b) * nn + src));
            acc1 += static_cast<double>(__ldcg(a.partials + static_cast<size_t>(b + S) * nn + src));
            acc2 += static_cast<double>(__ldcg(a.partials + static_cast<size_t>(b + 2u * S) * nn + src));
            acc3 += static_cast<double>(__ldcg(a.partials + static_cast<size_t>(b + 3u * S) * nn + src));
        }
        for (; b < G; b += S) acc0 += static_cast<double>(__ldcg(a.partials + static_cast<size_t>(b) * nn + src));
        red[w] = (acc0 + acc1) + (acc2 + acc3);
    }
    __syncthreads();
    for (int e = threadIdx.x; e < nn; e += blockDim.x) {
        int const i = e / n, j = e % n;
        float out = INFINITY;
        if (i != j) {
            int const lo = min(i, j), hi = max(i, j);
            int const pr = lo * n - lo * (lo + 1) / 2 + (hi - lo - 1);
            double total = 0.0;
            for (int seg = 0; seg < S; ++seg) total += red[pr * S + seg];
            float const f = static_cast<float>(total);
            out = (fabsf(f) <= FLT_MAX) ? f : INFINITY;
        }
        a.dist[e] = out;
    }
    if (threadIdx.x == 0) *a.counter = 0u;  // ready for the next launch
    __syncthreads();
    run_selection(a, *reinterpret_cast<SelectScratch*>(smem_raw));
}

// Upper bound on the CTAs of the next gb_pairwise_select launches of this thread (0 = fill the GPU). A distance pass that runs
// in the BACKGROUND of other work (the tail coordinates' distances during the backward pass) must not take every SM, or the
// work it hides behind stalls.
thread_local int g_pairwise_cta_limit = 0;

// ------------------------------------------------------------------------- kernel ---

template <int kG, bool PACKED = false>
__global__ void __launch_bounds__(kThreads) pairwise_kernel(const __grid_constant__ PairwiseArgs a) {
    static_assert(!PACKED || kG == 4, "packed FP32 accumulation is written for 4-row blocks");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ TileRing ring;
    __shared__ unsigned int is_last;
    int const n = a.n;
    int const nn = n * n;
    float* slots = reinterpret_cast<float*>(smem_raw);            // kConsumerWarps × n × n
    int const slot_floats = ((kConsumerWarps * nn + 31) / 32) * 32;
    float* stages = slots + slot_floats;                           // kStages × n × tile
    int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int const tile = a.tile;
    long long const full_tiles = a.use_tma ? a.d / tile : 0;

    for (int i = threadIdx.x; i < kConsumerWarps * nn; i += blockDim.x) slots[i] = 0.0f;
    ring_init(ring);

    int const B = (n + kG - 1) / kG;
    int const BP = B * (B + 1) / 2;

    if (warp == 0) {
        if (full_tiles > 0 && ptx::elect_one()) ring_produce(ring, stages, a.rows, n, tile, full_tiles, [](int) { return true; });
    } else {
        int const cw = warp - 1;  // consumer warp id
        float* slot = slots + cw * nn;
        using Acc = std::conditional_t<PACKED, pair2, float>;
        Acc acc[kG][kG];
#pragma unroll
        for (int r = 0; r < kG; ++r)
#pragma unroll
            for (int c = 0; c < kG; ++c) acc[r][c] = Acc(0);
        size_t const stage_floats = static_cast<size_t>(n) * tile;
        int const chunks = tile / 128;

        // Static ownership when there are at most as many block pairs as warps: warp cw owns
        // block pair (cw % BP) for the whole kernel and keeps its sums in registers.
        bool const fixed = BP <= kConsumerWarps;
        int my_bi = 0, my_bj = 0, my_first = 0, my_stride = 1;
        if (fixed) {
            // distribute the 8 warps over block pairs proportionally to their cost (16 vs 6 pairs):
            // each warp in turn joins the block pair with the highest cost per current owner
            int owners[kConsumerWarps];
            int count[kConsumerWarps];
            for (int p = 0; p < BP; ++p) count[p] = 0;
            for (int w = 0; w < kConsumerWarps; ++w) {
                int pick = 0;
                float worst = -1.0f;
                for (int p = 0; p < BP; ++p) {
                    int bi, bj;
                    block_pair_of(p, B, bi, bj);
                    float const load = count[p] == 0 ? 1e9f : (bi == bj ? 0.5f * kG * (kG - 1) : 1.0f * kG * kG) / count[p];
                    if (load > worst) {
                        worst = load;
                        pick = p;
                    }
                }
                owners[w] = pick;
                ++count[pick];
            }
            int const mine = owners[cw];
            block_pair_of(mine, B, my_bi, my_bj);
            my_stride = count[mine];
            for (int w = 0; w < cw; ++w) my_first += (owners[w] == mine) ? 1 : 0;
        }

        long long it = 0;
        for (long long t = blockIdx.x; t < full_tiles; t += gridDim.x, ++it) {
            int const s = static_cast<int>(it % kStages);
            uint32_t const phase = static_cast<uint32_t>((it / kStages) & 1);
            ptx::mbar_wait(&ring.full[s], phase);
            float const* src = stages + s * stage_floats + lane * 4;
            if (fixed) {
                if constexpr (PACKED) accumulate_pair_packed(acc, src, tile, n, my_bi, my_bj, my_first, my_stride, chunks);
                else accumulate_pair<kG>(acc, src, tile, n, my_bi, my_bj, my_first, my_stride, chunks);
            } else {
                for (int p = cw; p < BP; p += kConsumerWarps) {
                    int bi, bj;
                    block_pair_of(p, B, bi, bj);
                    if constexpr (PACKED) {
                        accumulate_pair_packed(acc, src, tile, n, bi, bj, 0, 1, chunks);
                        flush_packed(acc, bi, bj, n, slot);
                    } else {
                        accumulate_pair<kG>(acc, src, tile, n, bi, bj, 0, 1, chunks);
                        flush<kG>(acc, bi, bj, n, slot);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&ring.empty[s]);
        }
        if (fixed) {
            if constexpr (PACKED) flush_packed(acc, my_bi, my_bj, n, slot);
            else flush<kG>(acc, my_bi, my_bj, n, slot);
        }

        remainder_pass<kG>(a, full_tiles * tile, cw, lane, B, BP, slot);
    }
    finish_cta(a, slots, smem_raw, is_last);
}

// ---------------------------------------------------- 13..32 rows: register-resident sums ---
//
// With more block pairs than warps the kernel above flushes a warp's 16 sums after every (tile, block pair) — 80 shuffles
// against the 16 LDS.128 + 256 FP32 instructions of a 256-coordinate tile, and one lane issuing n one-kilobyte copies per
// stage. This variant (4-row blocks, B <= 8 → at most 36 block pairs) instead
//   * gives every warp up to QO off-diagonal and QD diagonal block pairs for the whole kernel (longest-processing-time
//     assignment made on the host, plan_multi: 16 against 6 row pairs) and keeps all their sums in registers — one flush per kernel. The slots are typed
//     so that the unrolled loop holds QO + QD copies of the block code, not 2·(QO + QD): measured at n = 23, ten copies ran
//     at 0.75 ms and twelve at 0.89 ms — the loop has to stay inside the instruction cache;
//   * pads each stage to 4·B rows whose phantom rows are zeroed once, so every block takes the predicate-free path;
//   * sizes the tile for one CTA per SM (≈200 KB of shared memory: 640 coordinates at n = 23) and lets lane i of the
//     producer warp issue row i's copy.
constexpr int kMultiRows = 32;
constexpr int kMultiPairs = kMultiRows / 4 * (kMultiRows / 4 + 1) / 2;

template <int QO, int QD, bool PACKED>
__global__ void __launch_bounds__(kThreads, 1) pairwise_multi_kernel(const __grid_constant__ PairwiseArgs a) {
    constexpr int kG = 4;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ TileRing ring;
    __shared__ unsigned int is_last;
    int const n = a.n;
    int const nn = n * n;
    float* slots = reinterpret_cast<float*>(smem_raw);
    int const slot_floats = ((kConsumerWarps * nn + 31) / 32) * 32;
    float* stages = slots + slot_floats;
    int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int const tile = a.tile;
    long long const full_tiles = a.use_tma ? a.d / tile : 0;
    int const B = (n + kG - 1) / kG;
    int const BP = B * (B + 1) / 2;
    int const padded = B * kG;
    size_t const stage_floats = static_cast<size_t>(padded) * tile;

    for (int i = threadIdx.x; i < kConsumerWarps * nn; i += blockDim.x) slots[i] = 0.0f;
    for (int s = 0; s < kStages; ++s)
        for (int i = threadIdx.x; i < (padded - n) * tile; i += blockDim.x) stages[s * stage_floats + static_cast<size_t>(n) * tile + i] = 0.0f;
    ring_init(ring);

    if (warp == 0) {
        uint32_t const row_bytes = static_cast<uint32_t>(tile) * 4u;
        float const* const row = lane < n ? a.rows.p[lane] : nullptr;
        long long it = 0;
        for (long long t = blockIdx.x; t < full_tiles; t += gridDim.x, ++it) {
            int const s = static_cast<int>(it % kStages);
            uint32_t const phase = static_cast<uint32_t>((it / kStages) & 1);
            ptx::mbar_wait(&ring.empty[s], phase ^ 1u);
            if (lane == 0) ptx::mbar_arrive_expect_tx(&ring.full[s], row_bytes * static_cast<uint32_t>(n));
            __syncwarp();
            if (lane < n) ptx::bulk_g2s(stages + s * stage_floats + static_cast<size_t>(lane) * tile, row + t * tile, row_bytes, &ring.full[s]);
        }
    } else {
        int const cw = warp - 1;
        float* slot = slots + cw * nn;
        using Acc = std::conditional_t<PACKED, pair2, float>;
        Acc acc_off[QO][kG][kG], acc_diag[QD][kG][kG];
        int off_a[QO], off_b[QO], diag_a[QD];
#pragma unroll
        for (int q = 0; q < QO; ++q) {
            off_a[q] = off_b[q] = 0;
#pragma unroll
            for (int r = 0; r < kG; ++r)
#pragma unroll
                for (int c = 0; c < kG; ++c) acc_off[q][r][c] = Acc(0);
        }
#pragma unroll
        for (int q = 0; q < QD; ++q) {
            diag_a[q] = 0;
#pragma unroll
            for (int r = 0; r < kG; ++r)
#pragma unroll
                for (int c = 0; c < kG; ++c) acc_diag[q][r][c] = Acc(0);
        }
        int n_off = 0, n_diag = 0;
        for (int p = 0; p < BP; ++p)
            if (a.owner[p] == cw) {
                int x, y;
                block_pair_of(p, B, x, y);
                if (x == y) {
#pragma unroll
                    for (int q = 0; q < QD; ++q)
                        if (q == n_diag) diag_a[q] = x;
                    ++n_diag;
                } else {
#pragma unroll
                    for (int q = 0; q < QO; ++q)
                        if (q == n_off) { off_a[q] = x; off_b[q] = y; }
                    ++n_off;
                }
            }
        int const chunks = tile / 128;
        long long it = 0;
        for (long long t = blockIdx.x; t < full_tiles; t += gridDim.x, ++it) {
            int const s = static_cast<int>(it % kStages);
            uint32_t const phase = static_cast<uint32_t>((it / kStages) & 1);
            ptx::mbar_wait(&ring.full[s], phase);
            float const* src = stages + s * stage_floats + lane * 4;
#pragma unroll
            for (int q = 0; q < QO; ++q)
                if (q < n_off)
                {
                    float const* pa = src + off_a[q] * kG * tile;
                    float const* pb = src + off_b[q] * kG * tile;
                    if constexpr (PACKED) accumulate_block_packed<true, false>(acc_off[q], pa, pb, tile, 4, 4, 0, 1, chunks);
                    else accumulate_block<kG, true, false>(acc_off[q], pa, pb, tile, kG, kG, 0, 1, chunks);
                }
#pragma unroll
            for (int q = 0; q < QD; ++q)
                if (q < n_diag) {
                    float const* pa = src + diag_a[q] * kG * tile;
                    if constexpr (PACKED) accumulate_block_packed<true, true>(acc_diag[q], pa, pa, tile, 4, 4, 0, 1, chunks);
                    else accumulate_block<kG, true, true>(acc_diag[q], pa, pa, tile, kG, kG, 0, 1, chunks);
                }
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&ring.empty[s]);
        }
#pragma unroll
        for (int q = 0; q < QO; ++q)
            if (q < n_off) {
                if constexpr (PACKED) flush_packed(acc_off[q], off_a[q], off_b[q], n, slot);
                else flush<kG>(acc_off[q], off_a[q], off_b[q], n, slot);
            }
#pragma unroll
        for (int q = 0; q < QD; ++q)
            if (q < n_diag) {
                if constexpr (PACKED) flush_packed(acc_diag[q], diag_a[q], diag_a[q], n, slot);
                else flush<kG>(acc_diag[q], diag_a[q], diag_a[q], n, slot);
            }
        remainder_pass<kG>(a, full_tiles * tile, cw, lane, B, BP, slot);
    }
    finish_cta(a, slots, smem_raw, is_last);
}

// GARFIELD_PAIRWISE_MULTI=0 sends 13..32 rows through the general kernel again (A/B measurements).
bool multi_enabled() {
    static int const on = [] { char const* v = std::getenv("GARFIELD_PAIRWISE_MULTI"); return (v != nullptr && v[0] == '0') ? 0 : 1; }();
    return on != 0;
}

// Tile of the register-resident variant: the largest multiple of 128 coordinates whose three padded stages fit ≈200 KB.
int multi_tile(int n) {
    int const padded = (n + 3) / 4 * 4;
    int const slot_bytes = ((kConsumerWarps * n * n + 31) / 32) * 32 * 4;
    int tile = (200 * 1024 - slot_bytes) / (kStages * padded * 4) / 128 * 128;
    if (tile < 128) tile = 128;
    if (tile > 1024) tile = 1024;
    return tile;
}

// Slot counts per block count B = ⌈n/4⌉ (B(B−1)/2 off-diagonal and B diagonal pairs on 8 warps):
//   B = 4: 6 + 4 → (1, 2);  B = 5: 10 + 5 → (2, 1);  B = 6: 15 + 6 → (2, 1);  B = 7: 21 + 7 → (3, 1);  B = 8: 28 + 8 → (4, 1)
// Longest-processing-time assignment of the block pairs to the 8 consumer warps under the slot limits (qo off-diagonal, qd
// diagonal pairs per warp); off-diagonal pairs (16 row pairs) first, then the diagonal ones (6). false = some pair finds no warp.
bool plan_multi(int n, int qo, int qd, unsigned char* owner) {
    int const B = (n + 3) / 4;
    float load[kConsumerWarps] = {};
    int held_off[kConsumerWarps] = {}, held_diag[kConsumerWarps] = {};
    for (int pass = 0; pass < 2; ++pass) {
        int p = 0;
        for (int bi = 0; bi < B; ++bi)
            for (int bj = bi; bj < B; ++bj, ++p) {      // same enumeration as block_pair_of
                bool const diag = bi == bj;
                if (diag != (pass == 1)) continue;
                int best = -1;
                for (int w = 0; w < kConsumerWarps; ++w) {
                    bool const room = diag ? held_diag[w] < qd : held_off[w] < qo;
                    if (room && (best < 0 || load[w] < load[best])) best = w;
                }
                if (best < 0) return false;
                owner[p] = static_cast<unsigned char>(best);
                load[best] += diag ? 6.0f : 16.0f;
                ++(diag ? held_diag[best] : held_off[best]);
            }
    }
    return true;
}
void multi_slots(int n, int& qo, int& qd) {
    int const B = (n + 3) / 4;
    qo = B <= 4 ? 1 : B <= 6 ? 2 : B == 7 ? 3 : 4;
    qd = B <= 4 ? 2 : 1;
}

using MultiKernel = void (*)(PairwiseArgs);
// (two CTAs per SM with 256-coordinate tiles were measured too: no faster, 0.74 against 0.72 ms at n = 23)
// GARFIELD_PW_PACKED=0: scalar FADD / FFMA everywhere (A/B measurements).
bool packed_enabled() {
    static int const on = [] { char const* v = std::getenv("GARFIELD_PW_PACKED"); return (v != nullptr && v[0] == '0') ? 0 : 1; }();
    return on != 0;
}
MultiKernel multi_kernel_for(int n) {
    int const B = (n + 3) / 4;
    if (packed_enabled()) {
        if (B <= 4) return pairwise_multi_kernel<1, 2, true>;
        if (B <= 6) return pairwise_multi_kernel<2, 1, true>;
        if (B == 7) return pairwise_multi_kernel<3, 1, true>;
    }
    if (B <= 4) return pairwise_multi_kernel<1, 2, false>;
    if (B <= 6) return pairwise_multi_kernel<2, 1, false>;
    if (B == 7) return pairwise_multi_kernel<3, 1, false>;
    return pairwise_multi_kernel<4, 1, false>;
}

}  // namespace
}  // namespace gb

extern "C" GB_API int64_t gb_pairwise_scratch_bytes(int n) {
    // partials for the largest grid this library launches (2 CTAs per SM) + the counter word
    int64_t const grid = 2 * static_cast<int64_t>(gb::device_sm_count());
    return 256 + grid * n * n * static_cast<int64_t>(sizeof(float));
}

extern "C" GB_API int gb_pairwise_select(float const* const* rows, int n, int64_t d, int mode, int f, int m, int neighbors, int squared,
                                         void* scratch, int64_t scratch_bytes, float* dist, float* W, int* info, void* stream) {
    using namespace gb;
    if (n < 1 || n > kMaxRows) return cuda_fail("pairwise: need 1 <= n <= 64 rows");
    if (d < 1) return cuda_fail("pairwise: empty rows");
    if (dist == nullptr || scratch == nullptr) return cuda_fail("pairwise: missing output/scratch buffer");
    if (mode != kSelectNone && (W == nullptr || info == nullptr)) return cuda_fail("pairwise: selection needs W and info buffers");
    if (char const* bad = selection_error(mode, n, f, m, neighbors)) {
        static thread_local char msg[160];
        std::snprintf(msg, sizeof(msg), "pairwise: %s", bad);
        return cuda_fail(msg);
    }

    PairwiseArgs a{};
    for (int i = 0; i < n; ++i) a.rows.p[i] = rows[i];
    a.n = n; a.d = d; a.dist = dist; a.W = W; a.info = info;
    a.sel.mode = mode; a.sel.f = f; a.sel.m = m; a.sel.neighbors = neighbors; a.sel.squared = squared;
    a.use_tma = rows_aligned(a.rows, n) ? 1 : 0;
    // 13..32 rows: register-resident sums. Measured at d = 23.5 M (general kernel -> this one): n = 13 0.61 -> 0.26 ms, 16 0.33 -> 0.26,
    // 20 1.09 -> 0.42, 23 1.58 -> 0.58, 24 0.86 -> 0.59, 28 2.61 -> 0.92, 32 5.79 -> 1.22.
    bool multi = n > 12 && n <= kMultiRows && multi_enabled();
    if (multi) {
        int qo, qd;
        multi_slots(n, qo, qd);
        multi = plan_multi(n, qo, qd, a.owner);      // no feasible plan (cannot happen for 13..32 rows): general kernel
    }
    // packed FP32 in the general kernel pays from 9 rows on (n = 10 0.256 -> 0.214 ms, 12 0.223 -> 0.207; n = 8 0.136 -> 0.142)
    bool const packed = n > 8 && packed_enabled();
    a.tile = multi ? multi_tile(n) : pick_tile(n);

    int const nn = n * n;
    int const slot_floats = ((kConsumerWarps * nn + 31) / 32) * 32;
    int const stage_rows = multi ? (n + 3) / 4 * 4 : n;
    size_t smem = sizeof(float) * (static_cast<size_t>(slot_floats) + static_cast<size_t>(kStages) * stage_rows * a.tile);
    if (smem < sizeof(SelectScratch)) smem = sizeof(SelectScratch);
    long long const tiles = (d + a.tile - 1) / a.tile;
    int const ctas_per_sm = (smem <= 100 * 1024 && !multi) ? 2 : 1;
    long long grid = static_cast<long long>(device_sm_count()) * ctas_per_sm;
    if (g_pairwise_cta_limit > 0 && grid > g_pairwise_cta_limit) grid = g_pairwise_cta_limit;   // background launch: leave SMs to others
    if (grid > tiles) grid = tiles;
    if (grid < 1) grid = 1;
    int64_t const need = 256 + grid * nn * static_cast<int64_t>(sizeof(float));
    if (scratch_bytes < need) return cuda_fail("pairwise: scratch buffer too small (see gb_pairwise_scratch_bytes)");
    a.counter = static_cast<unsigned int*>(scratch);
    a.partials = reinterpret_cast<float*>(static_cast<unsigned char*>(scratch) + 256);

    auto kern = multi ? multi_kernel_for(n) : (packed ? pairwise_kernel<4, true> : pairwise_kernel<4, false>);
    if (cuda_check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)), "pairwise smem attribute")) return 1;
    kern<<<static_cast<int>(grid), kThreads, smem, static_cast<cudaStream_t>(stream)>>>(a);
    return cuda_check(cudaGetLastError(), "pairwise launch");
}


// What gb_pairwise_select would launch for n rows (host-only; no device work): kind 0 = general kernel (4-row blocks, scalar),
// 1 = general kernel with packed FP32, 2 = register-resident kernel. For kind 2, `slots` receives (off-diagonal, diagonal) pairs
// per warp and `owner` the warp of each of the B(B+1)/2 block pairs (B = ⌈n/4⌉, pairs enumerated row-major over bi <= bj).
extern "C" GB_API int gb_pairwise_plan(int n, int* kind, int* tile, int* slots, unsigned char* owner) {
    using namespace gb;
    if (n < 1 || n > kMaxRows || kind == nullptr || tile == nullptr) return cuda_fail("pairwise_plan: bad arguments");
    bool multi = n > 12 && n <= kMultiRows && multi_enabled();
    int qo = 0, qd = 0;
    unsigned char local[kMultiPairs] = {};
    if (multi) {
        multi_slots(n, qo, qd);
        multi = plan_multi(n, qo, qd, local);
    }
    *kind = multi ? 2 : (n > 8 && packed_enabled() ? 1 : 0);
    *tile = multi ? multi_tile(n) : pick_tile(n);
    if (slots != nullptr) { slots[0] = qo; slots[1] = qd; }
    if (owner != nullptr && multi) {
        int const B = (n + 3) / 4;
        for (int p = 0; p < B * (B + 1) / 2; ++p) owner[p] = local[p];
    }
    return 0;
}

extern "C" GB_API int gb_pairwise_cta_limit(int max_ctas) {
    int const previous = gb::g_pairwise_cta_limit;
    gb::g_pairwise_cta_limit = max_ctas < 0 ? 0 : max_ctas;
    return previous;
}

// ------------------------------------------------------------------ selection only ---
//
// Coordinate-sharded aggregation (garfield_b200/parallel/engine.py): every GPU computes the
// PARTIAL squared distances of its coordinate slice, the n×n partials are summed across GPUs
// (n² floats), and each GPU then runs this one-CTA kernel on the summed matrix to obtain the
// (identical) averaging weights for its slice. Same selection code as the fused kernels.
namespace gb {
namespace {
struct SelectOnlyArgs {
    int n;
    float* dist;
    float* W;
    int* info;
    SelectArgs sel;
};
__global__ void __launch_bounds__(kThreads) select_only_kernel(const __grid_constant__ SelectOnlyArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    int const nn = a.n * a.n;
    for (int e = threadIdx.x; e < nn; e += blockDim.x) {  // normalise: +inf diagonal, non-finite → +inf
        float const v = a.dist[e];
        bool const diag = (e / a.n) == (e % a.n);
        a.dist[e] = (!diag && fabsf(v) <= FLT_MAX) ? v : INFINITY;
    }
    __syncthreads();
    run_selection(a, *reinterpret_cast<SelectScratch*>(smem_raw));
}
}  // namespace
}  // namespace gb

extern "C" GB_API int gb_select_only(float* dist, int n, int mode, int f, int m, int neighbors, int squared, float* W, int* info, void* stream) {
    using namespace gb;
    if (n < 1 || n > kMaxRows) return cuda_fail("select_only: need 1 <= n <= 64 rows");
    if (dist == nullptr || W == nullptr || info == nullptr) return cuda_fail("select_only: missing buffers");
    if (mode == kSelectNone) return cuda_fail("select_only: no selection mode given");
    if (char const* bad = selection_error(mode, n, f, m, neighbors)) {
        static thread_local char msg[160];
        std::snprintf(msg, sizeof(msg), "select_only: %s", bad);
        return cuda_fail(msg);
    }
    SelectOnlyArgs a{};
    a.n = n; a.dist = dist; a.W = W; a.info = info;
    a.sel.mode = mode; a.sel.f = f; a.sel.m = m; a.sel.neighbors = neighbors; a.sel.squared = squared;
    auto kern = select_only_kernel;
    size_t const smem = sizeof(SelectScratch);
    if (cuda_check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)), "select_only smem attribute")) return 1;
    kern<<<1, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(a);
    return cuda_check(cudaGetLastError(), "select_only launch");
}

// ------------------------------------------------------------- allocation-free variant ---
//
// Some hosts of this library may not allocate device memory at the moment they need an
// aggregation: a serving or training process that handed the whole HBM to its own pool at
// start-up (a later cudaMalloc fails or fragments that pool), a steady-state loop running under
// a fixed memory plan, or a latency-critical path that cannot afford cudaMalloc's device-wide
// synchronisation. For them the scratch of the pairwise kernel (per-CTA partial distance
// matrices + the completion counter) lives in this fixed array inside the library image instead
// of a caller-provided buffer. Every GPU that loads the library carries the array until exit.
//
// Limit: kStaticMaxRows gradients per call (the array holds 256 + 296·16·16·4 bytes, i.e. the
// partials of the largest grid this library launches on a B200 for n <= 16).
constexpr int kStaticMaxRows = 16;
constexpr long long kStaticScratchBytes = 256 + 296LL * kStaticMaxRows * kStaticMaxRows * static_cast<long long>(sizeof(float));
static __device__ __align__(256) unsigned char g_pairwise_static_scratch[kStaticScratchBytes];

namespace {
// Device address of the array, looked up once per device. In host code the symbol above is only
// a stand-in; the address a kernel can dereference comes from cudaGetSymbolAddress.
std::mutex g_static_mutex;
void* g_static_scratch_ptr[64] = {nullptr};

int static_scratch_for_current_device(void** out) {
    int dev = 0;
    if (gb::cuda_check(cudaGetDevice(&dev), "no-alloc pairwise: cudaGetDevice")) return 1;
    if (dev < 0 || dev >= 64) return gb::cuda_fail("no-alloc pairwise: device ordinal out of range (0..63)");
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (g_static_scratch_ptr[dev] == nullptr) {
        void* ptr = nullptr;
        if (gb::cuda_check(cudaGetSymbolAddress(&ptr, g_pairwise_static_scratch), "no-alloc pairwise: cudaGetSymbolAddress")) return 1;
        g_static_scratch_ptr[dev] = ptr;
    }
    *out = g_static_scratch_ptr[dev];
    return 0;
}
}  // namespace

// Same computation and outputs as gb_pairwise_select, but the caller passes no scratch buffer and
// nothing is allocated (no cudaMalloc / cudaFree, no allocating library call): the scratch is the
// library's own fixed device array.
//   * n is limited to 16 rows; larger n is refused with an error (use gb_pairwise_select).
//   * The first call on each device resolves the array's device address (with that device
//     current). Make that first call outside of any stream capture.
//   * All calls on a device share the one array and it keeps whatever the previous call left; the
//     words that must start at zero are cleared on `stream` before the kernel. Two calls that can
//     overlap on the GPU (different streams, no dependency between them) therefore race on the
//     scratch: serialise them.
extern "C" GB_API int gb_pairwise_select_noalloc(float const* const* rows, int n, int64_t d, int mode, int f, int m, int neighbors,
                                                 int squared, float* dist, float* W, int* info, void* stream) {
    using namespace gb;
    if (n < 1) return cuda_fail("no-alloc pairwise: need at least one row");
    if (n > kStaticMaxRows) return cuda_fail("no-alloc pairwise: at most 16 rows fit the built-in scratch array; use gb_pairwise_select with a caller-provided buffer for larger n");
    if (d < 1) return cuda_fail("no-alloc pairwise: empty rows");
    if (dist == nullptr) return cuda_fail("no-alloc pairwise: missing distance output buffer");
    if (mode != kSelectNone && (W == nullptr || info == nullptr)) return cuda_fail("no-alloc pairwise: selection needs W and info buffers");
    if (char const* bad = selection_error(mode, n, f, m, neighbors)) {
        static thread_local char msg[160];
        std::snprintf(msg, sizeof(msg), "no-alloc pairwise: %s", bad);
        return cuda_fail(msg);
    }
    void* scratch = nullptr;
    if (static_scratch_for_current_device(&scratch)) return 1;

    PairwiseArgs a{};
    for (int i = 0; i < n; ++i) a.rows.p[i] = rows[i];
    a.n = n; a.d = d; a.dist = dist; a.W = W; a.info = info;
    a.sel.mode = mode; a.sel.f = f; a.sel.m = m; a.sel.neighbors = neighbors; a.sel.squared = squared;
    a.tile = pick_tile(n);
    a.use_tma = rows_aligned(a.rows, n) ? 1 : 0;

    int const nn = n * n;
    int const slot_floats = ((kConsumerWarps * nn + 31) / 32) * 32;
    size_t smem = sizeof(float) * (static_cast<size_t>(slot_floats) + static_cast<size_t>(kStages) * n * a.tile);
    if (smem < sizeof(SelectScratch)) smem = sizeof(SelectScratch);
    long long const tiles = (d + a.tile - 1) / a.tile;
    int const ctas_per_sm = smem <= 100 * 1024 ? 2 : 1;
    long long grid = static_cast<long long>(device_sm_count()) * ctas_per_sm;
    long long const fit = (kStaticScratchBytes - 256) / (static_cast<long long>(nn) * static_cast<long long>(sizeof(float)));
    if (grid > fit) grid = fit;  // a GPU with more SMs than a B200 simply runs fewer CTAs
    if (grid > tiles) grid = tiles;
    if (grid < 1) grid = 1;
    a.counter = static_cast<unsigned int*>(scratch);
    a.partials = reinterpret_cast<float*>(static_cast<unsigned char*>(scratch) + 256);

    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    // the completion counter must be zero when the kernel starts, whatever an earlier call left behind
    if (cuda_check(cudaMemsetAsync(scratch, 0, 256, st), "no-alloc pairwise: clearing the counter")) return 1;
    auto kern = pairwise_kernel<4, false>;   // n <= 16 here: the fixed scratch array is sized for it
    if (cuda_check(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)), "no-alloc pairwise: smem attribute")) return 1;
    kern<<<static_cast<int>(grid), kThreads, smem, st>>>(a);
    return cuda_check(cudaGetLastError(), "no-alloc pairwise: launch");
}
