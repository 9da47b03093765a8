// coordwise_select kernel template (see gar_coordwise.cu for the description and citations).
// Instantiated per padded column size NP in gar_coordwise_np*.cu so the translation units build
// in parallel.
#pragma once

#include <cfloat>
#include <cmath>

#include "gar_common.cuh"

namespace gb {

template <int V>
struct Vec;
template <>
struct Vec<4> {
    using type = float4;
};
template <>
struct Vec<2> {
    using type = float2;
};
template <>
struct Vec<1> {
    using type = float;
};

template <int V>
__device__ __forceinline__ void vec_load(float (&dst)[V], const float* p) {
    typename Vec<V>::type const v = *reinterpret_cast<typename Vec<V>::type const*>(p);
    if constexpr (V == 4) {
        dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w;
    } else if constexpr (V == 2) {
        dst[0] = v.x; dst[1] = v.y;
    } else {
        dst[0] = v;
    }
}

// V floats from a 32-bit shared-memory address. The tile loads used to go through generic pointers (`src + i * tile + c`): the
// compiler then rebuilt the shared window base for EVERY row (S2UR + ULEA + UMOV + ... ≈ 15 instructions per LDS; FMNMX was 26 %
// of the median kernel's instructions at n = 23). With an explicit address a row costs one IMAD and one LDS.
template <int V>
__device__ __forceinline__ void lds_vec(float (&dst)[V], uint32_t addr) {
    if constexpr (V == 4) {
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(dst[0]), "=f"(dst[1]), "=f"(dst[2]), "=f"(dst[3]) : "r"(addr));
    } else if constexpr (V == 2) {
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(dst[0]), "=f"(dst[1]) : "r"(addr));
    } else {
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(dst[0]) : "r"(addr));
    }
}

template <int V>
__device__ __forceinline__ void vec_store(float* p, float const (&src)[V]) {
    if constexpr (V == 4) *reinterpret_cast<float4*>(p) = make_float4(src[0], src[1], src[2], src[3]);
    else if constexpr (V == 2) *reinterpret_cast<float2*>(p) = make_float2(src[0], src[1]);
    else *p = src[0];
}

// Parameter and momentum of V consecutive coordinates — what apply_update reads. Split off so that the streaming loop can issue
// these global loads one tile AHEAD: issued inside apply_update they start only after the tile's rows have been reduced, and every
// warp then sits out a full DRAM latency per tile (the rows come through the TMA ring; these two streams did not).
template <int V>
__device__ __forceinline__ void load_state(FusedUpdate const& u, long long x, float (&p)[V], float (&m)[V]) {
    if (u.mode == 0) return;
    vec_load<V>(p, u.param + x);
    if (u.momentum != 0.0f && !u.first_step) vec_load<V>(m, u.mom + x);
}

// SGD step (torch.optim.SGD semantics) + publication, on V consecutive coordinates; p / m = load_state of the same coordinates.
template <int V>
__device__ __forceinline__ void apply_update(FusedUpdate const& u, long long x, float (&g)[V], float* out, float (&p)[V], float const (&m)[V]) {
    if (out != nullptr) vec_store<V>(out + x, g);
    if (u.mode == 0) return;
#pragma unroll
    for (int e = 0; e < V; ++e) g[e] = fmaf(u.weight_decay, p[e], g[e]);
    if (u.momentum != 0.0f) {
        float buf[V];
        if (u.first_step) {
#pragma unroll
            for (int e = 0; e < V; ++e) buf[e] = g[e];
        } else {
#pragma unroll
            for (int e = 0; e < V; ++e) buf[e] = fmaf(u.momentum, m[e], (1.0f - u.dampening) * g[e]);
        }
        vec_store<V>(u.mom + x, buf);
#pragma unroll
        for (int e = 0; e < V; ++e) g[e] = u.nesterov ? fmaf(u.momentum, buf[e], g[e]) : buf[e];
    }
#pragma unroll
    for (int e = 0; e < V; ++e) p[e] = fmaf(-u.lr, g[e], p[e]);
    if (u.multicast != nullptr) {   // one multicast store reaches every GPU's parameter arena (this one included)
        if constexpr (V == 4) ptx::multimem_st(u.multicast + x, make_float4(p[0], p[1], p[2], p[3]));
        else if constexpr (V == 2) ptx::multimem_st(u.multicast + x, make_float2(p[0], p[1]));
        else ptx::multimem_st(u.multicast + x, p[0]);
        return;
    }
    vec_store<V>(u.param + x, p);
#pragma unroll
    for (int k = 0; k < kMaxMirrors; ++k)
        if (k < u.n_mirrors) vec_store<V>(u.mirror[k] + x, p);
}
template <int V>
__device__ __forceinline__ void apply_update(FusedUpdate const& u, long long x, float (&g)[V], float* out) {
    float p[V], m[V];
#pragma unroll
    for (int e = 0; e < V; ++e) p[e] = m[e] = 0.0f;
    load_state<V>(u, x, p, m);
    apply_update<V>(u, x, g, out, p, m);
}

// Reduce the `cnt` values of one coordinate held in v[0..NP). Entries at index >= cnt must hold
// FLT_MAX (finite padding that sorts last). A cheap all-finite test (Σ 0·v on the FMA pipe)
// selects the common path, which is the bare sorting network; columns containing NaN/±inf take
// the careful path (NaN propagation or non-finite filtering, depending on the rule).
template <int NP, int RULE, bool MERGER = true>
__device__ __forceinline__ float reduce_column(float (&v)[NP], int cnt, int rank, int upper, float p, unsigned long long seed,
                                               long long x, float first_row_value) {
    if constexpr (RULE == kRuleLinear) {
        return v[0];
    } else {
        float probe = 0.0f;
#pragma unroll
        for (int i = 0; i < NP; ++i) probe = fmaf(v[i], 0.0f, probe);
        bool has_nan = false;
        int finite = cnt;
        if (probe != 0.0f) {  // NaN: at least one non-finite value in this column
            finite = 0;
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                float const val = v[i];
                bool const live = i < cnt;
                bool const isnan_ = val != val;
                bool const fin = fabsf(val) <= FLT_MAX;
                has_nan |= live && isnan_;
                finite += (live && fin) ? 1 : 0;
                if constexpr (RULE == kRuleMedianFiniteUpper) v[i] = (live && fin) ? val : INFINITY;
                else v[i] = (live && !isnan_) ? val : INFINITY;
            }
        }
        if constexpr (RULE != kRuleMedianLower && RULE != kRuleCondense && RULE != kRuleMedianFiniteUpper) sort_regs<NP>(v);

        float result = 0.0f;
        if constexpr (RULE == kRuleMedianLower || RULE == kRuleCondense) {
            result = select_rank<NP>(v, (cnt - 1) >> 1);
            if (has_nan) result = NAN;
            if constexpr (RULE == kRuleCondense) {
                uint4 const r = philox4x32(seed, static_cast<unsigned long long>(x), 0u);
                result = (u01(r.x) < p) ? result : first_row_value;
            }
        } else if constexpr (RULE == kRuleMedianFiniteUpper) {
            result = select_rank<NP>(v, finite >> 1);
            if (finite == 0) result = 0.0f;
        } else if constexpr (RULE == kRuleTrimmedMean) {
            int const lo = rank, hi = cnt - rank;
            float sum = 0.0f;
#pragma unroll
            for (int i = 0; i < NP; ++i) sum += (i >= lo && i < hi) ? v[i] : 0.0f;
            result = has_nan ? NAN : sum / static_cast<float>(hi - lo);
        } else if constexpr (RULE == kRuleAveragedMedian) {
            float const med = pick_reg<NP>(v, upper ? (cnt >> 1) : ((cnt - 1) >> 1));
            // The β values closest to the median: v is sorted, so their distances to it form a bitonic sequence — a merger finds
            // the β-th smallest distance r and the sum takes what lies closer than r plus as many at distance r as are still
            // missing (this replaced a key-value sort of all NP pairs: 5 instructions per compare-exchange of the full network).
            if constexpr (!MERGER) {   // 8 mixed values (Bulyan at n = 8): measured 0.362 ms with the key-value sort, 0.376 with the merger
                float key[NP];
#pragma unroll
                for (int i = 0; i < NP; ++i) key[i] = (i < cnt) ? fabsf(v[i] - med) : INFINITY;
                sort_pairs<NP>(key, v);
                float sum = 0.0f;
#pragma unroll
                for (int i = 0; i < NP; ++i) sum += (i < rank) ? v[i] : 0.0f;
                result = has_nan ? NAN : sum / static_cast<float>(rank);
            } else {
                float key[NP], ordered[NP];
#pragma unroll
                for (int i = 0; i < NP; ++i) ordered[i] = key[i] = (i < cnt) ? fabsf(v[i] - med) : INFINITY;
                merge_bitonic<NP>(ordered);
                float const radius = pick_reg<NP>(ordered, rank - 1);
                float sum = 0.0f;
                int missing = rank;
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    bool const inside = key[i] < radius;
                    sum += inside ? v[i] : 0.0f;
                    missing -= inside ? 1 : 0;
                }
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    bool const edge = key[i] == radius && missing > 0;
                    sum += edge ? v[i] : 0.0f;
                    missing -= edge ? 1 : 0;
                }
                result = has_nan ? NAN : sum / static_cast<float>(rank);
            }
        }
        return result;
    }
}

// V consecutive coordinates starting at x: gather the columns through `loadv(i, dst[V])`, reduce.
template <int NP, int RULE, bool MIX, int V, class LoadV>
__device__ __forceinline__ void columns(CoordArgs const& a, float const* Ws, unsigned long long used_mask, long long x, LoadV loadv,
                                        float (&result)[V]) {
    float v[NP][V];
    float first[V];
    int cnt;
    if constexpr (MIX) {  // y_k = Σ_i W[k][i]·x_i; Ws is NP × n, zero rows beyond theta; unused rows skipped (may hold NaN)
        cnt = a.theta;
#pragma unroll
        for (int k = 0; k < NP; ++k)
#pragma unroll
            for (int e = 0; e < V; ++e) v[k][e] = 0.0f;
        uint32_t const ws_at = ptx::smem_u32(Ws);   // explicit shared addresses (see lds_vec)
        if constexpr (NP == 1) {   // weighted sum: the weights of 4 inputs per broadcast LDS.128; zero weight = input not read
            for (int i0 = 0; i0 < a.n; i0 += 4) {
                float w[4];
                lds_vec<4>(w, ws_at + static_cast<uint32_t>(i0) * 4u);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (w[j] != 0.0f) {  // warp-uniform; entries beyond n are zero
                        float xi[V];
                        loadv(i0 + j, xi);
#pragma unroll
                        for (int e = 0; e < V; ++e) v[0][e] = fmaf(w[j], xi[e], v[0][e]);
                    }
            }
        } else
        for (int i = 0; i < a.n; ++i) {
            if (!((used_mask >> i) & 1ull)) continue;  // warp-uniform
            float xi[V];
            loadv(i, xi);
            // Ws is stored row-of-W-per-input (n × NPW, NPW = NP rounded up to 4): the NP weights of input i are NPW/4 broadcast
            // LDS.128 instead of NP scalar loads (θ = 13, n = 23: 92 instead of 299 shared-memory wavefronts per coordinate)
            constexpr int NPW = NP == 1 ? 1 : (NP + 3) / 4 * 4;
            if constexpr (NP >= 4) {
#pragma unroll
                for (int k4 = 0; k4 < NP / 4; ++k4) {
                    float w[4];
                    lds_vec<4>(w, ws_at + static_cast<uint32_t>(i * NPW + k4 * 4) * 4u);
#pragma unroll
                    for (int e = 0; e < V; ++e) {
                        v[k4 * 4 + 0][e] = fmaf(w[0], xi[e], v[k4 * 4 + 0][e]);
                        v[k4 * 4 + 1][e] = fmaf(w[1], xi[e], v[k4 * 4 + 1][e]);
                        v[k4 * 4 + 2][e] = fmaf(w[2], xi[e], v[k4 * 4 + 2][e]);
                        v[k4 * 4 + 3][e] = fmaf(w[3], xi[e], v[k4 * 4 + 3][e]);
                    }
                }
            } else {
#pragma unroll
                for (int k = 0; k < NP; ++k) {
                    float const w = Ws[i * NPW + k];
#pragma unroll
                    for (int e = 0; e < V; ++e) v[k][e] = fmaf(w, xi[e], v[k][e]);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < NP; ++k)
            if (k >= cnt) {
#pragma unroll
                for (int e = 0; e < V; ++e) v[k][e] = FLT_MAX;
            }
#pragma unroll
        for (int e = 0; e < V; ++e) first[e] = 0.0f;
    } else {
        cnt = a.n;
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            if (i < a.n) {
                loadv(i, v[i]);
            } else {
#pragma unroll
                for (int e = 0; e < V; ++e) v[i][e] = FLT_MAX;
            }
        }
#pragma unroll
        for (int e = 0; e < V; ++e) first[e] = v[0][e];
    }
#pragma unroll
    for (int e = 0; e < V; ++e) {
        float col[NP];
#pragma unroll
        for (int i = 0; i < NP; ++i) col[i] = v[i][e];
        result[e] = reduce_column<NP, RULE, (NP >= 12 || !MIX)>(col, cnt, a.rank, a.upper, a.p, a.seed, x + e, first[e]);
    }
}

template <int NP, int RULE, bool MIX, int V>
__global__ void __launch_bounds__(kThreads) coordwise_kernel(const __grid_constant__ CoordArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ TileRing ring;
    constexpr int NPW = NP == 1 ? 1 : (NP + 3) / 4 * 4;
    int const n4 = (a.n + 3) / 4 * 4;
    float* Ws = reinterpret_cast<float*>(smem_raw);  // n4 × NPW weights (MIX only): Ws[i][k] = W[k][i], zero beyond theta and n
    int const w_floats = MIX ? ((NPW * n4 + 31) / 32) * 32 : 0;
    float* stages = Ws + w_floats;  // kStages × n × tile
    int const warp = threadIdx.x >> 5;
    int const tile = a.tile;
    long long const full_tiles = a.use_tma ? a.d / tile : 0;

    if constexpr (MIX)
        for (int e = threadIdx.x; e < NPW * n4; e += blockDim.x) {
            int const i = e / NPW, k = e % NPW;
            Ws[e] = (k < a.theta && i < a.n) ? a.W[k * a.n + i] : 0.0f;
        }
    ring_init(ring);  // contains the __syncthreads that also publishes Ws
    unsigned long long used_mask = ~0ull;  // rows with at least one non-zero weight
    if constexpr (MIX) {
        used_mask = 0ull;
        for (int i = 0; i < a.n; ++i) {
            bool used = false;
            for (int k = 0; k < a.theta; ++k) used |= (Ws[i * NPW + k] != 0.0f);
            used_mask |= used ? (1ull << i) : 0ull;
        }
    }

    if (warp == 0) {
        // (one copy per lane of the producer warp instead of one elected lane issuing all n was measured: no difference)
        if (full_tiles > 0 && ptx::elect_one())
            ring_produce(ring, stages, a.rows, a.n, tile, full_tiles, [&](int i) { return ((used_mask >> i) & 1ull) != 0ull; });
    } else {
        int const ctid = threadIdx.x - 32;
        constexpr int kConsumers = 32 * kConsumerWarps;
        size_t const stage_floats = static_cast<size_t>(a.n) * tile;
        long long it = 0;
        // every thread visits at most one column group per tile (the usual case: 1024-coordinate tiles, 4 coordinates per thread):
        // its parameter / momentum values are fetched one tile ahead
        bool const ahead = tile <= kConsumers * V && a.upd.mode != 0;
        bool const mine = ctid * V < tile;
        float p_next[V], m_next[V];
#pragma unroll
        for (int e = 0; e < V; ++e) p_next[e] = m_next[e] = 0.0f;
        if (ahead && mine && static_cast<long long>(blockIdx.x) < full_tiles)
            load_state<V>(a.upd, static_cast<long long>(blockIdx.x) * tile + ctid * V, p_next, m_next);
        for (long long t = blockIdx.x; t < full_tiles; t += gridDim.x, ++it) {
            int const s = static_cast<int>(it % kStages);
            uint32_t const phase = static_cast<uint32_t>((it / kStages) & 1);
            float p_now[V], m_now[V];
            if (ahead) {
#pragma unroll
                for (int e = 0; e < V; ++e) { p_now[e] = p_next[e]; m_now[e] = m_next[e]; }
                if (mine && t + gridDim.x < full_tiles) load_state<V>(a.upd, (t + gridDim.x) * tile + ctid * V, p_next, m_next);
            }
            ptx::mbar_wait(&ring.full[s], phase);
            uint32_t const src = ptx::smem_u32(stages + s * stage_floats);
            uint32_t const row_bytes = static_cast<uint32_t>(tile) * 4u;
            if (ahead) {
                if (mine) {
                    int const c = ctid * V;
                    long long const x = t * tile + c;
                    float r[V];
                    uint32_t const at = src + static_cast<uint32_t>(c) * 4u;
                    columns<NP, RULE, MIX, V>(a, Ws, used_mask, x, [&](int i, float (&dst)[V]) { lds_vec<V>(dst, at + static_cast<uint32_t>(i) * row_bytes); }, r);
                    apply_update<V>(a.upd, x, r, a.out, p_now, m_now);
                }
            } else
            for (int c = ctid * V; c < tile; c += kConsumers * V) {
                long long const x = t * tile + c;
                float r[V];
                uint32_t const at = src + static_cast<uint32_t>(c) * 4u;
                columns<NP, RULE, MIX, V>(a, Ws, used_mask, x, [&](int i, float (&dst)[V]) { lds_vec<V>(dst, at + static_cast<uint32_t>(i) * row_bytes); }, r);
                apply_update<V>(a.upd, x, r, a.out);
            }
            __syncwarp();
            if ((threadIdx.x & 31) == 0) ptx::mbar_arrive(&ring.empty[s]);
        }
        // remainder (and everything when buffers are not 16-byte aligned): direct global loads
        long long const done = full_tiles * tile;
        for (long long x = done + static_cast<long long>(blockIdx.x) * kConsumers + ctid; x < a.d;
             x += static_cast<long long>(gridDim.x) * kConsumers) {
            float r[1];
            columns<NP, RULE, MIX, 1>(a, Ws, used_mask, x, [&](int i, float (&dst)[1]) { dst[0] = a.rows.p[i][x]; }, r);
            apply_update<1>(a.upd, x, r, a.out);
        }
    }
}

template <int NP, int RULE, bool MIX, int V>
cudaError_t launch_coordwise(CoordArgs const& a, int grid, size_t smem, cudaStream_t stream) {
    auto kern = coordwise_kernel<NP, RULE, MIX, V>;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (err != cudaSuccess) return err;
    kern<<<grid, kThreads, smem, stream>>>(a);
    return cudaGetLastError();
}

// Per-NP dispatch over the rules (defined once per NP in gar_coordwise_np*.cu).
template <int NP>
cudaError_t dispatch_coordwise(CoordArgs const& a, int grid, size_t smem, cudaStream_t stream);

#define GB_DEFINE_COORDWISE_DISPATCH(NP, V, VMIX)                                                                          \
    template <>                                                                                                        \
    cudaError_t dispatch_coordwise<NP>(CoordArgs const& a, int grid, size_t smem, cudaStream_t stream) {               \
        bool const mix = a.theta > 0;                                                                                  \
        switch (a.rule) {                                                                                              \
            case kRuleMedianLower: return launch_coordwise<NP, kRuleMedianLower, false, V>(a, grid, smem, stream);     \
            case kRuleMedianFiniteUpper: return launch_coordwise<NP, kRuleMedianFiniteUpper, false, V>(a, grid, smem, stream); \
            case kRuleTrimmedMean: return launch_coordwise<NP, kRuleTrimmedMean, false, V>(a, grid, smem, stream);     \
            case kRuleCondense: return launch_coordwise<NP, kRuleCondense, false, V>(a, grid, smem, stream);           \
            case kRuleAveragedMedian:                                                                                  \
                return mix ? launch_coordwise<NP, kRuleAveragedMedian, true, VMIX>(a, grid, smem, stream)              \
                           : launch_coordwise<NP, kRuleAveragedMedian, false, V>(a, grid, smem, stream);               \
            default: return cudaErrorInvalidValue;                                                                     \
        }                                                                                                              \
    }

}  // namespace gb
