// conv_tc: implicit-GEMM convolutions (forward, data gradient, per-worker weight gradient) and, as their 1×1 special
// case, Linear layers, on the 5th-generation tensor cores.
//
// No im2col matrix is materialised: every 128-byte K-slab of an operand tile is ONE tiled TMA load of a 5-D box
// (32 channels × box_w × box_h × box_n × 1 group) out of the channels-last activation tensor, shifted by the filter
// tap's offset — out-of-bounds pixels are the convolution's zero padding (TMA zero fill), a convolution stride is the
// tensor map's traversal stride. The box lands in shared memory as [pixels][32 floats] rows with the 128-byte
// swizzle, which is both
//   * the K-major canonical tcgen05 operand (pixels = M rows, channels = K)      → forward and data-gradient GEMMs, and
//   * the MN-major canonical operand (pixels = K rows, channels = N columns)     → weight-gradient GEMMs,
// so the same loads feed all three. Weights are read in their stored OHWI layout [Cout][kh·kw·Cin]: K-major for the
// forward pass, MN-major (rows = Cout = K, columns = Cin = N) for the data gradient — no transposed copy exists.
//
//   fprop   Y[pix, Cout]      = Σ_tap  X[pix ⊕ tap, Cin]  · W[Cout, tap, Cin]ᵀ          (+ bias, ReLU)
//   dgrad   dX[pix, Cin]      = Σ_tap  dY[pix ⊖ tap, Cout] · W[Cout, tap, Cin]           (stride 1)
//   wgrad   dW_g[Cout, tap, Cin] = Σ_{pix of worker g}  dY[pix, Cout]ᵀ · X[pix ⊕ tap, Cin]   (G workers = grid.z)
//
// Accumulators live in TMEM (tcgen05.mma kind::tf32, fp32 accumulate — the precision of the library convolutions they
// replace), warp 0 = TMA producer, warp 1 = MMA issuer, warps 2-5 = epilogue (tcgen05.ld → shared-memory transpose →
// 128-byte row stores). Non-persistent small tiles (128 × 64 / 128 × 128): the flagship's GEMMs have M ≤ 12 800 rows.
//
// Reference counterpart: none in the reference's own code — it reaches cuDNN / cuBLAS through nn.Conv2d / nn.Linear
// (/root/reference/pytorch_impl/libs/garfieldpp/models/nets.py:40-78, tools.py:69-75, worker.py:85-91).
// All tensor maps cover this GPU's memory only.

#include <cuda.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "garfield_b200.h"
#include "launch.cuh"
#include "ptx.cuh"

namespace gb {
namespace {

constexpr int kConvThreads = 32 * 6;
constexpr int kMaxTaps = 32;
constexpr int kStagePitch = 36;       // floats per staged epilogue row
constexpr uint32_t kATile = 128 * 128;   // bytes reserved for the M-side operand of one stage

struct ConvGeom {
    int taps, kchunks;        // filter taps; 32-float chunks of the reduction extent per tap (fprop/dgrad) or pixel chunks (wgrad)
    int k_extent;             // reduction extent per tap: Cin (fprop), Cout (dgrad)
    int n_total;              // N extent: Cout (fprop), Cin (dgrad), Cin (wgrad, per tap)
    int m_total;              // M extent: output pixels (fprop / dgrad), Cout (wgrad)
    int bw, bh, bn;           // pixel box of one tile / chunk
    int sh, sw;               // traversal strides of the pixel-box operand
    int wo, ho, nimg;         // pixel grid tiled by the boxes
    int hin, win;             // spatial extent of the pixel-box operand (tap validity)
    int tiles_h;              // tiles per image along h (1 when a tile covers whole images)
    int n_tiles_per_tap;      // wgrad: N tiles per filter tap
    int n_store;              // wgrad: columns stored per tap (<= n_total; a zero-padded reduction matrix has fewer real ones)
    int relu;
    int stages;               // depth of the TMA -> MMA shared-memory ring (2..8)
    int split;                // split-K: CTAs of a cluster (along grid.z) that share one output tile (1 = none; fprop / dgrad only)
    int mn_lbo, mn_sbo, mn_layout, mn_kstep;   // MN-major operand descriptors (bytes; see mn_descriptor_defaults)
    long long ldc, c_group_stride;
    const float* addend;      // optional tensor of the output's shape and pitch added in the epilogue (gradient of a skip connection)
    signed char dh[kMaxTaps], dw[kMaxTaps];
};

constexpr int kMaxStages = 8;

struct ConvShared {
    uint64_t full[kMaxStages];
    uint64_t empty[kMaxStages];
    uint64_t acc_full;
    uint32_t tmem_base;
    uint32_t tap_mask;
};

__device__ __forceinline__ void tma_load_3d(void* dst, const void* map, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                     ptx::smem_u32(dst)),
                 "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(ptx::smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_load_5d(void* dst, const void* map, int c0, int c1, int c2, int c3, int c4, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];" ::"r"(
            ptx::smem_u32(dst)),
        "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "r"(ptx::smem_u32(bar))
        : "memory");
}

// ---- thread-block cluster helpers (split-K reduction through distributed shared memory) ----
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float4 ld_dsmem_f4(const float* local_ptr, uint32_t cta_rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(ptx::smem_u32(local_ptr)), "r"(cta_rank));
    float4 v;
    asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(remote) : "memory");
    return v;
}

// Does filter tap t touch any in-bounds pixel for rows h0 … h0+rows-1 of the pixel grid? (interval test: conservative)
__device__ __forceinline__ bool tap_in_range(ConvGeom const& g, int t, int h0, int rows) {
    int const lo_h = h0 * g.sh + g.dh[t], hi_h = (h0 + rows - 1) * g.sh + g.dh[t];
    int const lo_w = g.dw[t], hi_w = (g.wo - 1) * g.sw + g.dw[t];
    return hi_h >= 0 && lo_h < g.hin && hi_w >= 0 && lo_w < g.win;
}

// One warp moves its TMEM lane quadrant (32 rows × BN columns) to global memory: tcgen05.ld → bias / ReLU → transpose
// through shared memory → 128-byte row segments.
template <int BN>
__device__ __forceinline__ void store_quadrant(uint32_t acc, float* stage, float* __restrict__ C, long long ldc, int row0, int rows_end, int col0,
                                               int cols_end, const float* __restrict__ bias, int relu, bool zero, int lane,
                                               const float* __restrict__ addend = nullptr) {
    bool const vec_ok = (reinterpret_cast<uintptr_t>(C) & 15u) == 0 && (ldc & 3) == 0 && (col0 & 3) == 0;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t v[32];
        ptx::tmem_ld_32x32b_x16(acc + static_cast<uint32_t>(c0), reinterpret_cast<uint32_t(&)[16]>(v[0]));
        ptx::tmem_ld_32x32b_x16(acc + static_cast<uint32_t>(c0 + 16), reinterpret_cast<uint32_t(&)[16]>(v[16]));
        ptx::tmem_ld_wait();
        int const cbase = col0 + c0;
        if (cbase >= cols_end || row0 >= rows_end) continue;                // warp-uniform
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            float4 o = make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]), __uint_as_float(v[4 * q + 2]),
                                   __uint_as_float(v[4 * q + 3]));
            if (zero) o = make_float4(0.f, 0.f, 0.f, 0.f);
            if (bias != nullptr) {
                int const c = cbase + 4 * q;
                o.x += c < cols_end ? bias[c] : 0.f; o.y += c + 1 < cols_end ? bias[c + 1] : 0.f;
                o.z += c + 2 < cols_end ? bias[c + 2] : 0.f; o.w += c + 3 < cols_end ? bias[c + 3] : 0.f;
            }
            if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
            *reinterpret_cast<float4*>(stage + lane * kStagePitch + 4 * q) = o;
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int const r = i * 4 + (lane >> 3), c = (lane & 7) * 4;
            float4 const o = *reinterpret_cast<float4 const*>(stage + r * kStagePitch + c);
            int const grow = row0 + r, gcol = cbase + c;
            if (grow < rows_end && gcol < cols_end) {
                float* out = C + static_cast<long long>(grow) * ldc + gcol;
                if (vec_ok && gcol + 3 < cols_end) {
                    float4 v = o;
                    if (addend != nullptr) {
                        float4 const t = *reinterpret_cast<float4 const*>(addend + static_cast<long long>(grow) * ldc + gcol);
                        v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w;
                    }
                    *reinterpret_cast<float4*>(out) = v;
                } else {
                    const float* t = addend != nullptr ? addend + static_cast<long long>(grow) * ldc + gcol : nullptr;
                    out[0] = o.x + (t ? t[0] : 0.f);
                    if (gcol + 1 < cols_end) out[1] = o.y + (t ? t[1] : 0.f);
                    if (gcol + 2 < cols_end) out[2] = o.z + (t ? t[2] : 0.f);
                    if (gcol + 3 < cols_end) out[3] = o.w + (t ? t[3] : 0.f);
                }
            }
        }
        __syncwarp();
    }
}

// MODE 0: fprop (B = weights, K-major)   MODE 1: dgrad (B = weights, MN-major)   MODE 2: wgrad (A = dY, B = X, both MN-major)
// SPLIT: the K loop of one output tile is shared by the g.split CTAs of a cluster (grid.z); their TMEM partials meet in shared
// memory and every CTA reduces one row slab over DSMEM in CTA order (deterministic) before storing it.
template <int BN, int MODE, bool SPLIT = false>
__global__ void __launch_bounds__(kConvThreads)
conv_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const __grid_constant__ ConvGeom g,
               float* __restrict__ C, const float* __restrict__ bias) {
    extern __shared__ unsigned char smem_dyn[];
    __shared__ ConvShared sh;
    constexpr uint32_t kBBytes = BN * 128, kStageBytes = kATile + kBBytes;
    unsigned char* const smem_raw = smem_dyn + ((1024u - (ptx::smem_u32(smem_dyn) & 1023u)) & 1023u);   // 1024-byte aligned (swizzle atoms)
    int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int const STAGES = g.stages;

    // ---- tile coordinates ----
    int n0 = 0, h0 = 0, rows_valid, m0, tile_n, tap_fixed = 0, grp = 0;
    if (MODE < 2) {
        int const tile_m = blockIdx.x;
        tile_n = blockIdx.y;
        if (g.tiles_h == 1) {
            n0 = tile_m * g.bn;
            int const imgs = min(g.bn, g.nimg - n0);
            rows_valid = imgs * g.ho * g.wo;
            m0 = n0 * g.ho * g.wo;
        } else {
            n0 = tile_m / g.tiles_h;
            h0 = (tile_m % g.tiles_h) * g.bh;
            rows_valid = min(g.bh, g.ho - h0) * g.wo;
            m0 = (n0 * g.ho + h0) * g.wo;
        }
    } else {
        tap_fixed = blockIdx.x / g.n_tiles_per_tap;
        tile_n = blockIdx.x % g.n_tiles_per_tap;
        m0 = blockIdx.y * 128;
        rows_valid = min(128, g.m_total - m0);
        grp = blockIdx.z;
    }

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            ptx::mbar_init(&sh.full[s], 1);
            ptx::mbar_init(&sh.empty[s], 1);
        }
        ptx::mbar_init(&sh.acc_full, 1);
        ptx::fence_mbar_init();
        ptx::prefetch_tensormap(&map_a);
        ptx::prefetch_tensormap(&map_b);
        uint32_t mask = 0;
        if (MODE < 2) {
            int const rows_h = g.tiles_h == 1 ? g.ho : min(g.bh, g.ho - h0);
            for (int t = 0; t < g.taps; ++t)
                if (tap_in_range(g, t, h0, rows_h)) mask |= 1u << t;
        } else {
            mask = tap_in_range(g, tap_fixed, 0, g.ho) ? 1u : 0u;
        }
        sh.tap_mask = mask;
    }
    if (warp == 1) ptx::tmem_alloc<(BN < 32 ? 32 : BN)>(&sh.tmem_base);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    uint32_t const tmem = sh.tmem_base;
    uint32_t const tap_mask = sh.tap_mask;
    // Programmatic dependent launch: everything above (barriers, TMEM, descriptor prefetch) overlapped the tail of the kernel
    // before us; from here on we touch its results. Let the kernel after us start ITS prologue right away.
    cudaGridDependencySynchronize();
    cudaTriggerProgrammaticLaunchCompletion();
    uint32_t const a_bytes = MODE < 2 ? static_cast<uint32_t>(g.bn * g.bh * g.bw) * 128u : kATile;
    // this CTA's share [step_lo, step_hi) of the flattened (valid tap, K chunk) sequence
    int const steps_total = MODE < 2 ? __popc(tap_mask) * g.kchunks : (tap_mask ? g.kchunks : 0);
    int const zr = SPLIT ? static_cast<int>(blockIdx.z) : 0;
    int const step_lo = SPLIT ? static_cast<int>(static_cast<long long>(steps_total) * zr / g.split) : 0;
    int const step_hi = SPLIT ? static_cast<int>(static_cast<long long>(steps_total) * (zr + 1) / g.split) : steps_total;

    if (warp == 0) {
        if (ptx::elect_one()) {
            int s = 0;
            uint32_t phase = 0;
            auto advance = [&] { if (++s == STAGES) { s = 0; phase ^= 1u; } };
            if (MODE < 2) {
                int step = 0;
                for (int t = 0; t < g.taps; ++t) {
                    if (!((tap_mask >> t) & 1u)) continue;
                    if (SPLIT && (step + g.kchunks <= step_lo || step >= step_hi)) { step += g.kchunks; continue; }
                    for (int kc = 0; kc < g.kchunks; ++kc, ++step) {
                        if (SPLIT && (step < step_lo || step >= step_hi)) continue;
                        ptx::mbar_wait(&sh.empty[s], phase ^ 1u);
                        ptx::mbar_arrive_expect_tx(&sh.full[s], a_bytes + kBBytes);
                        unsigned char* dst = smem_raw + static_cast<size_t>(s) * kStageBytes;
                        tma_load_5d(dst, &map_a, kc * 32, g.dw[t], h0 * g.sh + g.dh[t], n0, 0, &sh.full[s]);
                        if (MODE == 0) {
                            ptx::tma_load_2d(dst + kATile, &map_b, t * g.k_extent + kc * 32, tile_n * BN, &sh.full[s]);
                        } else {
#pragma unroll
                            for (int j = 0; j < BN / 32; ++j)
                                ptx::tma_load_2d(dst + kATile + j * 4096, &map_b, t * g.n_total + tile_n * BN + 32 * j, kc * 32, &sh.full[s]);
                        }
                        advance();
                    }
                }
            } else if (tap_mask) {
                int const hw = g.ho * g.wo;
                for (int pc = 0; pc < g.kchunks; ++pc, advance()) {
                    int const p0 = pc * 32;
                    int const b0 = hw >= 32 ? p0 / hw : pc * g.bn;
                    int const hh = hw >= 32 ? (p0 % hw) / g.wo : 0;
                    ptx::mbar_wait(&sh.empty[s], phase ^ 1u);
                    ptx::mbar_arrive_expect_tx(&sh.full[s], kATile + kBBytes);
                    unsigned char* dst = smem_raw + static_cast<size_t>(s) * kStageBytes;
#pragma unroll
                    for (int i = 0; i < 4; ++i) tma_load_3d(dst + i * 4096, &map_a, m0 + 32 * i, p0, grp, &sh.full[s]);
#pragma unroll
                    for (int j = 0; j < BN / 32; ++j)
                        tma_load_5d(dst + kATile + j * 4096, &map_b, tile_n * BN + 32 * j, g.dw[tap_fixed], hh * g.sh + g.dh[tap_fixed], b0, grp,
                                    &sh.full[s]);
                }
            }
        }
    } else if (warp == 1) {
        if (ptx::elect_one()) {
            uint32_t const idesc = ptx::make_idesc(2u, 2u, 128, BN, MODE == 2 ? 1u : 0u, MODE >= 1 ? 1u : 0u);
            int const steps = step_hi - step_lo;
            int s = 0;
            uint32_t phase = 0;
            for (int it = 0; it < steps; ++it) {
                ptx::mbar_wait(&sh.full[s], phase);
                ptx::tc_fence_after();
                uint32_t const a_base = ptx::smem_u32(smem_raw + static_cast<size_t>(s) * kStageBytes);
                uint32_t const b_base = a_base + kATile;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    uint32_t const mn_step = static_cast<uint32_t>(k * g.mn_kstep);
                    uint64_t const adesc = MODE == 2 ? ptx::make_smem_desc(a_base + mn_step, g.mn_lbo, g.mn_sbo, g.mn_layout)
                                                     : ptx::make_smem_desc(a_base + k * 32u, 16u, 1024u, 2u);
                    uint64_t const bdesc = MODE >= 1 ? ptx::make_smem_desc(b_base + mn_step, g.mn_lbo, g.mn_sbo, g.mn_layout)
                                                     : ptx::make_smem_desc(b_base + k * 32u, 16u, 1024u, 2u);
                    ptx::mma_tf32_ss(tmem, adesc, bdesc, idesc, (it > 0 || k > 0) ? 1u : 0u);
                }
                ptx::mma_commit(&sh.empty[s]);
                if (++s == STAGES) { s = 0; phase ^= 1u; }
            }
            ptx::mma_commit(&sh.acc_full);
        }
    } else {
        int const quad = warp & 3;
        bool const nothing = step_hi <= step_lo;         // no MMA was issued: the accumulator is undefined, the tile is zero
        ptx::mbar_wait(&sh.acc_full, 0u);
        ptx::tc_fence_after();
        uint32_t const acc = tmem + (static_cast<uint32_t>(quad * 32) << 16);
        if constexpr (SPLIT) {
            // partial tile -> shared memory [128][BN + 4]; the reduction follows the cluster barrier below
            float* const red = reinterpret_cast<float*>(smem_raw);
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                uint32_t v[32];
                ptx::tmem_ld_32x32b_x16(acc + static_cast<uint32_t>(c0), reinterpret_cast<uint32_t(&)[16]>(v[0]));
                ptx::tmem_ld_32x32b_x16(acc + static_cast<uint32_t>(c0 + 16), reinterpret_cast<uint32_t(&)[16]>(v[16]));
                ptx::tmem_ld_wait();
                float* row = red + (quad * 32 + lane) * (BN + 4) + c0;
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    *reinterpret_cast<float4*>(row + 4 * q) = nothing ? make_float4(0.f, 0.f, 0.f, 0.f)
                                                                      : make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]),
                                                                                    __uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3]));
            }
        } else {
        float* const stage = reinterpret_cast<float*>(smem_raw) + quad * 32 * kStagePitch;   // the pipeline's smem is idle by now
        float* out = C;
        int col0 = tile_n * BN, cols_end = g.n_total;
        if (MODE == 2) {
            out = C + static_cast<long long>(grp) * g.c_group_stride + static_cast<long long>(tap_fixed) * g.n_store;
            cols_end = g.n_store;
        }
        store_quadrant<BN>(acc, stage, out, g.ldc, m0 + quad * 32, m0 + rows_valid, col0, cols_end, MODE == 0 ? bias : nullptr, g.relu, nothing,
                           lane, MODE == 1 ? g.addend : nullptr);
        }
    }
    if constexpr (SPLIT) {
        __syncthreads();
        cluster_sync_all();                               // every CTA's partial tile is in its shared memory
        if (warp >= 2) {
            float const* const red = reinterpret_cast<float const*>(smem_raw);
            int const rows_per = 128 / g.split;
            int const et = threadIdx.x - 64;                  // 0..127
            int const col0 = tile_n * BN, cols_end = g.n_total;
            bool const vec_ok = (reinterpret_cast<uintptr_t>(C) & 15u) == 0 && (g.ldc & 3) == 0;
            for (int e = et; e < rows_per * (BN / 4); e += 128) {
                int const r = zr * rows_per + e / (BN / 4), c = (e % (BN / 4)) * 4;
                float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int src = 0; src < g.split; ++src) {     // fixed order: bitwise reproducible
                    float4 const v = ld_dsmem_f4(red + r * (BN + 4) + c, static_cast<uint32_t>(src));
                    sum.x += v.x; sum.y += v.y; sum.z += v.z; sum.w += v.w;
                }
                int const grow = m0 + r, gcol = col0 + c;
                if (r >= rows_valid || gcol >= cols_end) continue;
                float o[4] = {sum.x, sum.y, sum.z, sum.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (gcol + k >= cols_end) break;
                    if (MODE == 0 && bias != nullptr) o[k] += bias[gcol + k];
                    if (MODE == 1 && g.addend != nullptr) o[k] += g.addend[static_cast<long long>(grow) * g.ldc + gcol + k];
                    if (g.relu) o[k] = fmaxf(o[k], 0.f);
                }
                float* out = C + static_cast<long long>(grow) * g.ldc + gcol;
                if (vec_ok && gcol + 3 < cols_end) {
                    *reinterpret_cast<float4*>(out) = make_float4(o[0], o[1], o[2], o[3]);
                } else {
                    for (int k = 0; k < 4 && gcol + k < cols_end; ++k) out[k] = o[k];
                }
            }
        }
        cluster_sync_all();                               // nobody leaves while a peer may still read its tile
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<(BN < 32 ? 32 : BN)>(tmem);
}

using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                              const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn conv_encode_fn() {
    static EncodeFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult status;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &status) == cudaSuccess &&
            status == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeFn>(ptr);
    });
    return fn;
}

// MN-major tf32 operands (the reduction index runs over shared-memory ROWS, 32 consecutive floats of a row are 32 M/N
// indices) exist in one canonical form only: 128-byte rows swizzled in 32-byte units with a 4-row period
// (TMA: SWIZZLE_128B_ATOM_32B; descriptor layout type 1), atoms of 4 rows 512 bytes apart (SBO), 32-float column blocks
// one TMA box = 4096 bytes apart (LBO); one MMA (K = 8) consumes 8 rows = 1024 bytes.
// GARFIELD_MN_DESC="lbo,sbo,layout,kstep,tma_swizzle" overrides them (bring-up experiments only).
struct MnDesc {
    int lbo = 4096, sbo = 512, layout = 1, kstep = 1024, tma = static_cast<int>(CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
};
MnDesc const& mn_desc() {
    static MnDesc d = [] {
        MnDesc v;
        if (char const* env = std::getenv("GARFIELD_MN_DESC")) std::sscanf(env, "%d,%d,%d,%d,%d", &v.lbo, &v.sbo, &v.layout, &v.kstep, &v.tma);
        return v;
    }();
    return d;
}
CUtensorMapSwizzle mn_tma_swizzle() { return static_cast<CUtensorMapSwizzle>(mn_desc().tma); }
void mn_descriptor_defaults(ConvGeom& g) {
    MnDesc const& d = mn_desc();
    g.mn_lbo = d.lbo; g.mn_sbo = d.sbo; g.mn_layout = d.layout; g.mn_kstep = d.kstep;
}

thread_local char g_map_error[512] = "";

// fp32 tensor of `rank` dims (dim 0 contiguous), strides in floats for dims 1…, 128-byte swizzle, zero fill outside.
bool make_map_nd(CUtensorMap* map, const void* base, int rank, const long long* dims, const long long* strides, const int* box, const int* estr,
                 bool mn_major = false) {
    EncodeFn fn = conv_encode_fn();
    if (fn == nullptr) return false;
    cuuint64_t gd[5], gs[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) {
        gd[i] = static_cast<cuuint64_t>(dims[i] < 1 ? 1 : dims[i]);
        bx[i] = static_cast<cuuint32_t>(box[i]);
        es[i] = static_cast<cuuint32_t>(estr[i]);
        if (i > 0) gs[i - 1] = static_cast<cuuint64_t>(strides[i]) * 4u;
    }
    auto encode = [&] {
        return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gd, gs, bx, es,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? mn_tma_swizzle() : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    };
    CUresult rc = encode();
    if (rc == CUDA_ERROR_INVALID_CONTEXT) {
        // a thread that has made no runtime call yet (autograd's backward threads) has no current context: bind the primary one
        cudaFree(nullptr);
        rc = encode();
    }
    if (rc != CUDA_SUCCESS) {
        int n = std::snprintf(g_map_error, sizeof(g_map_error), "cuTensorMapEncodeTiled rc=%d base=%p rank=%d", static_cast<int>(rc), base, rank);
        for (int i = 0; i < rank && n < static_cast<int>(sizeof(g_map_error)) - 64; ++i)
            n += std::snprintf(g_map_error + n, sizeof(g_map_error) - n, " [dim %lld stride %lld box %d step %d]", dims[i], strides[i], box[i], estr[i]);
    }
    return rc == CUDA_SUCCESS;
}

// Pixel-box map over a channels-last tensor [G][N][H][W][pitch] (C valid channels): box = 32 channels × bw·sw × bh·sh × bn × 1.
bool make_pixel_map(CUtensorMap* map, const float* base, int C, int pitch, int W, int H, int N, int G, int bw, int bh, int bn, int sw, int sh,
                    bool mn_major = false) {
    long long dims[5] = {C, W, H, N, G};
    long long strides[5] = {1, pitch, static_cast<long long>(W) * pitch, static_cast<long long>(H) * W * pitch,
                            static_cast<long long>(N) * H * W * pitch};
    int box[5] = {32, bw * sw, bh * sh, bn, 1};
    int estr[5] = {1, sw, sh, 1, 1};
    return make_map_nd(map, base, 5, dims, strides, box, estr, mn_major);
}

// GARFIELD_PDL=0 switches programmatic dependent launches off (plain stream order).
bool pdl_enabled() {
    static int const on = [] { char const* v = std::getenv("GARFIELD_PDL"); return (v != nullptr && v[0] == '0') ? 0 : 1; }();
    return on != 0;
}

// GARFIELD_SPLIT_K=0 switches the cluster split-K of small-grid, deep-K problems off.
bool split_k_enabled() {
    static int const on = [] { char const* v = std::getenv("GARFIELD_SPLIT_K"); return (v != nullptr && v[0] == '0') ? 0 : 1; }();
    return on != 0;
}

int env_int(char const* name, int fallback) {
    char const* v = std::getenv(name);
    return v != nullptr && v[0] != 0 ? std::atoi(v) : fallback;
}
// Largest cluster (2, 4 or 8 CTAs per output tile) and the fewest K steps a CTA must keep for the split to pay for its cluster
// barriers and the DSMEM reduction (tuned on B200: profiles/README.md).
int split_k_max() { static int const v = env_int("GARFIELD_SPLIT_MAX", 4); return v; }
int split_k_min_steps() { static int const v = env_int("GARFIELD_SPLIT_MIN_STEPS", 12); return v; }

template <int BN, int MODE>
cudaError_t launch_conv(CUtensorMap const& ma, CUtensorMap const& mb, ConvGeom& g, float* C, const float* bias, dim3 grid, cudaStream_t st) {
    constexpr size_t kStage = kATile + BN * 128;
    constexpr int kDeep = static_cast<int>((200u * 1024u) / kStage) > kMaxStages ? kMaxStages : static_cast<int>((200u * 1024u) / kStage);
    constexpr bool kCanSplit = MODE < 2 && BN == 64;
    auto kern = conv_tc_kernel<BN, MODE, false>;
    auto kern_split = conv_tc_kernel<BN, MODE, kCanSplit>;
    static bool configured = false;      // per instantiation
    if (!configured) {
        cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kDeep * kStage + 1024));
        if (err == cudaSuccess && kCanSplit)
            err = cudaFuncSetAttribute(kern_split, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kDeep * kStage + 1024));
        if (err != cudaSuccess) return err;
        configured = true;
    }
    long long const ctas = static_cast<long long>(grid.x) * grid.y * grid.z;
    int steps = MODE < 2 ? g.taps * g.kchunks : g.kchunks;
    // Split-K: a small grid with a long reduction loop leaves most SMs idle while a few CTAs walk K serially. Up to 8 CTAs (one
    // cluster) share an output tile; their partials are reduced through distributed shared memory in CTA order.
    int split = 1;
    if (kCanSplit && split_k_enabled())
        for (int s2 = split_k_max(); s2 >= 2; s2 >>= 1)
            if (ctas * s2 <= device_sm_count() && steps / s2 >= split_k_min_steps()) { split = s2; break; }
    g.split = split;
    if (split > 1) {
        grid.z = static_cast<unsigned>(split);
        steps = (steps + split - 1) / split;
    }
    // Latency-bound shapes (at most one CTA per SM, long reduction loops) get the deepest ring that fits: the bytes in flight per
    // SM set the rate. Larger grids keep two CTAs per SM so that one CTA's prologue / epilogue overlaps the other's main loop.
    int stages = ctas * split <= device_sm_count() ? kDeep : 4;
    if (BN == 128 && stages == 4) stages = 3;
    if (stages > steps) stages = steps < 2 ? 2 : steps;
    g.stages = stages;
    size_t const smem = static_cast<size_t>(stages) * kStage + 1024;      // >= the split-K reduction tile (128 x 68 floats) for stages >= 2
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kConvThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int count = 0;
    if (pdl_enabled()) {
        attr[count].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[count].val.programmaticStreamSerializationAllowed = 1;
        ++count;
    }
    if (split > 1) {
        attr[count].id = cudaLaunchAttributeClusterDimension;
        attr[count].val.clusterDim.x = 1;
        attr[count].val.clusterDim.y = 1;
        attr[count].val.clusterDim.z = static_cast<unsigned>(split);
        ++count;
    }
    cfg.attrs = attr;
    cfg.numAttrs = count;
    if (split > 1) return cudaLaunchKernelEx(&cfg, kern_split, ma, mb, g, C, bias);
    return cudaLaunchKernelEx(&cfg, kern, ma, mb, g, C, bias);
}

int map_fail(char const* what) {
    static thread_local char msg[640];
    std::snprintf(msg, sizeof(msg), "%s: %s", what, g_map_error);
    return cuda_fail(msg);
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int fill_taps(ConvGeom& g, int kh, int kw, int pad_h, int pad_w, bool flip) {
    if (kh * kw > kMaxTaps) return 1;
    g.taps = kh * kw;
    for (int r = 0; r < kh; ++r)
        for (int s = 0; s < kw; ++s) {
            g.dh[r * kw + s] = static_cast<signed char>(flip ? pad_h - r : r - pad_h);
            g.dw[r * kw + s] = static_cast<signed char>(flip ? pad_w - s : s - pad_w);
        }
    return 0;
}

// Tile the [nimg][ho][wo] pixel grid with ≤128-pixel boxes: whole rows, then whole images.
int pick_pixel_tile(ConvGeom& g, int nimg, int ho, int wo) {
    if (wo > 128 || wo < 1 || ho < 1 || nimg < 1) return 1;
    g.wo = wo; g.ho = ho; g.nimg = nimg;
    g.bw = wo;
    g.bh = 128 / wo < ho ? 128 / wo : ho;
    if (g.bh == ho) {
        int const per = 128 / (wo * ho);
        g.bn = per < nimg ? per : nimg;
        g.tiles_h = 1;
    } else {
        g.bn = 1;
        g.tiles_h = (ho + g.bh - 1) / g.bh;
    }
    return 0;
}

}  // namespace
}  // namespace gb

using namespace gb;

// y[n, ho, wo, :cout] (pixel pitch y_pitch) = conv(x[n, h, w, :cin] (pixel pitch x_pitch), w[cout, kh, kw, cin]) (+ bias) (ReLU)
extern "C" GB_API int gb_conv_fprop_tc(float const* x, float const* w, float const* bias, float* y, int n, int h, int wd, int cin, int x_pitch,
                                       int cout, int y_pitch, int kh, int kw, int stride_h, int stride_w, int pad_h, int pad_w, int relu,
                                       void* stream) {
    if (n < 1 || h < 1 || wd < 1 || cin < 1 || cout < 1 || kh < 1 || kw < 1 || stride_h < 1 || stride_w < 1 || pad_h < 0 || pad_w < 0)
        return cuda_fail("conv_fprop_tc: bad geometry");
    if ((kh * kw * cin) % 4 || x_pitch % 4 || x_pitch < cin || y_pitch < cout || !aligned16(x) || !aligned16(w))
        return cuda_fail("conv_fprop_tc: kh*kw*cin and the input pixel pitch must be multiples of 4 floats, pointers 16-byte aligned");
    if (stride_h > 8 || stride_w > 8) return cuda_fail("conv_fprop_tc: stride > 8");
    int const ho = (h + 2 * pad_h - kh) / stride_h + 1, wo = (wd + 2 * pad_w - kw) / stride_w + 1;
    ConvGeom g{};
    if (fill_taps(g, kh, kw, pad_h, pad_w, false)) return cuda_fail("conv_fprop_tc: more than 32 filter taps");
    if (pick_pixel_tile(g, n, ho, wo)) return cuda_fail("conv_fprop_tc: output width must be 1..128");
    if (g.bw * stride_w > 256 || g.bh * stride_h > 256) return cuda_fail("conv_fprop_tc: strided box exceeds 256");
    g.k_extent = cin; g.kchunks = (cin + 31) / 32; g.n_total = cout; g.m_total = n * ho * wo;
    g.sh = stride_h; g.sw = stride_w; g.hin = h; g.win = wd; g.relu = relu; g.ldc = y_pitch;
    int const tiles_m = g.tiles_h == 1 ? (n + g.bn - 1) / g.bn : n * g.tiles_h;
    int const bn_tile = (cout > 64 && static_cast<long long>(tiles_m) * ((cout + 127) / 128) >= device_sm_count()) ? 128 : 64;
    CUtensorMap ma, mb;
    if (!make_pixel_map(&ma, x, cin, x_pitch, wd, h, n, 1, g.bw, g.bh, g.bn, stride_w, stride_h))
        return map_fail("conv_fprop_tc: activation tensor map rejected");
    long long const ktot = static_cast<long long>(kh) * kw * cin;
    long long wd2[2] = {ktot, cout}, ws2[2] = {1, ktot};
    int wb[2] = {32, bn_tile}, we[2] = {1, 1};
    if (!make_map_nd(&mb, w, 2, wd2, ws2, wb, we)) return map_fail("conv_fprop_tc: weight tensor map rejected");
    dim3 const grid(static_cast<unsigned>(tiles_m), static_cast<unsigned>((cout + bn_tile - 1) / bn_tile));
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    cudaError_t const err = bn_tile == 128 ? launch_conv<128, 0>(ma, mb, g, y, bias, grid, st) : launch_conv<64, 0>(ma, mb, g, y, bias, grid, st);
    return cuda_check(err, "conv_fprop_tc launch");
}

// dx[n, h, w, :cin] (pixel pitch dx_pitch) = Σ_taps dy[n, h + pad − r, w + pad − s, :cout] · w[cout, r, s, cin]   (stride-1 convolution)
// (+ addend[n, h, w, :cin], same pitch: the gradient arriving over a skip connection, summed in the epilogue)
extern "C" GB_API int gb_conv_dgrad_tc(float const* dy, float const* w, float* dx, int n, int h, int wd, int cin, int dx_pitch, int cout,
                                       int dy_pitch, int kh, int kw, int pad_h, int pad_w, float const* addend, void* stream) {
    if (n < 1 || h < 1 || wd < 1 || cin < 1 || cout < 1 || kh < 1 || kw < 1 || pad_h < 0 || pad_w < 0) return cuda_fail("conv_dgrad_tc: bad geometry");
    if ((kh * kw * cin) % 4 || dy_pitch % 4 || dy_pitch < cout || dx_pitch < cin || !aligned16(dy) || !aligned16(w))
        return cuda_fail("conv_dgrad_tc: kh*kw*cin and the gradient pixel pitch must be multiples of 4 floats, pointers 16-byte aligned");
    int const ho = h + 2 * pad_h - kh + 1, wo = wd + 2 * pad_w - kw + 1;
    if (ho < 1 || wo < 1) return cuda_fail("conv_dgrad_tc: empty output gradient");
    ConvGeom g{};
    if (fill_taps(g, kh, kw, pad_h, pad_w, true)) return cuda_fail("conv_dgrad_tc: more than 32 filter taps");
    if (pick_pixel_tile(g, n, h, wd)) return cuda_fail("conv_dgrad_tc: input width must be 1..128");
    g.k_extent = cout; g.kchunks = (cout + 31) / 32; g.n_total = cin; g.m_total = n * h * wd;
    g.sh = 1; g.sw = 1; g.hin = ho; g.win = wo; g.relu = 0; g.ldc = dx_pitch; g.addend = addend;
    int const tiles_m = g.tiles_h == 1 ? (n + g.bn - 1) / g.bn : n * g.tiles_h;
    int const bn_tile = (cin > 64 && static_cast<long long>(tiles_m) * ((cin + 127) / 128) >= device_sm_count()) ? 128 : 64;
    CUtensorMap ma, mb;
    if (!make_pixel_map(&ma, dy, cout, dy_pitch, wo, ho, n, 1, g.bw, g.bh, g.bn, 1, 1)) return map_fail("conv_dgrad_tc: gradient tensor map rejected");
    long long const ktot = static_cast<long long>(kh) * kw * cin;
    long long wd2[2] = {ktot, cout}, ws2[2] = {1, ktot};
    int wb[2] = {32, 32}, we[2] = {1, 1};
    mn_descriptor_defaults(g);
    if (!make_map_nd(&mb, w, 2, wd2, ws2, wb, we, true)) return map_fail("conv_dgrad_tc: weight tensor map rejected");
    dim3 const grid(static_cast<unsigned>(tiles_m), static_cast<unsigned>((cin + bn_tile - 1) / bn_tile));
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    cudaError_t const err = bn_tile == 128 ? launch_conv<128, 1>(ma, mb, g, dx, nullptr, grid, st) : launch_conv<64, 1>(ma, mb, g, dx, nullptr, grid, st);
    return cuda_check(err, "conv_dgrad_tc launch");
}

// dw[g][cout, kh, kw, cin] (g-th matrix at dw + g·dw_group_stride, row stride dw_ld, dw_cols <= cin columns stored per tap)
//   = Σ over the pixels of images g·b … g·b+b−1 of dyᵀ · patches(x)
extern "C" GB_API int gb_conv_wgrad_tc(float const* x, float const* dy, float* dw, int groups, int b, int h, int wd, int cin, int x_pitch, int cout,
                                       int dy_pitch, int kh, int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dw_cols, int64_t dw_ld,
                                       int64_t dw_group_stride, void* stream) {
    if (groups < 1 || b < 1 || h < 1 || wd < 1 || cin < 1 || cout < 1 || kh < 1 || kw < 1 || stride_h < 1 || stride_w < 1 || pad_h < 0 || pad_w < 0)
        return cuda_fail("conv_wgrad_tc: bad geometry");
    if (x_pitch % 4 || dy_pitch % 4 || x_pitch < cin || dy_pitch < cout || !aligned16(x) || !aligned16(dy))
        return cuda_fail("conv_wgrad_tc: pixel pitches must be multiples of 4 floats, pointers 16-byte aligned");
    if (stride_h > 8 || stride_w > 8) return cuda_fail("conv_wgrad_tc: stride > 8");
    int const ho = (h + 2 * pad_h - kh) / stride_h + 1, wo = (wd + 2 * pad_w - kw) / stride_w + 1;
    int const hw = ho * wo;
    if (wo > 32 || (32 % wo) != 0 || !((hw % 32) == 0 || (32 % hw) == 0))
        return cuda_fail("conv_wgrad_tc: output plane must tile into 32-pixel chunks (width a power of two <= 32)");
    ConvGeom g{};
    if (fill_taps(g, kh, kw, pad_h, pad_w, false)) return cuda_fail("conv_wgrad_tc: more than 32 filter taps");
    g.wo = wo; g.ho = ho; g.nimg = b;
    g.bw = wo; g.bh = hw >= 32 ? 32 / wo : ho; g.bn = hw >= 32 ? 1 : 32 / hw;
    g.tiles_h = 1;
    long long const pixels = static_cast<long long>(b) * hw;
    g.kchunks = static_cast<int>(hw >= 32 ? pixels / 32 : (b + g.bn - 1) / g.bn);
    g.k_extent = 0; g.n_total = cin; g.m_total = cout;
    g.sh = stride_h; g.sw = stride_w; g.hin = h; g.win = wd; g.relu = 0;
    if (dw_cols < 1 || dw_cols > cin || dw_ld < static_cast<long long>(kh) * kw * dw_cols) return cuda_fail("conv_wgrad_tc: bad output layout");
    g.n_store = dw_cols; g.ldc = dw_ld; g.c_group_stride = dw_group_stride;
    int const bn_tile = (cin > 64 && static_cast<long long>(groups) * g.taps * ((cin + 127) / 128) * ((cout + 127) / 128) >= device_sm_count()) ? 128 : 64;
    g.n_tiles_per_tap = (cin + bn_tile - 1) / bn_tile;
    CUtensorMap ma, mb;
    long long ad[3] = {cout, pixels, groups}, as[3] = {1, dy_pitch, pixels * dy_pitch};
    int ab[3] = {32, 32, 1}, ae[3] = {1, 1, 1};
    mn_descriptor_defaults(g);
    if (!make_map_nd(&ma, dy, 3, ad, as, ab, ae, true)) return map_fail("conv_wgrad_tc: gradient tensor map rejected");
    if (!make_pixel_map(&mb, x, cin, x_pitch, wd, h, b, groups, g.bw, g.bh, g.bn, stride_w, stride_h, true))
        return map_fail("conv_wgrad_tc: activation tensor map rejected");
    dim3 const grid(static_cast<unsigned>(g.taps * g.n_tiles_per_tap), static_cast<unsigned>((cout + 127) / 128), static_cast<unsigned>(groups));
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    cudaError_t const err = bn_tile == 128 ? launch_conv<128, 2>(ma, mb, g, dw, nullptr, grid, st) : launch_conv<64, 2>(ma, mb, g, dw, nullptr, grid, st);
    return cuda_check(err, "conv_wgrad_tc launch");
}
