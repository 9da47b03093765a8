// gram_select_tc: pairwise squared distances of n <= 32 gradient rows from a Gram matrix
// accumulated on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulator in TMEM),
// followed by the same in-kernel selection logic as the SIMT kernel (gar_select.cuh).
//
// Why: the exact SIMT kernel (gar_pairwise.cu) needs n(n-1)/2 · 2 FP32 ops per coordinate and is
// ALU-bound beyond n ≈ 16 (2.2 ms at n = 23, d = 23.5 M against a 0.35 ms HBM floor). The Gram
// form costs one MMA per 8 coordinates for ALL pairs at once, so the pass is memory-bound again.
// Counterpart of the reference's K1+K2 launches (native/py_krum/krum.cu:87-98,
// py_bulyan/bulyan.cu:286-293) — the north-star "TMA → tcgen05/TMEM Gram tile" path.
//
// Operand layout trick (no swizzle, K-major "interleaved" canonical layout): a core matrix is
// 8 rows × 16 bytes stored contiguously. A plain 1-D TMA bulk copy of one peer's contiguous
// chunk of CH floats therefore lands as CH/32 core matrices if we declare
//     MMA row  = 8·peer + j      (j = which 16-byte piece of each 128-byte line)
//     K index  = 4·line + k      (k = float within the piece)
// with LBO (K-adjacent core matrices) = 128 B and SBO (next 8-row group = next peer) = 4·CH B.
// Row (8p+j) and row (8p'+j) then hold THE SAME coordinates of peers p and p', so
//     G[p][p'] = Σ_j D[8p+j][8p'+j]
// — the j ≠ j' entries of D are by-products that are simply not read. No data movement other
// than the bulk copies is needed, and rows may be peer-mapped memory of other GPUs.
//
// M = 128 rows per MMA = 16 peers; n <= 16 uses one M-tile, n <= 32 two (D tiles in disjoint TMEM
// column ranges); N = 8·⌈n/2⌉·2 columns. The accumulators stay in TMEM for the whole kernel; one
// epilogue per CTA reads them with tcgen05.ld, folds the 8 j-lanes with shuffles and writes the
// CTA's partial Gram matrix; the last CTA reduces the partials in fixed order (double), forms
// D_ij = G_ii + G_jj − 2·G_ij (clamped at 0) and runs the selection.
//
// Precision: operands are read as TF32 (10-bit mantissa), products accumulate in FP32 in TMEM.
// Squared norms and dot products carry a relative error of a few 1e-4, so after the
// G_ii + G_jj − 2·G_ij cancellation the distances are good to ~1e-3·(‖x_i‖² + ‖x_j‖²); selection
// outcomes match the exact kernel unless two scores are closer than that. The exact SIMT kernel
// remains the default for n <= 16 and the oracle in the tests.

#include <cfloat>
#include <cstdio>
#include <cmath>

#include "gar_common.cuh"
#include "gar_select.cuh"
#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

constexpr int kTcThreads = 32 * 6;     // warp 0 producer, warp 1 MMA, warps 2-5 epilogue (TMEM quadrants 2,3,0,1)
constexpr int kTcStages = 3;
constexpr int kTcMaxRows = 32;

struct GramArgs {
    RowTable rows;
    int n;
    int chunk;              // floats per peer per stage (multiple of 64)
    int n_cols;             // N of the MMA: 8 * (n rounded up to even)  (multiple of 16, <= 256)
    int m_tiles;            // 1 (n <= 16) or 2
    long long d;
    float* partials;        // gridDim.x × n × n partial Gram matrices
    unsigned int* counter;
    float* dist;
    float* W;
    int* info;
    SelectArgs sel;
};

struct TcShared {
    uint64_t full[kTcStages];
    uint64_t empty[kTcStages];
    uint64_t accum_done;
    uint32_t tmem_base;
    unsigned int is_last;
};

template <int COLS>
__global__ void __launch_bounds__(kTcThreads, 1) gram_tc_kernel(const __grid_constant__ GramArgs a) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    __shared__ TcShared sh;
    float* stages = reinterpret_cast<float*>(smem_raw);
    int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int const n = a.n, nn = n * n;
    int const chunk = a.chunk;
    int const peers_pad = a.m_tiles * 16;                     // row groups the A descriptors span
    size_t const stage_floats = static_cast<size_t>(peers_pad) * chunk;
    long long const full_chunks = a.d / chunk;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kTcStages; ++s) {
            ptx::mbar_init(&sh.full[s], 1);
            ptx::mbar_init(&sh.empty[s], 1);
        }
        ptx::mbar_init(&sh.accum_done, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 1) ptx::tmem_alloc<COLS>(&sh.tmem_base);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    uint32_t const tmem = sh.tmem_base;

    long long my_chunks = 0;
    for (long long t = blockIdx.x; t < full_chunks; t += gridDim.x) ++my_chunks;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer ---
        if (ptx::elect_one()) {
            uint32_t const row_bytes = static_cast<uint32_t>(chunk) * 4u;
            long long it = 0;
            for (long long t = blockIdx.x; t < full_chunks; t += gridDim.x, ++it) {
                int const s = static_cast<int>(it % kTcStages);
                uint32_t const phase = static_cast<uint32_t>((it / kTcStages) & 1);
                ptx::mbar_wait(&sh.empty[s], phase ^ 1u);
                ptx::mbar_arrive_expect_tx(&sh.full[s], row_bytes * static_cast<uint32_t>(n));
                float* dst = stages + s * stage_floats;
                for (int p = 0; p < n; ++p) ptx::bulk_g2s(dst + static_cast<size_t>(p) * chunk, a.rows.p[p] + t * chunk, row_bytes, &sh.full[s]);
            }
        }
    } else if (warp == 1) {
        // -------------------------------------------------------------------- MMA issuer ---
        if (ptx::elect_one()) {
            uint32_t const idesc = ptx::make_idesc(2u, 2u, 128u, static_cast<uint32_t>(a.n_cols));
            uint32_t const sbo = static_cast<uint32_t>(chunk) * 4u;   // next 8-row group = next peer
            int const ksteps = chunk / 64;                             // one MMA consumes 2 core matrices = 256 B per row group
            long long it = 0;
            for (long long t = blockIdx.x; t < full_chunks; t += gridDim.x, ++it) {
                int const s = static_cast<int>(it % kTcStages);
                uint32_t const phase = static_cast<uint32_t>((it / kTcStages) & 1);
                ptx::mbar_wait(&sh.full[s], phase);
                ptx::tc_fence_after();
                uint32_t const base = ptx::smem_u32(stages + s * stage_floats);
                for (int k = 0; k < ksteps; ++k) {
                    uint64_t const bdesc = ptx::make_smem_desc(base + static_cast<uint32_t>(k) * 256u, 128u, sbo, 0u);
                    for (int mt = 0; mt < a.m_tiles; ++mt) {
                        uint64_t const adesc = ptx::make_smem_desc(base + static_cast<uint32_t>(mt) * 16u * sbo + static_cast<uint32_t>(k) * 256u, 128u, sbo, 0u);
                        ptx::mma_tf32_ss(tmem + static_cast<uint32_t>(mt * a.n_cols), adesc, bdesc, idesc, (it > 0 || k > 0) ? 1u : 0u);
                    }
                }
                ptx::mma_commit(&sh.empty[s]);      // slot reusable once these MMAs have read it
            }
            ptx::mma_commit(&sh.accum_done);         // all accumulation finished
        }
    } else {
        // ---------------------------------------------------------------------- epilogue ---
        float* mine = a.partials + static_cast<size_t>(blockIdx.x) * nn;
        int const quad = warp & 3;                   // TMEM lane quadrant this warp may read
        int const j = lane & 7;
        if (my_chunks > 0) {
            ptx::mbar_wait(&sh.accum_done, 0u);
            ptx::tc_fence_after();
            for (int mt = 0; mt < a.m_tiles; ++mt) {
                int const p = mt * 16 + quad * 4 + (lane >> 3);   // peer of this lane's D row
                for (int pc = 0; pc < a.n_cols / 8; ++pc) {        // column group = peer p'
                    uint32_t v[8];
                    uint32_t const taddr = tmem + (static_cast<uint32_t>(quad * 32) << 16) + static_cast<uint32_t>(mt * a.n_cols + pc * 8);
                    ptx::tmem_ld_32x32b_x8(taddr, v);
                    ptx::tmem_ld_wait();
                    float x = __uint_as_float(v[0]);
#pragma unroll
                    for (int c = 1; c < 8; ++c) x = (j == c) ? __uint_as_float(v[c]) : x;   // D[8p+j][8p'+j]
                    x += __shfl_xor_sync(0xffffffffu, x, 1);
                    x += __shfl_xor_sync(0xffffffffu, x, 2);
                    x += __shfl_xor_sync(0xffffffffu, x, 4);
                    if (j == 0 && p < n && pc < n) mine[p * n + pc] = x;
                }
            }
        } else {
            for (int e = threadIdx.x - 64; e < nn; e += kTcThreads - 64) mine[e] = 0.0f;
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<COLS>(tmem);

    __threadfence();
    if (threadIdx.x == 0) sh.is_last = (atomicAdd(a.counter, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!sh.is_last) return;
    __threadfence();

    // last CTA: Gram = Σ partials (fixed order, double) + exact FP32 tail; distances; selection
    double* G = reinterpret_cast<double*>(smem_raw);              // n × n (ring is idle now)
    long long const tail0 = full_chunks * chunk;
    for (int e = threadIdx.x; e < nn; e += blockDim.x) {
        int const i = e / n, jj = e % n;
        double total = 0.0;
        for (unsigned b = 0; b < gridDim.x; ++b) total += static_cast<double>(__ldcg(a.partials + static_cast<size_t>(b) * nn + e));
        double tail = 0.0;
        for (long long x = tail0; x < a.d; ++x) tail += static_cast<double>(a.rows.p[i][x]) * static_cast<double>(a.rows.p[jj][x]);
        G[e] = total + tail;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < nn; e += blockDim.x) {
        int const i = e / n, jj = e % n;
        float out = INFINITY;
        if (i != jj) {
            double const sq = G[i * n + i] + G[jj * n + jj] - (G[i * n + jj] + G[jj * n + i]);   // symmetric in (i, j)
            float const f = static_cast<float>(sq > 0.0 ? sq : 0.0);
            out = (sq == sq && fabsf(f) <= FLT_MAX) ? f : INFINITY;
        }
        a.dist[e] = out;
    }
    if (threadIdx.x == 0) *a.counter = 0u;
    __threadfence_block();
    __syncthreads();
    run_selection(a, *reinterpret_cast<SelectScratch*>(smem_raw));
}

}  // namespace
}  // namespace gb

extern "C" GB_API int64_t gb_gram_scratch_bytes(int n) {
    return 256 + static_cast<int64_t>(gb::device_sm_count()) * n * n * static_cast<int64_t>(sizeof(float));
}

// Same contract as gb_pairwise_select (outputs dist / W / info, caller scratch with a zero first
// word) for 2 <= n <= 32 rows that are 16-byte aligned; distances come from the TF32 tensor-core
// Gram matrix. Returns an error (and launches nothing) when the inputs do not qualify, so callers
// can fall back to gb_pairwise_select.
extern "C" GB_API int gb_gram_select_tc(float const* const* rows, int n, int64_t d, int mode, int f, int m, int neighbors, int squared,
                                        void* scratch, int64_t scratch_bytes, float* dist, float* W, int* info, void* stream) {
    using namespace gb;
    if (n < 2 || n > kTcMaxRows) return cuda_fail("gram_tc: supports 2 <= n <= 32 rows");
    if (d < 1) return cuda_fail("gram_tc: empty rows");
    if (dist == nullptr || scratch == nullptr) return cuda_fail("gram_tc: missing output/scratch buffer");
    if (mode != kSelectNone && (W == nullptr || info == nullptr)) return cuda_fail("gram_tc: selection needs W and info buffers");
    if (char const* bad = selection_error(mode, n, f, m, neighbors)) {
        static thread_local char msg[160];
        std::snprintf(msg, sizeof(msg), "gram_tc: %s", bad);
        return cuda_fail(msg);
    }
    GramArgs a{};
    for (int i = 0; i < n; ++i) a.rows.p[i] = rows[i];
    if (!rows_aligned(a.rows, n)) return cuda_fail("gram_tc: rows must be 16-byte aligned");
    a.n = n; a.d = d; a.dist = dist; a.W = W; a.info = info;
    a.sel.mode = mode; a.sel.f = f; a.sel.m = m; a.sel.neighbors = neighbors; a.sel.squared = squared;
    a.m_tiles = n <= 16 ? 1 : 2;
    a.n_cols = 8 * ((n + 1) / 2 * 2);
    a.chunk = n <= 16 ? 1024 : 512;           // 16 (32) row groups × 4 KiB (2 KiB) = 64 KiB per stage
    int const cols = a.m_tiles * a.n_cols;     // TMEM columns in use
    size_t smem = static_cast<size_t>(kTcStages) * a.m_tiles * 16 * a.chunk * sizeof(float);
    if (smem < sizeof(SelectScratch)) smem = sizeof(SelectScratch);
    long long const chunks = d / a.chunk;
    long long grid = device_sm_count();
    if (grid > chunks) grid = chunks;
    if (grid < 1) grid = 1;
    int64_t const need = 256 + grid * n * n * static_cast<int64_t>(sizeof(float));
    if (scratch_bytes < need) return cuda_fail("gram_tc: scratch buffer too small (see gb_gram_scratch_bytes)");
    a.counter = static_cast<unsigned int*>(scratch);
    a.partials = reinterpret_cast<float*>(static_cast<unsigned char*>(scratch) + 256);
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    cudaError_t err;
#define GB_LAUNCH_TC(COLS)                                                                                                   \
    do {                                                                                                                     \
        auto kern = gram_tc_kernel<COLS>;                                                                                    \
        err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));              \
        if (err == cudaSuccess) {                                                                                            \
            kern<<<static_cast<int>(grid), kTcThreads, smem, st>>>(a);                                                       \
            err = cudaGetLastError();                                                                                        \
        }                                                                                                                    \
    } while (0)
    if (cols <= 32) GB_LAUNCH_TC(32);
    else if (cols <= 64) GB_LAUNCH_TC(64);
    else if (cols <= 128) GB_LAUNCH_TC(128);
    else if (cols <= 256) GB_LAUNCH_TC(256);
    else GB_LAUNCH_TC(512);
#undef GB_LAUNCH_TC
    return cuda_check(err, "gram_tc launch");
}
