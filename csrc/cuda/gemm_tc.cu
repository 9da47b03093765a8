// gemm_tc: C[M,N] = A[M,K] · B[N,K]ᵀ (+ bias[N]) (ReLU), fp32 output, on the 5th-generation tensor
// cores: TMA 2-D tiled loads (128-byte swizzle) → 4-stage mbarrier ring → tcgen05.mma
// (kind::tf32 for fp32 operands, kind::f16 for bf16 operands) accumulating in TMEM → tcgen05.ld
// epilogue with fused bias / ReLU. Warp-specialised: warp 0 = TMA producer, warp 1 = MMA issuer
// (one elected thread), warps 2-5 = epilogue (one TMEM lane quadrant each). Persistent CTAs, two TMEM
// accumulators: the epilogue of one tile overlaps the main loop of the next.
//
// This is the framework's own GEMM for nn.Linear forward / dgrad / wgrad (the reference relies on
// cuBLAS through ATen: garfieldpp/models/nets.py:47-57, 68-77). Both operands are
// K-major ("NT"), which is nn.Linear's natural layout: y = x[M,K] · W[N,K]ᵀ.
//
// Tile: 128 (M) × BN (N ∈ {128, 256}) × 128 bytes of K per stage (32 tf32 or 64 bf16 elements =
// 4 MMAs of K = 32 bytes). Out-of-range rows/columns are zero-filled by TMA and masked in the epilogue,
// so M, N, K need not be tile multiples (K·sizeof(T) must be a multiple of 16 bytes for the tensor map).

#include <cuda.h>

#include <cstdlib>
#include <cstring>
#include <mutex>

#include "garfield_b200.h"
#include "launch.cuh"
#include "ptx.cuh"

namespace gb {
namespace {

constexpr int kGemmThreads = 32 * 6;
constexpr int kGemmStages = 4;
constexpr int kBM = 128;
constexpr int kStagePitch = 36;  // floats per staged epilogue row (32 + 4: 16-byte aligned, bank-conflict-free)
constexpr int kRowBytes = 128;  // bytes of K per stage and per smem row (one swizzle atom wide)

// Epilogue of one warp: its 32 TMEM lanes (rows row0 .. row0+31 of C) × BN accumulator columns starting at column col_base →
// bias / ReLU → global memory, staged through shared memory so that the stores are 128-byte row segments.
template <int BN>
__device__ __forceinline__ void store_accumulator(uint32_t acc, int row0, int col_base, float* __restrict__ C, const float* __restrict__ bias, int M,
                                                  int N, long long ldc, int relu, float* stage, int lane) {
    bool const vec_ok = (reinterpret_cast<uintptr_t>(C) & 15u) == 0 && (ldc & 3) == 0;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t v[32];
        ptx::tmem_ld_32x32b_x16(acc + static_cast<uint32_t>(c0), reinterpret_cast<uint32_t(&)[16]>(v[0]));
        ptx::tmem_ld_32x32b_x16(acc + static_cast<uint32_t>(c0 + 16), reinterpret_cast<uint32_t(&)[16]>(v[16]));
        ptx::tmem_ld_wait();
        int const col0 = col_base + c0;
        if (col0 >= N || row0 >= M) continue;                         // warp-uniform
        // thread = one row × 32 columns in registers → stage through shared memory (pitch 36 floats: conflict-free
        // 16-byte accesses both ways) so that global stores are 128-byte row segments instead of 32 scattered 16-byte pieces
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            float4 o = make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]), __uint_as_float(v[4 * q + 2]),
                                   __uint_as_float(v[4 * q + 3]));
            if (bias != nullptr) {
                int const c = col0 + 4 * q;
                o.x += c < N ? bias[c] : 0.f; o.y += c + 1 < N ? bias[c + 1] : 0.f;
                o.z += c + 2 < N ? bias[c + 2] : 0.f; o.w += c + 3 < N ? bias[c + 3] : 0.f;
            }
            if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
            *reinterpret_cast<float4*>(stage + lane * kStagePitch + 4 * q) = o;
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int const r = i * 4 + (lane >> 3), c = (lane & 7) * 4;
            float4 const o = *reinterpret_cast<float4 const*>(stage + r * kStagePitch + c);
            int const grow = row0 + r, gcol = col0 + c;
            if (grow < M && gcol < N) {
                float* out = C + static_cast<long long>(grow) * ldc + gcol;
                if (vec_ok && gcol + 3 < N) {
                    *reinterpret_cast<float4*>(out) = o;
                } else {
                    out[0] = o.x;
                    if (gcol + 1 < N) out[1] = o.y;
                    if (gcol + 2 < N) out[2] = o.z;
                    if (gcol + 3 < N) out[3] = o.w;
                }
            }
        }
        __syncwarp();
    }
}

// Persistent form: one CTA per SM walks the tile list; TMEM holds TWO accumulators (2·BN columns) so the epilogue of
// tile j (tcgen05.ld → bias/ReLU → global stores) overlaps the TMA + MMA main loop of tile j+1. The smem ring runs on
// across tile boundaries. Consecutive CTAs take tiles that share the same B tile (column-major rasterisation), so a
// wave reads each B tile once from L2.
struct GemmSharedP {
    uint64_t full[kGemmStages];
    uint64_t empty[kGemmStages];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint32_t tmem_base;
};

template <int BN, bool TF32>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_persistent_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, float* __restrict__ C,
                          const float* __restrict__ bias, int M, int N, int K, long long ldc, int relu, int tiles_m, int tiles_n) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    __shared__ GemmSharedP sh;
    __shared__ __align__(16) float epi_stage[4][32 * kStagePitch];   // per epilogue warp: 32 rows × 32 columns (+ padding)
    constexpr uint32_t kABytes = kBM * kRowBytes, kBBytes = BN * kRowBytes, kStageBytes = kABytes + kBBytes;
    constexpr int kElemsPerStage = TF32 ? 32 : 64;
    int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int const k_steps = (K + kElemsPerStage - 1) / kElemsPerStage;
    int const total = tiles_m * tiles_n;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kGemmStages; ++s) {
            ptx::mbar_init(&sh.full[s], 1);
            ptx::mbar_init(&sh.empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(&sh.acc_full[b], 1);
            ptx::mbar_init(&sh.acc_empty[b], 4);     // one arrival per epilogue warp
        }
        ptx::fence_mbar_init();
        ptx::prefetch_tensormap(&map_a);
        ptx::prefetch_tensormap(&map_b);
    }
    if (warp == 1) ptx::tmem_alloc<2 * BN>(&sh.tmem_base);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    uint32_t const tmem = sh.tmem_base;

    if (warp == 0) {
        if (ptx::elect_one()) {
            int it = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                int const tile_m = t % tiles_m, tile_n = t / tiles_m;
                for (int kk = 0; kk < k_steps; ++kk, ++it) {
                    int const s = it % kGemmStages;
                    uint32_t const phase = static_cast<uint32_t>((it / kGemmStages) & 1);
                    ptx::mbar_wait(&sh.empty[s], phase ^ 1u);
                    ptx::mbar_arrive_expect_tx(&sh.full[s], kStageBytes);
                    unsigned char* dst = smem_raw + static_cast<size_t>(s) * kStageBytes;
                    ptx::tma_load_2d(dst, &map_a, kk * kElemsPerStage, tile_m * kBM, &sh.full[s]);
                    ptx::tma_load_2d(dst + kABytes, &map_b, kk * kElemsPerStage, tile_n * BN, &sh.full[s]);
                }
            }
        }
    } else if (warp == 1) {
        if (ptx::elect_one()) {
            uint32_t const idesc = TF32 ? ptx::make_idesc(2u, 2u, kBM, BN) : ptx::make_idesc(1u, 1u, kBM, BN);
            int it = 0, j = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x, ++j) {
                int const buf = j & 1;
                ptx::mbar_wait(&sh.acc_empty[buf], static_cast<uint32_t>(((j >> 1) & 1) ^ 1));   // epilogue drained this accumulator
                ptx::tc_fence_after();
                uint32_t const acc = tmem + static_cast<uint32_t>(buf * BN);
                for (int kk = 0; kk < k_steps; ++kk, ++it) {
                    int const s = it % kGemmStages;
                    uint32_t const phase = static_cast<uint32_t>((it / kGemmStages) & 1);
                    ptx::mbar_wait(&sh.full[s], phase);
                    ptx::tc_fence_after();
                    uint32_t const a_base = ptx::smem_u32(smem_raw + static_cast<size_t>(s) * kStageBytes);
                    uint32_t const b_base = a_base + kABytes;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        uint64_t const adesc = ptx::make_smem_desc(a_base + k * 32u, 16u, 1024u, 2u);
                        uint64_t const bdesc = ptx::make_smem_desc(b_base + k * 32u, 16u, 1024u, 2u);
                        if (TF32) ptx::mma_tf32_ss(acc, adesc, bdesc, idesc, (kk > 0 || k > 0) ? 1u : 0u);
                        else ptx::mma_f16_ss(acc, adesc, bdesc, idesc, (kk > 0 || k > 0) ? 1u : 0u);
                    }
                    ptx::mma_commit(&sh.empty[s]);
                }
                ptx::mma_commit(&sh.acc_full[buf]);
            }
        }
    } else {
        int const quad = warp & 3;
        int j = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x, ++j) {
            int const tile_m = t % tiles_m, tile_n = t / tiles_m;
            int const buf = j & 1;
            ptx::mbar_wait(&sh.acc_full[buf], static_cast<uint32_t>((j >> 1) & 1));
            ptx::tc_fence_after();
            uint32_t const acc = tmem + static_cast<uint32_t>(buf * BN) + (static_cast<uint32_t>(quad * 32) << 16);
            int const row0 = tile_m * kBM + quad * 32;                       // first row of this warp's TMEM lane quadrant
            store_accumulator<BN>(acc, row0, tile_n * BN, C, bias, M, N, ldc, relu, &epi_stage[quad][0], lane);
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&sh.acc_empty[buf]);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<2 * BN>(tmem);
}

// ------------------------------------------------------------------ CTA-pair form ---
//
// Large problems: two CTAs of a cluster (one TPC) compute ONE 256 × 256 tile with tcgen05.mma.cta_group::2. Each CTA stages its
// own 128 rows of A and 128 of the tile's 256 B rows — 32 KB per stage and CTA where the single-CTA kernel moves 48 KB for a
// 128 × 256 tile — so the L2 → shared-memory traffic per FLOP drops by a third and the ring gets 6 stages. The leader (cluster
// rank 0) owns the full / accumulator-empty barriers: the TMA loads of both CTAs complete on the leader's full barrier, the
// leader's MMA thread issues for the pair and its commits arrive (multicast) on both CTAs' empty / accumulator-full barriers;
// each CTA drains its own 128 TMEM lanes. All loads are plain (no multicast) tiled loads of this GPU's operands.
constexpr int kPairStages = 6;
constexpr int kPairBN = 256;
struct GemmSharedPair {
    uint64_t full[kPairStages];
    uint64_t empty[kPairStages];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint32_t tmem_base;
};

// Tile order of the pair kernel: groups of kRasterM rows of tiles, column-major inside a group, so that the ≈74 tiles in flight
// cover about 8 × 9 tiles — (8 + 9) operand panels in L2 instead of 74 + 1 (8192² × 4096 bf16: 34 MB instead of 150 MB).
constexpr int kRasterM = 8;
__device__ __forceinline__ void pair_tile_of(int t, int tiles_m, int tiles_n, int& tile_m, int& tile_n) {
    int const per_group = kRasterM * tiles_n;
    int const group = t / per_group, within = t - group * per_group;
    int const height = min(kRasterM, tiles_m - group * kRasterM);
    tile_m = group * kRasterM + within % height;
    tile_n = within / height;
}

template <bool TF32>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_pair_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, float* __restrict__ C,
                    const float* __restrict__ bias, int M, int N, int K, long long ldc, int relu, int tiles_m, int tiles_n) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    __shared__ GemmSharedPair sh;
    __shared__ __align__(16) float epi_stage[4][32 * kStagePitch];
    constexpr int BN = kPairBN;
    constexpr uint32_t kABytes = kBM * kRowBytes, kBBytes = (BN / 2) * kRowBytes, kStageBytes = kABytes + kBBytes;
    constexpr int kElemsPerStage = TF32 ? 32 : 64;
    int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t const rank = ptx::cluster_ctarank();
    bool const leader = rank == 0;
    int const k_steps = (K + kElemsPerStage - 1) / kElemsPerStage;
    int const total = tiles_m * tiles_n;                  // 256 × 256 pair tiles
    int const pair = blockIdx.x >> 1, pairs = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kPairStages; ++s) {
            ptx::mbar_init(&sh.full[s], 1);               // leader's producer arrives; bytes of both CTAs' loads
            ptx::mbar_init(&sh.empty[s], 1);              // the leader's commit
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(&sh.acc_full[b], 1);           // the leader's commit
            ptx::mbar_init(&sh.acc_empty[b], 8);          // 4 epilogue warps of each CTA (used in the leader)
        }
        ptx::fence_mbar_init();
        ptx::prefetch_tensormap(&map_a);
        ptx::prefetch_tensormap(&map_b);
    }
    if (warp == 1) ptx::tmem_alloc_pair<2 * BN>(&sh.tmem_base);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();                                  // the peer's barriers exist before anything signals them
    ptx::tc_fence_after();
    uint32_t const tmem = sh.tmem_base;

    if (warp == 0) {
        if (ptx::elect_one()) {
            int it = 0;
            for (int t = pair; t < total; t += pairs) {
                int tile_m, tile_n;
                pair_tile_of(t, tiles_m, tiles_n, tile_m, tile_n);
                for (int kk = 0; kk < k_steps; ++kk, ++it) {
                    int const s = it % kPairStages;
                    uint32_t const phase = static_cast<uint32_t>((it / kPairStages) & 1);
                    ptx::mbar_wait(&sh.empty[s], phase ^ 1u);
                    if (leader) ptx::mbar_arrive_expect_tx(&sh.full[s], 2u * kStageBytes);
                    uint32_t const bar = ptx::map_to_cta(&sh.full[s], 0u);
                    unsigned char* dst = smem_raw + static_cast<size_t>(s) * kStageBytes;
                    ptx::tma_load_2d_pair(dst, &map_a, kk * kElemsPerStage, tile_m * 2 * kBM + static_cast<int>(rank) * kBM, bar);
                    ptx::tma_load_2d_pair(dst + kABytes, &map_b, kk * kElemsPerStage, tile_n * BN + static_cast<int>(rank) * (BN / 2), bar);
                }
            }
        }
    } else if (warp == 1) {
        if (leader && ptx::elect_one()) {
            uint32_t const idesc = TF32 ? ptx::make_idesc(2u, 2u, 2 * kBM, BN) : ptx::make_idesc(1u, 1u, 2 * kBM, BN);
            int it = 0, j = 0;
            for (int t = pair; t < total; t += pairs, ++j) {
                int const buf = j & 1;
                ptx::mbar_wait(&sh.acc_empty[buf], static_cast<uint32_t>(((j >> 1) & 1) ^ 1));   // both CTAs drained this accumulator
                ptx::tc_fence_after();
                uint32_t const acc = tmem + static_cast<uint32_t>(buf * BN);
                for (int kk = 0; kk < k_steps; ++kk, ++it) {
                    int const s = it % kPairStages;
                    uint32_t const phase = static_cast<uint32_t>((it / kPairStages) & 1);
                    ptx::mbar_wait(&sh.full[s], phase);
                    ptx::tc_fence_after();
                    uint32_t const a_base = ptx::smem_u32(smem_raw + static_cast<size_t>(s) * kStageBytes);
                    uint32_t const b_base = a_base + kABytes;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        uint64_t const adesc = ptx::make_smem_desc(a_base + k * 32u, 16u, 1024u, 2u);
                        uint64_t const bdesc = ptx::make_smem_desc(b_base + k * 32u, 16u, 1024u, 2u);
                        if (TF32) ptx::mma_tf32_ss_pair(acc, adesc, bdesc, idesc, (kk > 0 || k > 0) ? 1u : 0u);
                        else ptx::mma_f16_ss_pair(acc, adesc, bdesc, idesc, (kk > 0 || k > 0) ? 1u : 0u);
                    }
                    ptx::mma_commit_pair(&sh.empty[s], 3);
                }
                ptx::mma_commit_pair(&sh.acc_full[buf], 3);
            }
        }
    } else {
        int const quad = warp & 3;
        int j = 0;
        for (int t = pair; t < total; t += pairs, ++j) {
            int tile_m, tile_n;
            pair_tile_of(t, tiles_m, tiles_n, tile_m, tile_n);
            int const buf = j & 1;
            ptx::mbar_wait(&sh.acc_full[buf], static_cast<uint32_t>((j >> 1) & 1));
            ptx::tc_fence_after();
            uint32_t const acc = tmem + static_cast<uint32_t>(buf * BN) + (static_cast<uint32_t>(quad * 32) << 16);
            int const row0 = tile_m * 2 * kBM + static_cast<int>(rank) * kBM + quad * 32;
            store_accumulator<BN>(acc, row0, tile_n * BN, C, bias, M, N, ldc, relu, &epi_stage[quad][0], lane);
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_cluster_relaxed(ptx::map_to_cta(&sh.acc_empty[buf], 0u));
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();                                  // nobody frees TMEM or exits while the peer may still signal or read
    if (warp == 1) ptx::tmem_dealloc_pair<2 * BN>(tmem);
}

using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn encode_fn() {
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

// Row-major [rows, K] matrix with K contiguous; box = (128 bytes of K) × box_rows, 128B swizzle.
bool make_map(CUtensorMap* map, const void* base, long long rows, long long k, long long ld_elems, int elem_bytes, int box_rows) {
    EncodeFn fn = encode_fn();
    if (fn == nullptr) return false;
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(k), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld_elems) * static_cast<cuuint64_t>(elem_bytes)};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(kRowBytes / elem_bytes), static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    CUtensorMapDataType const dtype = elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    return fn(map, dtype, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int BN, bool TF32>
cudaError_t launch_gemm(CUtensorMap const& ma, CUtensorMap const& mb, float* C, const float* bias, int M, int N, int K, long long ldc,
                        int relu, cudaStream_t stream) {
    auto kern = gemm_tc_persistent_kernel<BN, TF32>;
    size_t const smem = static_cast<size_t>(kGemmStages) * (kBM + BN) * kRowBytes;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (err != cudaSuccess) return err;
    int const tiles_n = (N + BN - 1) / BN, tiles_m = (M + kBM - 1) / kBM;
    long long const total = static_cast<long long>(tiles_m) * tiles_n;
    int const grid = static_cast<int>(total < device_sm_count() ? total : device_sm_count());
    kern<<<grid, kGemmThreads, smem, stream>>>(ma, mb, C, bias, M, N, K, ldc, relu, tiles_m, tiles_n);
    return cudaGetLastError();
}

template <bool TF32>
cudaError_t launch_gemm_pair(CUtensorMap const& ma, CUtensorMap const& mb, float* C, const float* bias, int M, int N, int K, long long ldc,
                             int relu, cudaStream_t stream) {
    auto kern = gemm_tc_pair_kernel<TF32>;
    size_t const smem = static_cast<size_t>(kPairStages) * (kBM + kPairBN / 2) * kRowBytes;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (err != cudaSuccess) return err;
    int const tiles_n = (N + kPairBN - 1) / kPairBN, tiles_m = (M + 2 * kBM - 1) / (2 * kBM);
    long long const total = static_cast<long long>(tiles_m) * tiles_n;
    long long const max_pairs = device_sm_count() / 2;
    unsigned const pairs = static_cast<unsigned>(total < max_pairs ? total : max_pairs);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * pairs);
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, ma, mb, C, bias, M, N, K, ldc, relu, tiles_m, tiles_n);
}

// GARFIELD_GEMM_PAIR=0 keeps the single-CTA kernel for every shape (A/B measurements).
bool pair_enabled() {
    static int const on = [] { char const* v = std::getenv("GARFIELD_GEMM_PAIR"); return (v != nullptr && v[0] == '0') ? 0 : 1; }();
    return on != 0;
}

}  // namespace
}  // namespace gb

// C[M,N] (fp32, row stride ldc) = A[M,K] · B[N,K]ᵀ (+ bias) (ReLU). dtype: 0 = fp32 operands (TF32 MMA),
// 1 = bf16 operands. lda / ldb are row strides in elements; base pointers and row strides must be
// 16-byte aligned.
extern "C" GB_API int gb_gemm_nt(void const* A, void const* B, float* C, float const* bias, int M, int N, int K, int64_t lda, int64_t ldb,
                                 int64_t ldc, int dtype, int relu, void* stream) {
    using namespace gb;
    if (M < 1 || N < 1 || K < 1) return cuda_fail("gemm: empty problem");
    if (dtype != 0 && dtype != 1) return cuda_fail("gemm: dtype must be 0 (fp32/tf32) or 1 (bf16)");
    int const eb = dtype == 0 ? 4 : 2;
    if ((reinterpret_cast<uintptr_t>(A) & 15u) || (reinterpret_cast<uintptr_t>(B) & 15u) || (lda * eb) % 16 || (ldb * eb) % 16)
        return cuda_fail("gemm: operands must be 16-byte aligned with row strides that are multiples of 16 bytes");
    CUtensorMap ma, mb;
    cudaStream_t const st0 = static_cast<cudaStream_t>(stream);
    // CTA pairs once there are at least as many 256 × 256 tiles as pairs on the GPU (smaller problems fill more SMs with 128-row tiles)
    long long const pair_tiles = static_cast<long long>((M + 255) / 256) * ((N + 255) / 256);
    if (pair_enabled() && M > 128 && N > 128 && pair_tiles >= device_sm_count() / 2) {
        if (!make_map(&ma, A, M, K, lda, eb, kBM) || !make_map(&mb, B, N, K, ldb, eb, kPairBN / 2))
            return cuda_fail("gemm: cuTensorMapEncodeTiled failed (driver entry point unavailable or bad shape)");
        cudaError_t const e = dtype == 0 ? launch_gemm_pair<true>(ma, mb, C, bias, M, N, K, ldc, relu, st0)
                                         : launch_gemm_pair<false>(ma, mb, C, bias, M, N, K, ldc, relu, st0);
        if (e == cudaSuccess) return 0;
        // a device that cannot co-schedule CTA pairs (partitioned GPU, cluster launch refused): the single-CTA kernel below
        if (e != cudaErrorInvalidClusterSize && e != cudaErrorLaunchOutOfResources && e != cudaErrorInvalidConfiguration && e != cudaErrorNotSupported)
            return cuda_check(e, "gemm pair launch");
        (void)cudaGetLastError();
    }
    int const bn = N > 128 ? 256 : 128;
    if (!make_map(&ma, A, M, K, lda, eb, kBM) || !make_map(&mb, B, N, K, ldb, eb, bn))
        return cuda_fail("gemm: cuTensorMapEncodeTiled failed (driver entry point unavailable or bad shape)");
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    cudaError_t err;
    if (dtype == 0) err = bn == 256 ? launch_gemm<256, true>(ma, mb, C, bias, M, N, K, ldc, relu, st) : launch_gemm<128, true>(ma, mb, C, bias, M, N, K, ldc, relu, st);
    else err = bn == 256 ? launch_gemm<256, false>(ma, mb, C, bias, M, N, K, ldc, relu, st) : launch_gemm<128, false>(ma, mb, C, bias, M, N, K, ldc, relu, st);
    return cuda_check(err, "gemm launch");
}
