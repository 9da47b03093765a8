// gossip_contract: the decentralized (LEARN) exchange fused with the robust contract step — several agreement ROUNDS inside
// one persistent kernel per GPU.
//
// Reference (/root/reference/pytorch_impl/applications/LEARN/trainer.py:208-222, 244-258): every node repeatedly pulls
// the latest vector of every peer over RPC, aggregates them with the GAR and republishes the result — ⌈log2(i+1)⌉
// rounds per step for the aggregated gradients, one round for the models ("contract"). Each pull is an RPC round trip
// through host memory, each aggregation a separate call.
//
// Here one launch per GPU runs all rounds. Vectors live in double-buffered, peer-mapped arenas. Round r:
//   * every CTA streams tiles of ALL n nodes' vectors of parity r (local rows and peer rows over NVLink, plain TMA 1-D bulk
//     copies) through the mbarrier ring, applies the per-coordinate rule (register sorting network) and stores the result
//     into the parity r+1 buffer of EVERY node this GPU hosts (they all compute the same vector — one pass, k stores);
//   * a grid-wide barrier (all CTAs are co-resident: cooperative launch) closes the round locally; CTA 0 then publishes
//     "round r done" with a system-scope release store into this GPU's peer-mapped flag and polls the other GPUs' flags
//     (acquire loads, with a deadline — a peer that stalls costs one deadline, its stale buffer is just another Byzantine
//     input), and releases the grid into round r+1. The producer warp issues a generic→async proxy fence before it
//     starts pulling the peers' fresh buffers.
// No host involvement, no NCCL, no kernel relaunch between rounds.
//
// Only meaningful with ONE such kernel per GPU: kernels that wait on each other must run on different GPUs.

#include "gar_coordwise.cuh"
#include "garfield_b200.h"
#include "launch.cuh"

namespace gb {
namespace {

constexpr int kMaxLocalOut = 8;

struct GossipArgs {
    CoordArgs base;              // rule parameters; base.rows = inputs of parity 0
    RowTable rows1;              // inputs of parity 1
    float* out[2][kMaxLocalOut]; // out[p][k]: buffer of local node k that receives the result of a round reading parity p
    int n_out;
    int rounds, first;           // number of rounds; parity of the first round's inputs
    uint32_t* seq;               // local: flag sequence base, advanced by rounds − 1 at exit
    uint32_t* my_flag;           // peer-visible
    uint32_t* peer_flag[kMaxRows];
    int n_peers;
    unsigned int* counter;       // local grid-barrier counter, zero at launch
    unsigned long long timeout_ns;
    uint32_t* late;
};

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Returns false when the local deadline passed (a co-residency or logic fault must not hang the GPU).
__device__ __forceinline__ bool spin_until(const unsigned int* counter, unsigned int target, unsigned long long deadline) {
    while (ld_acquire_gpu(counter) < target) {
        if (ptx::globaltimer_ns() > deadline) return false;
        __nanosleep(64);
    }
    return true;
}

template <int NP, int RULE, int V>
__global__ void __launch_bounds__(kThreads) gossip_kernel(const __grid_constant__ GossipArgs g) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ TileRing ring;
    float* stages = reinterpret_cast<float*>(smem_raw);
    CoordArgs const& a = g.base;
    int const warp = threadIdx.x >> 5;
    int const tile = a.tile;
    long long const full_tiles = a.use_tma ? a.d / tile : 0;
    uint32_t const seq0 = *g.seq;
    unsigned int const G = gridDim.x;
    unsigned long long const local_deadline = ptx::globaltimer_ns() + 30ull * 1000000000ull;
    ring_init(ring);

    size_t const stage_floats = static_cast<size_t>(a.n) * tile;
    uint32_t const row_bytes = static_cast<uint32_t>(tile) * 4u;
    long long it = 0;     // ring position: runs on across rounds (producer and consumers walk the same tile sequence)

    for (int r = 0; r < g.rounds; ++r) {
        int const parity = (g.first + r) & 1;
        RowTable const& rows = parity ? g.rows1 : a.rows;
        if (warp == 0) {
            if (full_tiles > 0 && ptx::elect_one()) {
                if (r > 0) asm volatile("fence.proxy.async;" ::: "memory");   // peers' generic-proxy stores → our bulk (async-proxy) loads
                for (long long t = blockIdx.x; t < full_tiles; t += G, ++it) {
                    int const s = static_cast<int>(it % kStages);
                    ptx::mbar_wait(&ring.empty[s], static_cast<uint32_t>(((it / kStages) & 1) ^ 1));
                    ptx::mbar_arrive_expect_tx(&ring.full[s], row_bytes * static_cast<uint32_t>(a.n));
                    float* dst = stages + s * stage_floats;
                    for (int i = 0; i < a.n; ++i) ptx::bulk_g2s(dst + static_cast<size_t>(i) * tile, rows.p[i] + t * tile, row_bytes, &ring.full[s]);
                }
            }
        } else {
            int const ctid = threadIdx.x - 32;
            constexpr int kConsumers = 32 * kConsumerWarps;
            for (long long t = blockIdx.x; t < full_tiles; t += G, ++it) {
                int const s = static_cast<int>(it % kStages);
                ptx::mbar_wait(&ring.full[s], static_cast<uint32_t>((it / kStages) & 1));
                uint32_t const src = ptx::smem_u32(stages + s * stage_floats);
                uint32_t const row_bytes = static_cast<uint32_t>(tile) * 4u;
                for (int c = ctid * V; c < tile; c += kConsumers * V) {
                    long long const x = t * tile + c;
                    float res[V];
                    uint32_t const at = src + static_cast<uint32_t>(c) * 4u;
                    columns<NP, RULE, false, V>(a, nullptr, ~0ull, x, [&](int i, float (&dst)[V]) { lds_vec<V>(dst, at + static_cast<uint32_t>(i) * row_bytes); }, res);
                    for (int k = 0; k < g.n_out; ++k) vec_store<V>(g.out[parity][k] + x, res);
                }
                __syncwarp();
                if ((threadIdx.x & 31) == 0) ptx::mbar_arrive(&ring.empty[s]);
            }
            long long const done = full_tiles * tile;
            for (long long x = done + static_cast<long long>(blockIdx.x) * kConsumers + ctid; x < a.d; x += static_cast<long long>(G) * kConsumers) {
                float res[1];
                columns<NP, RULE, false, 1>(a, nullptr, ~0ull, x, [&](int i, float (&dst)[1]) { dst[0] = rows.p[i][x]; }, res);
                for (int k = 0; k < g.n_out; ++k) g.out[parity][k][x] = res[0];
            }
        }
        if (r + 1 == g.rounds) break;
        // ---- end of round r: local grid barrier, cross-GPU flag exchange by CTA 0, release ----
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned int const all_in = static_cast<unsigned int>(r) * (G + 1u) + G, go = static_cast<unsigned int>(r + 1) * (G + 1u);
            __threadfence();
            atomicAdd(g.counter, 1u);
            if (blockIdx.x == 0) {
                spin_until(g.counter, all_in, local_deadline);          // every CTA of this GPU has stored its part of round r
                uint32_t const value = seq0 + static_cast<uint32_t>(r) + 1u;
                __threadfence_system();
                if (g.my_flag != nullptr) ptx::st_release_sys(g.my_flag, value);
                unsigned long long const deadline = ptx::globaltimer_ns() + g.timeout_ns;
                for (int p = 0; p < g.n_peers; ++p) {
                    bool ready = false;
                    while (!(ready = static_cast<int32_t>(ptx::ld_acquire_sys(g.peer_flag[p]) - value) >= 0)) {
                        if (ptx::globaltimer_ns() > deadline) break;
                        __nanosleep(100);
                    }
                    if (!ready && g.late != nullptr) atomicAdd(g.late, 1u);
                }
                __threadfence_system();
                atomicAdd(g.counter, 1u);                               // release the grid into the next round
            }
            spin_until(g.counter, go, local_deadline);
            __threadfence();
        }
        __syncthreads();
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && g.rounds > 1) *g.seq = seq0 + static_cast<uint32_t>(g.rounds - 1);
}

template <int NP, int RULE, int V>
cudaError_t launch_gossip(GossipArgs& g, size_t smem, long long tiles, cudaStream_t stream) {
    auto kern = gossip_kernel<NP, RULE, V>;
    cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (err != cudaSuccess) return err;
    int per_sm = 0;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem);
    if (err != cudaSuccess) return err;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    if (per_sm > 2) per_sm = 2;
    long long grid = static_cast<long long>(device_sm_count()) * per_sm;       // all CTAs co-resident: the rounds need a grid barrier
    if (grid > tiles) grid = tiles;
    if (grid < 1) grid = 1;
    void* params[] = {&g};
    return cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kern), dim3(static_cast<unsigned>(grid)), dim3(kThreads), params, smem, stream);
}

template <int NP>
cudaError_t dispatch_gossip(GossipArgs& g, size_t smem, long long tiles, cudaStream_t stream) {
    switch (g.base.rule) {
        case kRuleMedianLower: return launch_gossip<NP, kRuleMedianLower, 4>(g, smem, tiles, stream);
        case kRuleTrimmedMean: return launch_gossip<NP, kRuleTrimmedMean, 4>(g, smem, tiles, stream);
        case kRuleAveragedMedian: return launch_gossip<NP, kRuleAveragedMedian, 4>(g, smem, tiles, stream);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace
}  // namespace gb

// rows0 / rows1: the n nodes' vectors of parity 0 / 1 (local or peer-mapped); out0[k] / out1[k]: buffers of the n_out local
// nodes that receive the result of a round reading parity 0 / 1 (i.e. local nodes' parity-1 / parity-0 vectors).
// `counter` (4 bytes, local) must be zero at launch; `seq` (local) carries the flag sequence across launches.
extern "C" GB_API int gb_gossip_rounds(float const* const* rows0, float const* const* rows1, int n, int64_t d, int rule, int rank, int upper,
                                       float* const* out0, float* const* out1, int n_out, int rounds, int first_parity, uint32_t* seq,
                                       uint32_t* my_flag, uint32_t* const* peer_flags, int n_peers, uint32_t* counter, uint64_t timeout_ns,
                                       uint32_t* late, void* stream) {
    using namespace gb;
    if (n < 1 || n > 16) return cuda_fail("gossip: need 1 <= n <= 16 nodes (larger groups take the launch-per-round path)");
    if (d < 1 || rounds < 1 || n_out < 1 || n_out > kMaxLocalOut || n_peers < 0 || n_peers > kMaxRows) return cuda_fail("gossip: bad arguments");
    if (rule != kRuleMedianLower && rule != kRuleTrimmedMean && rule != kRuleAveragedMedian)
        return cuda_fail("gossip: per-coordinate rules only (median, trimmed mean, averaged median)");
    if (rule == kRuleTrimmedMean && (rank < 0 || 2 * rank >= n)) return cuda_fail("gossip: trimmed mean needs 0 <= 2f < n");
    if (rule == kRuleAveragedMedian && (rank < 1 || rank > n)) return cuda_fail("gossip: averaged median needs 1 <= beta <= n");
    if (seq == nullptr || counter == nullptr || (n_peers > 0 && my_flag == nullptr)) return cuda_fail("gossip: missing flag / counter buffers");
    GossipArgs g{};
    bool aligned = true;
    for (int i = 0; i < n; ++i) {
        g.base.rows.p[i] = rows0[i];
        g.rows1.p[i] = rows1[i];
        aligned = aligned && ((reinterpret_cast<uintptr_t>(rows0[i]) | reinterpret_cast<uintptr_t>(rows1[i])) & 15u) == 0;
    }
    for (int k = 0; k < n_out; ++k) {
        g.out[0][k] = out0[k];
        g.out[1][k] = out1[k];
        aligned = aligned && ((reinterpret_cast<uintptr_t>(out0[k]) | reinterpret_cast<uintptr_t>(out1[k])) & 15u) == 0;
    }
    g.base.n = n; g.base.rule = rule; g.base.rank = rank; g.base.upper = upper; g.base.theta = 0; g.base.d = d;
    g.base.tile = pick_tile(n);
    g.base.use_tma = aligned ? 1 : 0;
    g.n_out = n_out; g.rounds = rounds; g.first = first_parity & 1; g.seq = seq; g.my_flag = my_flag; g.n_peers = n_peers;
    for (int p = 0; p < n_peers; ++p) g.peer_flag[p] = peer_flags[p];
    g.counter = counter; g.timeout_ns = timeout_ns; g.late = late;
    size_t const smem = sizeof(float) * static_cast<size_t>(kStages) * n * g.base.tile;
    long long const tiles = (d + g.base.tile - 1) / g.base.tile;
    cudaStream_t const st = static_cast<cudaStream_t>(stream);
    cudaError_t err;
    int const np = n <= 4 ? 4 : (n + 3) / 4 * 4;
    switch (np) {
        case 4: err = dispatch_gossip<4>(g, smem, tiles, st); break;
        case 8: err = dispatch_gossip<8>(g, smem, tiles, st); break;
        case 12: err = dispatch_gossip<12>(g, smem, tiles, st); break;
        default: err = dispatch_gossip<16>(g, smem, tiles, st); break;
    }
    return cuda_check(err, "gossip launch");
}
