// On-device selection logic shared by the distance kernels (SIMT pairwise and tcgen05 Gram):
// given the final n×n squared-distance matrix in global memory, emit the averaging weights W.
// Executed by the LAST CTA of the producing kernel; `Args` needs the fields n, dist, W, info, sel.
#pragma once

#include <cfloat>
#include <cmath>

#include "../cpu/combinations.hpp"
#include "gar_common.cuh"

namespace gb {

// Host-side validation shared by every entry point that runs the selection (gb_pairwise_select, gb_pairwise_select_noalloc,
// gb_select_only, gb_gram_select_tc): out-of-range (n, f, m, neighbors) would index the score order and the weight matrix
// out of bounds on the device. Returns nullptr when the parameters are valid, else what is wrong.
inline char const* selection_error(int mode, int n, int f, int m, int neighbors) {
    switch (mode) {
        case kSelectNone: return nullptr;
        case kSelectKrum:
            return (f < 1 || n < 2 * f + 3 || m < 1 || m > n - f - 2 || neighbors < 1 || neighbors > n - 1) ? "invalid Krum parameters" : nullptr;
        case kSelectBulyanPython:
        case kSelectBulyanNative:
            return (f < 1 || n < 4 * f + 3 || m < 1 || m > n - f - 2) ? "invalid Bulyan parameters" : nullptr;
        case kSelectMda: {
            if (f < 1 || n < 2 * f + 1) return "invalid MDA parameters";
            bool over = false;
            unsigned long long const total = binomial(static_cast<unsigned>(n), static_cast<unsigned>(n - f), &over);
            return (over || total > (1ull << 26)) ? "too many MDA subsets for the in-kernel search (limit 2^26)" : nullptr;
        }
        case kSelectClosest: return (n < 2 || m < 1 || m > n - 1) ? "invalid closest-to-last parameters" : nullptr;
        default: return "unknown selection mode";
    }
}

// ------------------------------------------------------------------------ selection ---

struct SelectScratch {      // lives in dynamic shared memory of the last CTA (ring is idle by then)
    double D[kMaxRows * kMaxRows];       // distances used for scoring (sqrt applied when !squared)
    double kept[kMaxRows * kMaxRows];    // Bulyan native: distances kept after pruning
    double score[kMaxRows];
    int order[kMaxRows];
    double best_diam[kThreads];
    unsigned long long best_pos[kThreads];
};

// lexicographic (non-finite last, score, index) comparison
__device__ __forceinline__ bool before(double sa, int a, double sb, int b) {
    bool const fa = isfinite(sa), fb = isfinite(sb);
    if (fa != fb) return fa;
    if (fa && sa != sb) return sa < sb;
    return a < b;
}

// order[] = indices of rows [0, cnt) by increasing score (stable); executed by one warp.
__device__ inline void rank_rows(SelectScratch& S, int cnt) {
    int const lane = threadIdx.x & 31;
    for (int i = lane; i < cnt; i += 32) {
        int rank = 0;
        for (int j = 0; j < cnt; ++j) rank += before(S.score[j], j, S.score[i], i) ? 1 : 0;
        S.order[rank] = i;
    }
    __syncwarp();
}

// score_i = Σ of the k smallest D[i][j], j != i, j < cnt (ascending summation order).
// `keep_out` (optional) receives the kept distances (Bulyan native pruning).
__device__ inline void neighbour_scores(SelectScratch& S, int n, int cnt, int k, bool prune) {
    int const lane = threadIdx.x & 31;
    for (int i = lane; i < cnt; i += 32) {
        double vals[kMaxRows];
        int ids[kMaxRows];
        int len = 0;
        for (int j = 0; j < cnt; ++j) {
            if (j == i) continue;
            double const v = S.D[i * n + j];
            int pos = len++;
            while (pos > 0 && (vals[pos - 1] > v)) {  // insertion sort, stable in j
                vals[pos] = vals[pos - 1];
                ids[pos] = ids[pos - 1];
                --pos;
            }
            vals[pos] = v;
            ids[pos] = j;
        }
        double s = 0.0;
        for (int t = 0; t < k && t < len; ++t) s += vals[t];
        S.score[i] = s;
        if (prune) {
            for (int j = 0; j < cnt; ++j) S.kept[i * n + j] = 0.0;
            for (int t = 0; t < k && t < len; ++t) S.kept[i * n + ids[t]] = vals[t];
        }
    }
    __syncwarp();
}

// Runs in the last CTA. D2 (global, float) holds the final squared distances.
template <class Args>
__device__ void run_selection(Args const& a, SelectScratch& S) {
    int const n = a.n;
    SelectArgs const sel = a.sel;
    int const tid = threadIdx.x;
    if (sel.mode == kSelectNone) return;

    for (int e = tid; e < n * n; e += blockDim.x) {
        auto const v = a.dist[e];      // float (fp32 kernels) or double (fp64 kernels)
        double dv = sel.squared ? static_cast<double>(v) : sqrt(static_cast<double>(v));
        // native MDA counts a non-finite distance as the largest finite one instead of skipping the subset
        if (sel.mode == kSelectMda && sel.squared && !isfinite(dv) && (e / n) != (e % n)) dv = DBL_MAX;
        S.D[e] = dv;
    }
    int const theta = (sel.mode == kSelectBulyanPython || sel.mode == kSelectBulyanNative) ? n - 2 * sel.f - 2 : 1;
    for (int e = tid; e < theta * n; e += blockDim.x) a.W[e] = 0.0f;
    __syncthreads();

    if (sel.mode == kSelectMda) {
        // every thread scans a contiguous range of subset ranks; strict '<' keeps the first minimum
        unsigned const k = static_cast<unsigned>(n - sel.f);
        unsigned long long const total = binomial(static_cast<unsigned>(n), k);
        unsigned long long const lo = total * tid / blockDim.x, hi = total * (tid + 1ull) / blockDim.x;
        double best = INFINITY;
        unsigned long long best_pos = ~0ull;
        if (lo < hi) {
            int idx[kMaxRows];
            combination_unrank(static_cast<unsigned>(n), k, lo, idx);
            for (unsigned long long pos = lo; pos < hi; ++pos) {
                double diam = 0.0;
                for (unsigned x = 0; x + 1 < k; ++x)
                    for (unsigned y = x + 1; y < k; ++y) diam = fmax(diam, S.D[idx[x] * n + idx[y]]);
                if (diam < best) {
                    best = diam;
                    best_pos = pos;
                }
                combination_next(static_cast<unsigned>(n), k, idx);
            }
        }
        S.best_diam[tid] = best;
        S.best_pos[tid] = best_pos;
        __syncthreads();
        if (tid == 0) {
            double win = INFINITY;
            unsigned long long win_pos = ~0ull;
            for (unsigned t = 0; t < blockDim.x; ++t)  // ranges are in rank order: first strict minimum wins
                if (S.best_diam[t] < win) {
                    win = S.best_diam[t];
                    win_pos = S.best_pos[t];
                }
            a.info[0] = 1;
            if (win_pos != ~0ull) {
                int idx[kMaxRows];
                combination_unrank(static_cast<unsigned>(n), k, win_pos, idx);
                for (unsigned x = 0; x < k; ++x) {
                    a.W[idx[x]] = 1.0f / static_cast<float>(k);
                    a.info[1 + x] = idx[x];
                }
            }
        }
        return;
    }

    if (tid >= 32) return;  // the remaining modes need one warp
    int const lane = tid;

    if (sel.mode == kSelectClosest) {
        int const cnt = n - 1;  // the last row is the reference vector
        for (int i = lane; i < cnt; i += 32) S.score[i] = S.D[i * n + (n - 1)];
        __syncwarp();
        rank_rows(S, cnt);
        for (int r = lane; r < sel.m; r += 32) {
            a.W[S.order[r]] = 1.0f / static_cast<float>(sel.m);
            a.info[1 + r] = S.order[r];
        }
        if (lane == 0) a.info[0] = 1;
        return;
    }

    if (sel.mode == kSelectKrum) {
        neighbour_scores(S, n, n, sel.neighbors, false);
        rank_rows(S, n);
        for (int r = lane; r < sel.m; r += 32) {
            a.W[S.order[r]] = 1.0f / static_cast<float>(sel.m);
            a.info[1 + r] = S.order[r];
        }
        if (lane == 0) a.info[0] = 1;
        return;
    }

    int const m_max = n - sel.f - 2;
    if (sel.mode == kSelectBulyanPython) {
        neighbour_scores(S, n, n, sel.m, false);  // python Bulyan sums the m nearest
        rank_rows(S, n);                          // scores are never updated (reference quirk Q2)
        for (int r = 0; r < theta; ++r) {
            int const take = min(sel.m, m_max - r);
            for (int t = lane; t < take; t += 32) a.W[r * n + S.order[r + t]] = 1.0f / static_cast<float>(take);
        }
        for (int t = lane; t < min(sel.m, m_max); t += 32) a.info[1 + t] = S.order[t];
        if (lane == 0) a.info[0] = theta;
        return;
    }

    // kSelectBulyanNative: real score updates with pruned distances
    neighbour_scores(S, n, n, m_max, true);
    for (int r = 0; r < theta; ++r) {
        rank_rows(S, n);
        int const take = sel.m - r;
        for (int t = lane; t < take; t += 32) a.W[r * n + S.order[t]] = 1.0f / static_cast<float>(take);
        if (r == 0)
            for (int t = lane; t < take; t += 32) a.info[1 + t] = S.order[t];
        __syncwarp();
        if (r + 1 >= theta) break;
        int const gone = S.order[0];
        __syncwarp();
        for (int i = lane; i < n; i += 32) {
            if (i == gone) S.score[i] = INFINITY;
            else if (isfinite(S.score[i])) S.score[i] -= S.kept[i * n + gone];
        }
        __syncwarp();
    }
    if (lane == 0) a.info[0] = theta;
}


}  // namespace gb
