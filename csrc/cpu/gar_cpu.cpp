// C++ host implementations of the "native" aggregation rules (float and double).
//
// Capability parity with the reference's CPU natives
//   krum    /root/reference/pytorch_impl/libs/native/py_krum/krum.cpp:50-118
//   bulyan  /root/reference/pytorch_impl/libs/native/py_bulyan/bulyan.cpp:53-193
//   median  /root/reference/pytorch_impl/libs/native/py_median/median.cpp:42-77
//   brute   /root/reference/pytorch_impl/libs/native/py_brute/brute.cpp:47-113
// and their helpers (operations.hpp:41-79). Exposed through a plain C ABI (no torch headers);
// garfield_b200/native binds them with ctypes.
//
// Design: all pairwise squared distances come from ONE blocked sweep over the n×d data (each
// cache-resident coordinate block updates all n(n-1)/2 accumulators) instead of n(n-1)/2
// separate passes; selection logic works on the n×n matrix; the output pass is blocked the
// same way. Partial sums are kept per block in double and combined in block order, so results
// do not depend on the thread count.

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numeric>
#include <string>
#include <vector>

#include "strided_view.hpp"
#include "combinations.hpp"
#include "garfield_b200.h"
#include "threadpool.hpp"

namespace {

thread_local std::string tl_error;

int fail(char const* what) {
    tl_error = what;
    return 1;
}

constexpr int kNetworkRows = 32;   // rows handled by the vectorised sorting-network median
constexpr int kLanes = 64;         // coordinates sorted together (one cache line of fp32 per row)
constexpr std::size_t kBlock = 4096;  // coordinates per work item (n rows × 16 KiB fp32 stays in L1/L2)

template <class T>
bool finite(T v) { return std::isfinite(v); }

// Strict weak order: finite ascending, every non-finite value after all finite ones.
template <class T>
bool finite_less(T x, T y) {
    if (!finite(x)) return false;
    if (!finite(y)) return true;
    return x < y;
}

// dist[i*n+j] = Σ_x (g_i[x] − g_j[x])², symmetric, diagonal untouched (caller fills).
template <class T>
void pairwise_sqdist(T const* const* in, int n, std::int64_t d, std::vector<double>& dist) {
    std::size_t const pairs = static_cast<std::size_t>(n) * (n - 1) / 2;
    std::size_t const blocks = (static_cast<std::size_t>(d) + kBlock - 1) / kBlock;
    std::vector<double> partial(blocks * pairs, 0.0);
    gb::parallel_for(0, blocks, [&](std::size_t lo, std::size_t hi) {
        for (std::size_t b = lo; b < hi; ++b) {
            std::size_t const x0 = b * kBlock;
            std::size_t const x1 = std::min<std::size_t>(static_cast<std::size_t>(d), x0 + kBlock);
            double* acc = partial.data() + b * pairs;
            std::size_t p = 0;
            for (int i = 0; i < n - 1; ++i) {
                T const* gi = in[i];
                for (int j = i + 1; j < n; ++j, ++p) {
                    T const* gj = in[j];
                    T lane[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // 8 lanes of input-precision partial sums (a SIMD-shaped loop)
                    std::size_t x = x0;
                    for (; x + 8 <= x1; x += 8)
                        for (int l = 0; l < 8; ++l) {
                            T const a = gi[x + l] - gj[x + l];
                            lane[l] += a * a;
                        }
                    for (; x < x1; ++x) {
                        T const a = gi[x] - gj[x];
                        lane[0] += a * a;
                    }
                    double total = 0.0;
                    for (int l = 0; l < 8; ++l) total += static_cast<double>(lane[l]);
                    acc[p] = total;
                }
            }
        }
    });
    dist.assign(static_cast<std::size_t>(n) * n, 0.0);
    std::size_t p = 0;
    for (int i = 0; i < n - 1; ++i)
        for (int j = i + 1; j < n; ++j, ++p) {
            double total = 0.0;
            for (std::size_t b = 0; b < blocks; ++b) total += partial[b * pairs + p];
            dist[static_cast<std::size_t>(i) * n + j] = dist[static_cast<std::size_t>(j) * n + i] = total;
        }
}

// out[x] = Σ_k w[k] · in[sel[k]][x]
template <class T>
void weighted_sum(T const* const* in, std::int64_t d, int const* sel, T const* w, int count, T* out) {
    std::size_t const blocks = (static_cast<std::size_t>(d) + kBlock - 1) / kBlock;
    gb::parallel_for(0, blocks, [&](std::size_t lo, std::size_t hi) {
        for (std::size_t b = lo; b < hi; ++b) {
            std::size_t const x0 = b * kBlock;
            std::size_t const x1 = std::min<std::size_t>(static_cast<std::size_t>(d), x0 + kBlock);
            T const* first = in[sel[0]];
            for (std::size_t x = x0; x < x1; ++x) out[x] = first[x] * w[0];
            for (int k = 1; k < count; ++k) {
                T const* g = in[sel[k]];
                T const wk = w[k];
                for (std::size_t x = x0; x < x1; ++x) out[x] += g[x] * wk;
            }
        }
    });
}

template <class T>
void selection_average(T const* const* in, std::int64_t d, int const* sel, int count, T* out) {
    // sum then divide (matches the reference's accumulate-then-divide rounding)
    std::size_t const blocks = (static_cast<std::size_t>(d) + kBlock - 1) / kBlock;
    T const denom = static_cast<T>(count);
    gb::parallel_for(0, blocks, [&](std::size_t lo, std::size_t hi) {
        for (std::size_t b = lo; b < hi; ++b) {
            std::size_t const x0 = b * kBlock;
            std::size_t const x1 = std::min<std::size_t>(static_cast<std::size_t>(d), x0 + kBlock);
            T const* first = in[sel[0]];
            for (std::size_t x = x0; x < x1; ++x) out[x] = first[x];
            for (int k = 1; k < count; ++k) {
                T const* g = in[sel[k]];
                for (std::size_t x = x0; x < x1; ++x) out[x] += g[x];
            }
            for (std::size_t x = x0; x < x1; ++x) out[x] /= denom;
        }
    });
}

// score_i = Σ of the `keep` smallest finite-ordered distances touching i; order[i] lists the
// other gradients by increasing distance (ties by index).
void neighbour_scores(std::vector<double> const& dist, int n, int keep, std::vector<double>& scores,
                      std::vector<std::vector<int>>& order) {
    scores.assign(n, 0.0);
    order.assign(n, {});
    for (int i = 0; i < n; ++i) {
        auto& row = order[i];
        for (int j = 0; j < n; ++j)
            if (j != i) row.push_back(j);
        double const* di = dist.data() + static_cast<std::size_t>(i) * n;
        std::stable_sort(row.begin(), row.end(), [&](int a, int b) { return finite_less(di[a], di[b]); });
        double s = 0.0;
        for (int k = 0; k < keep; ++k) s += di[row[k]];
        scores[i] = s;
    }
}

std::vector<int> by_score(std::vector<double> const& scores) {
    std::vector<int> idx(scores.size());
    std::iota(idx.begin(), idx.end(), 0);
    std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return finite_less(scores[a], scores[b]); });
    return idx;
}

template <class T>
int krum(T const* const* in, int n, std::int64_t d, int f, int m, T* out) {
    if (n < 1 || d < 1) return fail("krum: expected at least one non-empty gradient");
    if (f < 1 || n < 2 * f + 3) return fail("krum: need f >= 1 and n >= 2f+3");
    if (m < 1 || m > n - f - 2) return fail("krum: need 1 <= m <= n-f-2");
    std::vector<double> dist;
    pairwise_sqdist(in, n, d, dist);
    std::vector<double> scores;
    std::vector<std::vector<int>> order;
    neighbour_scores(dist, n, n - f - 2, scores, order);
    std::vector<int> best = by_score(scores);
    selection_average(in, d, best.data(), m, out);
    return 0;
}

template <class T>
int bulyan(T const* const* in, int n, std::int64_t d, int f, int m, T* out) {
    if (n < 1 || d < 1) return fail("bulyan: expected at least one non-empty gradient");
    if (f < 1 || n < 4 * f + 3) return fail("bulyan: need f >= 1 and n >= 4f+3");
    if (m < 1 || m > n - f - 2) return fail("bulyan: need 1 <= m <= n-f-2");
    int const keep = n - f - 2;
    int const theta = n - 2 * f - 2;
    int const beta = theta - 2 * f;
    std::vector<double> dist;
    pairwise_sqdist(in, n, d, dist);
    std::vector<double> scores;
    std::vector<std::vector<int>> order;
    neighbour_scores(dist, n, keep, scores, order);
    // kept[i][j] = dist(i,j) when j is one of i's `keep` nearest, else 0 ("pruned")
    std::vector<double> kept(static_cast<std::size_t>(n) * n, 0.0);
    for (int i = 0; i < n; ++i)
        for (int k = 0; k < keep; ++k) {
            int const j = order[i][k];
            kept[static_cast<std::size_t>(i) * n + j] = dist[static_cast<std::size_t>(i) * n + j];
        }
    // θ selection rounds → θ×n weights
    std::vector<T> weights(static_cast<std::size_t>(theta) * n, T(0));
    for (int k = 0; k < theta; ++k) {
        std::vector<int> rank = by_score(scores);
        int const take = m - k;
        for (int r = 0; r < take; ++r) weights[static_cast<std::size_t>(k) * n + rank[r]] = T(1) / static_cast<T>(take);
        if (k + 1 >= theta) break;
        int const gone = rank[0];
        scores[gone] = std::numeric_limits<double>::infinity();
        for (int i = 0; i < n; ++i)
            if (i != gone && std::isfinite(scores[i])) scores[i] -= kept[static_cast<std::size_t>(i) * n + gone];
    }
    // phase 2, fused per coordinate block: θ round-averages in a small local tile, then the
    // mean of the β values closest to their (upper) median
    std::size_t const blocks = (static_cast<std::size_t>(d) + kBlock - 1) / kBlock;
    gb::parallel_for(0, blocks, [&](std::size_t lo, std::size_t hi) {
        std::vector<T> tile(static_cast<std::size_t>(theta) * kBlock);
        std::vector<T> keys(theta <= kNetworkRows ? static_cast<std::size_t>(theta) * kBlock : 0);
        for (std::size_t b = lo; b < hi; ++b) {
            std::size_t const x0 = b * kBlock;
            std::size_t const len = std::min<std::size_t>(static_cast<std::size_t>(d), x0 + kBlock) - x0;
            std::fill(tile.begin(), tile.end(), T(0));
            for (int k = 0; k < theta; ++k) {
                T* row = tile.data() + static_cast<std::size_t>(k) * kBlock;
                for (int i = 0; i < n; ++i) {
                    T const w = weights[static_cast<std::size_t>(k) * n + i];
                    if (w == T(0)) continue;
                    T const* g = in[i] + x0;
                    for (std::size_t x = 0; x < len; ++x) row[x] += g[x] * w;
                }
            }
            if (theta <= kNetworkRows) {
                // Small θ: two odd-even transposition networks over whole tile rows (contiguous, branch-free min/max and
                // selects → SIMD): sort the θ mixes per coordinate, take the median row, then sort (|v − med|, v) pairs by
                // key and average the first β values.
                for (int round = 0; round < theta; ++round)
                    for (int k = round & 1; k + 1 < theta; k += 2) {
                        T* __restrict__ a = tile.data() + static_cast<std::size_t>(k) * kBlock;
                        T* __restrict__ b = a + kBlock;
                        for (std::size_t x = 0; x < len; ++x) {
                            T const u = a[x], v = b[x];
                            T const lo2 = v < u ? v : u;
                            T const hi2 = v < u ? u : v;
                            a[x] = lo2;
                            b[x] = hi2;
                        }
                    }
                T const* __restrict__ med = tile.data() + static_cast<std::size_t>(theta / 2) * kBlock;
                for (int k = 0; k < theta; ++k) {
                    T const* __restrict__ v = tile.data() + static_cast<std::size_t>(k) * kBlock;
                    T* __restrict__ key = keys.data() + static_cast<std::size_t>(k) * kBlock;
                    for (std::size_t x = 0; x < len; ++x) key[x] = std::fabs(v[x] - med[x]);
                }
                for (int round = 0; round < theta; ++round)
                    for (int k = round & 1; k + 1 < theta; k += 2) {
                        T* __restrict__ ka = keys.data() + static_cast<std::size_t>(k) * kBlock;
                        T* __restrict__ kb = ka + kBlock;
                        T* __restrict__ va = tile.data() + static_cast<std::size_t>(k) * kBlock;
                        T* __restrict__ vb = va + kBlock;
                        for (std::size_t x = 0; x < len; ++x) {
                            bool const swap = kb[x] < ka[x];
                            T const k0 = swap ? kb[x] : ka[x], k1 = swap ? ka[x] : kb[x];
                            T const v0 = swap ? vb[x] : va[x], v1 = swap ? va[x] : vb[x];
                            ka[x] = k0; kb[x] = k1;
                            va[x] = v0; vb[x] = v1;
                        }
                    }
                T* __restrict__ dst = out + x0;
                T const* __restrict__ first = tile.data();
                for (std::size_t x = 0; x < len; ++x) dst[x] = first[x];
                for (int k = 1; k < beta; ++k) {
                    T const* __restrict__ v = tile.data() + static_cast<std::size_t>(k) * kBlock;
                    for (std::size_t x = 0; x < len; ++x) dst[x] += v[x];
                }
                T const denom = static_cast<T>(beta);
                for (std::size_t x = 0; x < len; ++x) dst[x] /= denom;
                continue;
            }
            // larger θ: order statistics run in place along the (strided) θ axis of the tile — no per-coordinate gather
            gb::StridedView<T> const rounds(tile.data(), {static_cast<std::size_t>(theta), kBlock});
            for (std::size_t x = 0; x < len; ++x) {
                auto const column = rounds.at(1, x).line(0);
                std::nth_element(column.begin(), column.begin() + theta / 2, column.end());
                T const med = column.begin()[theta / 2];
                std::nth_element(column.begin(), column.begin() + (beta - 1), column.end(), [&](T a, T c) {
                    return std::fabs(a - med) < std::fabs(c - med);
                });
                T sum = 0;
                for (int k = 0; k < beta; ++k) sum += column.begin()[k];
                out[x0 + x] = sum / static_cast<T>(beta);
            }
        }
    });
    return 0;
}

template <class T>
int median(T const* const* in, int n, std::int64_t d, T* out) {
    if (n < 1 || d < 1) return fail("median: expected at least one non-empty gradient");
    if (n <= kNetworkRows) {
        // Small n (the usual case): sort kLanes coordinates at once with an odd-even transposition network over the n rows.
        // Non-finite values become +inf (they sink to the end), the finite count picks element count/2 per coordinate.
        // Every inner loop runs over a contiguous lane array with branch-free min/max, which the compiler vectorises.
        std::size_t const groups = (static_cast<std::size_t>(d) + kLanes - 1) / kLanes;
        gb::parallel_for(0, groups, [&](std::size_t lo, std::size_t hi) {
            alignas(64) T vals[kNetworkRows][kLanes];
            alignas(64) int count[kLanes];
            T const inf = std::numeric_limits<T>::infinity();
            for (std::size_t g = lo; g < hi; ++g) {
                std::size_t const x0 = g * kLanes;
                int const width = static_cast<int>(std::min<std::size_t>(kLanes, static_cast<std::size_t>(d) - x0));
                for (int l = 0; l < kLanes; ++l) count[l] = 0;
                for (int i = 0; i < n; ++i) {
                    T const* row = in[i] + x0;
                    for (int l = 0; l < width; ++l) {
                        T const v = row[l];
                        bool const ok = v - v == T(0);                     // false for NaN and ±inf
                        vals[i][l] = ok ? v : inf;
                        count[l] += ok ? 1 : 0;
                    }
                    for (int l = width; l < kLanes; ++l) vals[i][l] = inf;
                }
                for (int round = 0; round < n; ++round)
                    for (int i = round & 1; i + 1 < n; i += 2) {
                        T* __restrict__ a = vals[i];
                        T* __restrict__ b = vals[i + 1];
                        for (int l = 0; l < kLanes; ++l) {
                            T const u = a[l], v = b[l];
                            T const lo = v < u ? v : u;
                            T const hi = v < u ? u : v;
                            a[l] = lo;
                            b[l] = hi;
                        }
                    }
                for (int l = 0; l < width; ++l) out[x0 + l] = count[l] == 0 ? T(0) : vals[count[l] / 2][l];
            }
        }, 64);
        return 0;
    }
    gb::parallel_for(0, static_cast<std::size_t>(d), [&](std::size_t lo, std::size_t hi) {
        std::vector<T> column(n);
        for (std::size_t x = lo; x < hi; ++x) {
            int len = 0;
            for (int i = 0; i < n; ++i) {
                T const v = in[i][x];
                if (finite(v)) column[len++] = v;
            }
            if (len == 0) {
                out[x] = 0;
                continue;
            }
            std::nth_element(column.begin(), column.begin() + len / 2, column.begin() + len);
            out[x] = column[len / 2];
        }
    }, 8192);
    return 0;
}

// Python-semantics coordinate-wise rules on the host, n <= kNetworkRows, with the same SIMD-shaped network as `median`:
//   rule 0  lower median, NaN-propagating (torch.median)            — aggregators/median.py:31-39
//   rule 1  f-trimmed mean (arg = f)                                — drop the f smallest / f largest, average the rest
//   rule 2  mean of the arg = beta values closest to the lower median, ties: lower value first
// Returns 0, or 2 when a column held a value the fast path does not reproduce bit-for-bit semantics for (NaN for rule 1,
// any non-finite for rule 2): the caller then recomputes with the reference formulation. 3 = n too large.
template <class T>
int coordwise_host(T const* const* in, int n, std::int64_t d, int rule, int arg, T* out) {
    if (n < 1 || d < 1) return fail("coordwise: expected at least one non-empty gradient");
    if (rule < 0 || rule > 2) return fail("coordwise: unknown rule");
    if (rule == 1 && (arg < 0 || n < 2 * arg + 1)) return fail("coordwise: need 0 <= f and n >= 2f+1");
    if (rule == 2 && (arg < 1 || arg > n)) return fail("coordwise: need 1 <= beta <= n");
    if (n > kNetworkRows) return 3;
    std::atomic<int> odd{0};
    std::size_t const groups = (static_cast<std::size_t>(d) + kLanes - 1) / kLanes;
    gb::parallel_for(0, groups, [&](std::size_t lo, std::size_t hi) {
        alignas(64) T vals[kNetworkRows][kLanes];
        alignas(64) T keys[kNetworkRows][kLanes];
        alignas(64) int nans[kLanes];
        T const inf = std::numeric_limits<T>::infinity();
        T const nan = std::numeric_limits<T>::quiet_NaN();
        int const mid = (n - 1) / 2;
        for (std::size_t g = lo; g < hi; ++g) {
            std::size_t const x0 = g * kLanes;
            int const width = static_cast<int>(std::min<std::size_t>(kLanes, static_cast<std::size_t>(d) - x0));
            int strange = 0;
            for (int l = 0; l < kLanes; ++l) nans[l] = 0;
            for (int i = 0; i < n; ++i) {
                T const* row = in[i] + x0;
                for (int l = 0; l < width; ++l) {
                    T const v = row[l];
                    bool const is_nan = v != v;
                    vals[i][l] = is_nan ? inf : v;             // NaN sorts last, like torch.sort
                    nans[l] += is_nan ? 1 : 0;
                    strange |= (v - v == T(0)) ? 0 : 1;         // NaN or ±inf
                }
                for (int l = width; l < kLanes; ++l) vals[i][l] = T(0);
            }
            for (int round = 0; round < n; ++round)
                for (int i = round & 1; i + 1 < n; i += 2) {
                    T* __restrict__ a = vals[i];
                    T* __restrict__ b = vals[i + 1];
                    for (int l = 0; l < kLanes; ++l) {
                        T const u = a[l], v = b[l];
                        T const small = v < u ? v : u;
                        T const large = v < u ? u : v;
                        a[l] = small;
                        b[l] = large;
                    }
                }
            if (rule == 0) {
                for (int l = 0; l < width; ++l) out[x0 + l] = nans[l] > 0 ? nan : vals[mid][l];
            } else if (rule == 1) {
                int any_nan = 0;
                for (int l = 0; l < width; ++l) any_nan |= nans[l];
                if (any_nan) odd.store(1, std::memory_order_relaxed);
                T* __restrict__ dst = out + x0;
                T const denom = static_cast<T>(n - 2 * arg);
                for (int l = 0; l < width; ++l) dst[l] = vals[arg][l];
                for (int k = arg + 1; k < n - arg; ++k)
                    for (int l = 0; l < width; ++l) dst[l] += vals[k][l];
                for (int l = 0; l < width; ++l) dst[l] /= denom;
            } else {
                if (strange) odd.store(1, std::memory_order_relaxed);
                for (int k = 0; k < n; ++k)
                    for (int l = 0; l < kLanes; ++l) keys[k][l] = std::fabs(vals[k][l] - vals[mid][l]);
                // stable (strict-comparison, adjacent-exchange) sort by distance to the median: equal keys keep the
                // ascending value order of the first sort, i.e. "ties: lower value first"
                for (int round = 0; round < n; ++round)
                    for (int i = round & 1; i + 1 < n; i += 2) {
                        T* __restrict__ ka = keys[i];
                        T* __restrict__ kb = keys[i + 1];
                        T* __restrict__ va = vals[i];
                        T* __restrict__ vb = vals[i + 1];
                        for (int l = 0; l < kLanes; ++l) {
                            bool const swap = kb[l] < ka[l];
                            T const k0 = swap ? kb[l] : ka[l], k1 = swap ? ka[l] : kb[l];
                            T const v0 = swap ? vb[l] : va[l], v1 = swap ? va[l] : vb[l];
                            ka[l] = k0; kb[l] = k1;
                            va[l] = v0; vb[l] = v1;
                        }
                    }
                T* __restrict__ dst = out + x0;
                T const denom = static_cast<T>(arg);
                for (int l = 0; l < width; ++l) dst[l] = vals[0][l];
                for (int k = 1; k < arg; ++k)
                    for (int l = 0; l < width; ++l) dst[l] += vals[k][l];
                for (int l = 0; l < width; ++l) dst[l] /= denom;
            }
        }
    }, 64);
    return odd.load() ? 2 : 0;
}

template <class T>
int brute(T const* const* in, int n, std::int64_t d, int f, T* out) {
    if (n < 1 || d < 1) return fail("brute: expected at least one non-empty gradient");
    if (f < 1 || n < 2 * f + 1) return fail("brute: need f >= 1 and n >= 2f+1");
    int const k = n - f;
    bool overflow = false;
    std::uint64_t const total = gb::binomial(static_cast<unsigned>(n), static_cast<unsigned>(k), &overflow);
    if (overflow || total == 0) return fail("brute: too many combinations");
    std::vector<double> dist;
    pairwise_sqdist(in, n, d, dist);
    for (auto& v : dist)
        if (!std::isfinite(v)) v = std::numeric_limits<double>::max();
    // every chunk keeps its first strict minimum; chunks are merged in rank order so the
    // lexicographically first minimal subset wins regardless of the thread count
    std::size_t const chunks = std::min<std::uint64_t>(total, 4ull * gb::ThreadPool::global().size());
    std::vector<double> best_diam(chunks, std::numeric_limits<double>::infinity());
    std::vector<std::uint64_t> best_pos(chunks, 0);
    gb::parallel_for(0, chunks, [&](std::size_t lo, std::size_t hi) {
        std::vector<int> idx(k);
        for (std::size_t c = lo; c < hi; ++c) {
            std::uint64_t const start = total * c / chunks, stop = total * (c + 1) / chunks;
            if (start >= stop) continue;
            gb::combination_unrank(static_cast<unsigned>(n), static_cast<unsigned>(k), start, idx.data());
            for (std::uint64_t pos = start; pos < stop; ++pos) {
                double diam = 0.0;
                for (int a = 0; a < k - 1; ++a) {
                    double const* row = dist.data() + static_cast<std::size_t>(idx[a]) * n;
                    for (int b = a + 1; b < k; ++b) diam = std::max(diam, row[idx[b]]);
                }
                if (diam < best_diam[c]) {
                    best_diam[c] = diam;
                    best_pos[c] = pos;
                }
                gb::combination_next(static_cast<unsigned>(n), static_cast<unsigned>(k), idx.data());
            }
        }
    }, 1);
    std::size_t win = 0;
    for (std::size_t c = 1; c < chunks; ++c)
        if (best_diam[c] < best_diam[win]) win = c;
    std::vector<int> idx(k);
    gb::combination_unrank(static_cast<unsigned>(n), static_cast<unsigned>(k), best_pos[win], idx.data());
    selection_average(in, d, idx.data(), k, out);
    return 0;
}

template <class T>
int sqdist_matrix(T const* const* in, int n, std::int64_t d, double* out) {
    if (n < 1 || d < 1) return fail("pairwise_sqdist: expected at least one non-empty gradient");
    std::vector<double> dist;
    if (n > 1) pairwise_sqdist(in, n, d, dist); else dist.assign(1, 0.0);
    for (int i = 0; i < n; ++i) dist[static_cast<std::size_t>(i) * n + i] = std::numeric_limits<double>::infinity();
    std::memcpy(out, dist.data(), sizeof(double) * static_cast<std::size_t>(n) * n);
    return 0;
}

template <class F>
int guarded(F&& body) {
    try {
        return body();
    } catch (std::exception const& err) {
        tl_error = err.what();
        return 2;
    } catch (...) {
        tl_error = "unknown C++ exception";
        return 2;
    }
}

}  // namespace

extern "C" {

GB_API char const* gb_cpu_last_error(void) { return tl_error.c_str(); }
GB_API int gb_cpu_num_threads(void) { return static_cast<int>(gb::ThreadPool::global().size()); }

GB_API int gb_cpu_krum_f32(float const* const* in, int n, int64_t d, int f, int m, float* out) { return guarded([&] { return krum(in, n, d, f, m, out); }); }
GB_API int gb_cpu_krum_f64(double const* const* in, int n, int64_t d, int f, int m, double* out) { return guarded([&] { return krum(in, n, d, f, m, out); }); }
GB_API int gb_cpu_bulyan_f32(float const* const* in, int n, int64_t d, int f, int m, float* out) { return guarded([&] { return bulyan(in, n, d, f, m, out); }); }
GB_API int gb_cpu_bulyan_f64(double const* const* in, int n, int64_t d, int f, int m, double* out) { return guarded([&] { return bulyan(in, n, d, f, m, out); }); }
GB_API int gb_cpu_median_f32(float const* const* in, int n, int64_t d, float* out) { return guarded([&] { return median(in, n, d, out); }); }
GB_API int gb_cpu_median_f64(double const* const* in, int n, int64_t d, double* out) { return guarded([&] { return median(in, n, d, out); }); }
GB_API int gb_cpu_coordwise_f32(float const* const* in, int n, int64_t d, int rule, int arg, float* out) { return guarded([&] { return coordwise_host(in, n, d, rule, arg, out); }); }
GB_API int gb_cpu_coordwise_f64(double const* const* in, int n, int64_t d, int rule, int arg, double* out) { return guarded([&] { return coordwise_host(in, n, d, rule, arg, out); }); }
GB_API int gb_cpu_brute_f32(float const* const* in, int n, int64_t d, int f, float* out) { return guarded([&] { return brute(in, n, d, f, out); }); }
GB_API int gb_cpu_brute_f64(double const* const* in, int n, int64_t d, int f, double* out) { return guarded([&] { return brute(in, n, d, f, out); }); }
GB_API int gb_cpu_pairwise_sqdist_f32(float const* const* in, int n, int64_t d, double* out) { return guarded([&] { return sqdist_matrix(in, n, d, out); }); }
GB_API int gb_cpu_pairwise_sqdist_f64(double const* const* in, int n, int64_t d, double* out) { return guarded([&] { return sqdist_matrix(in, n, d, out); }); }

GB_API int gb_cpu_selection_average_f32(float const* const* in, int64_t d, int const* sel, int count, float* out) {
    return guarded([&] { if (count < 1) return fail("selection_average: empty selection"); selection_average(in, d, sel, count, out); return 0; });
}
GB_API int gb_cpu_selection_average_f64(double const* const* in, int64_t d, int const* sel, int count, double* out) {
    return guarded([&] { if (count < 1) return fail("selection_average: empty selection"); selection_average(in, d, sel, count, out); return 0; });
}
GB_API int gb_cpu_weighted_sum_f32(float const* const* in, int64_t d, int const* sel, float const* w, int count, float* out) {
    return guarded([&] { if (count < 1) return fail("weighted_sum: empty selection"); weighted_sum(in, d, sel, w, count, out); return 0; });
}

GB_API uint64_t gb_combinations_count(unsigned n, unsigned k, int* overflow) {
    bool over = false;
    uint64_t const v = gb::binomial(n, k, &over);
    if (overflow) *overflow = over ? 1 : 0;
    return v;
}
GB_API void gb_combinations_seek(unsigned n, unsigned k, uint64_t pos, int* idx) { gb::combination_unrank(n, k, pos, idx); }
GB_API int gb_combinations_next(unsigned n, unsigned k, int* idx) { return gb::combination_next(n, k, idx) ? 1 : 0; }

// Test hook: marks every index of [0, total) once per covering chunk; returns the number of chunks.
GB_API int gb_cpu_parallel_for_probe(int64_t total, int64_t grain, int32_t* hits) {
    std::atomic<int> chunks{0};
    gb::parallel_for(0, static_cast<std::size_t>(total), [&](std::size_t lo, std::size_t hi) {
        chunks.fetch_add(1);
        for (std::size_t i = lo; i < hi; ++i) __atomic_fetch_add(&hits[i], 1, __ATOMIC_RELAXED);
    }, static_cast<std::size_t>(grain));
    return chunks.load();
}

}  // extern "C"


// k-th smallest along `axis` of a row-major array (in place, like std::nth_element), one result per line:
// exercises gb::StridedView / StrideIterator (role of the reference's include/array.hpp) through the C ABI.
extern "C" GB_API int gb_cpu_axis_nth_f32(float* data, int rank, int64_t const* shape, int axis, int64_t k, float* out) {
    if (rank < 1 || rank > gb::StridedView<float>::kMaxRank || axis < 0 || axis >= rank) return fail("axis_nth: rank must be 1..4 and axis inside it");
    std::size_t dims[4] = {1, 1, 1, 1};
    std::size_t lines = 1;
    for (int a = 0; a < rank; ++a) {
        if (shape[a] < 1) return fail("axis_nth: empty dimension");
        dims[a] = static_cast<std::size_t>(shape[a]);
        if (a != axis) lines *= dims[a];
    }
    if (k < 0 || k >= shape[axis]) return fail("axis_nth: k outside the axis");
    gb::StridedView<float> const view(data, rank, dims);
    gb::parallel_for(0, lines, [&](std::size_t lo, std::size_t hi) {
        for (std::size_t line = lo; line < hi; ++line) {
            gb::StridedView<float> v = view;     // fix every other axis, last to first, so earlier positions stay valid
            std::size_t rest = line;
            for (int a = rank - 1; a >= 0; --a) {
                if (a == axis) continue;
                v = v.at(a, rest % dims[a]);
                rest /= dims[a];
            }
            auto const range = v.line(0);        // the only axis left is `axis`
            std::nth_element(range.begin(), range.begin() + k, range.end());
            out[line] = range.begin()[k];
        }
    });
    return 0;
}
