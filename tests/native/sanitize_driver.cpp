// Standalone driver of the C++ host runtime for ThreadSanitizer / AddressSanitizer+UBSan builds (tests/test_sanitizers.py):
// every host rule on random inputs (the thread pool runs underneath), the parallel_for probe, the combinations
// iterator, the strided-view order statistics and the multi-replica register under concurrent readers and writers.
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <random>
#include <thread>
#include <vector>

#include "garfield_b200.h"

static int failures = 0;
#define CHECK(cond)                                                      \
    do {                                                                 \
        if (!(cond)) { std::fprintf(stderr, "CHECK failed: %s (line %d): %s\n", #cond, __LINE__, gb_cpu_last_error()); ++failures; } \
    } while (0)

int main() {
    std::mt19937 rng(7);
    std::normal_distribution<float> normal(0.f, 1.f);
    int const n = 11, f = 2;
    int64_t const d = 50021;    // not a multiple of the block size
    std::vector<std::vector<float>> rows(n, std::vector<float>(d));
    for (auto& r : rows) for (auto& v : r) v = normal(rng);
    rows[3][17] = NAN; rows[5][4000] = INFINITY;
    std::vector<float const*> ptrs;
    for (auto& r : rows) ptrs.push_back(r.data());
    std::vector<float> out(d);
    CHECK(gb_cpu_median_f32(ptrs.data(), n, d, out.data()) == 0);
    CHECK(gb_cpu_krum_f32(ptrs.data(), n, d, f, n - f - 2, out.data()) == 0);
    CHECK(gb_cpu_bulyan_f32(ptrs.data(), n, d, f, n - f - 2, out.data()) == 0);
    CHECK(gb_cpu_brute_f32(ptrs.data(), n, d, f, out.data()) == 0);
    CHECK(gb_cpu_coordwise_f32(ptrs.data(), n, d, 0, 0, out.data()) == 0);                 // lower median: NaN columns are exact
    CHECK(std::isnan(out[17]) && !std::isnan(out[16]));
    CHECK(gb_cpu_coordwise_f32(ptrs.data(), n, d, 1, f, out.data()) == 2);                 // trimmed mean declines the NaN column
    CHECK(gb_cpu_coordwise_f32(ptrs.data(), n, d, 2, n - f, out.data()) == 2);             // and so does mean-around-median
    CHECK(gb_cpu_coordwise_f32(ptrs.data() + 6, 5, d, 1, 1, out.data()) == 0);             // rows 6..10 are finite
    CHECK(gb_cpu_coordwise_f32(ptrs.data() + 6, 5, d, 2, 3, out.data()) == 0);
    std::vector<double> dist(static_cast<size_t>(n) * n);
    CHECK(gb_cpu_pairwise_sqdist_f32(ptrs.data(), n, d, dist.data()) == 0);
    int const sel[3] = {0, 4, 9};
    CHECK(gb_cpu_selection_average_f32(ptrs.data(), d, sel, 3, out.data()) == 0);
    CHECK(std::fabs(out[1] - (rows[0][1] + rows[4][1] + rows[9][1]) / 3.f) < 1e-5f);
    CHECK(gb_cpu_krum_f32(ptrs.data(), 4, d, 1, 1, out.data()) != 0);          // n < 2f+3 is rejected, not UB

    std::vector<int32_t> hits(100003, 0);
    CHECK(gb_cpu_parallel_for_probe(static_cast<int64_t>(hits.size()), 0, hits.data()) >= 1);
    for (int32_t h : hits) if (h != 1) { ++failures; break; }

    int overflow = 0;
    CHECK(gb_combinations_count(10, 4, &overflow) == 210 && overflow == 0);
    int idx[4];
    gb_combinations_seek(10, 4, 209, idx);
    CHECK(idx[0] == 6 && idx[3] == 9 && gb_combinations_next(10, 4, idx) == 0);

    std::vector<float> cube(5 * 7 * 9);
    for (auto& v : cube) v = normal(rng);
    int64_t const shape[3] = {5, 7, 9};
    std::vector<float> kth(5 * 9);
    CHECK(gb_cpu_axis_nth_f32(cube.data(), 3, shape, 1, 3, kth.data()) == 0);

    // multi-replica register: 2 writers, 3 readers, then close
    void* reg = gb_mreg_create(3, 64, 0);
    CHECK(reg != nullptr);
    std::atomic<int> reads{0};
    std::vector<std::thread> threads;
    for (int w = 0; w < 2; ++w)
        threads.emplace_back([reg, w] {
            float payload[16];
            for (int i = 0; i < 200; ++i) {
                for (float& v : payload) v = static_cast<float>(w * 1000 + i);
                if (gb_mreg_write(reg, payload, 1000000000LL) == 1) return;
            }
        });
    for (int r = 0; r < 3; ++r)
        threads.emplace_back([reg, &reads] {
            float got[16];
            uint64_t seq = 0;
            for (int i = 0; i < 200; ++i) {
                int const rc = gb_mreg_read(reg, got, &seq, 2000000LL);
                if (rc == 1) return;
                if (rc == 0) {
                    for (int k = 1; k < 16; ++k) if (got[k] != got[0]) { std::fprintf(stderr, "torn read\n"); ++failures; return; }
                    reads.fetch_add(1);
                }
            }
        });
    for (size_t t = 0; t < 2; ++t) threads[t].join();
    gb_mreg_close(reg);
    for (size_t t = 2; t < threads.size(); ++t) threads[t].join();
    gb_mreg_destroy(reg);
    CHECK(reads.load() > 0);
    std::printf("sanitize driver: %d failure(s), %d reads, %d pool threads\n", failures, reads.load(), gb_cpu_num_threads());
    return failures == 0 ? 0 : 1;
}
