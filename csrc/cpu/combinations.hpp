// k-combinations of {0..n-1} in lexicographic order: count, unrank ("seek") and successor.
//
// Role parity: /root/reference/pytorch_impl/libs/native/include/combinations.hpp:57-178. The
// same unranking routine is compiled for the device (select_mda kernel) — keep it free of
// host-only constructs. Works on a caller-provided index array, no allocation.
#pragma once

#include <cstdint>

#if defined(__CUDACC__)
#define GB_HD __host__ __device__ __forceinline__
#else
#define GB_HD inline
#endif

namespace gb {

// C(n, k) with overflow detection; returns 0 and sets *overflow when it does not fit 64 bits.
GB_HD std::uint64_t binomial(unsigned n, unsigned k, bool* overflow = nullptr) {
    if (overflow) *overflow = false;
    if (k > n) return 0;
    if (k > n - k) k = n - k;
    std::uint64_t value = 1;
    for (unsigned i = 1; i <= k; ++i) {
        std::uint64_t const factor = n - k + i;
        // value * factor / i is exact at every step; guard the multiplication
        if (value > UINT64_MAX / factor) {
            if (overflow) *overflow = true;
            return 0;
        }
        value = value * factor / i;
    }
    return value;
}

// Write into idx[0..k) the combination of rank `pos` (0-based, lexicographic).
GB_HD void combination_unrank(unsigned n, unsigned k, std::uint64_t pos, int* idx) {
    unsigned candidate = 0;
    for (unsigned slot = 0; slot < k; ++slot) {
        for (;; ++candidate) {
            // combinations that keep `candidate` at this slot: choose the rest among the larger values
            std::uint64_t const block = binomial(n - candidate - 1, k - slot - 1);
            if (pos < block) break;
            pos -= block;
        }
        idx[slot] = static_cast<int>(candidate++);
    }
}

// Advance idx to its lexicographic successor; false when idx was the last combination.
GB_HD bool combination_next(unsigned n, unsigned k, int* idx) {
    for (int slot = static_cast<int>(k) - 1; slot >= 0; --slot) {
        if (idx[slot] < static_cast<int>(n - k) + slot) {
            ++idx[slot];
            for (unsigned later = static_cast<unsigned>(slot) + 1; later < k; ++later) idx[later] = idx[later - 1] + 1;
            return true;
        }
    }
    return false;
}

}  // namespace gb
