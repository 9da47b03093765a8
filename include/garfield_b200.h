/* garfield_b200 — C ABI of the native libraries.
 *
 *   libgarfield_b200_cpu.so   host runtime + CPU aggregation rules   (csrc/cpu, g++)
 *   libgarfield_b200_cuda.so  sm_100a kernels + device runtime       (csrc/cuda, nvcc)
 *
 * Every function returns 0 on success unless stated otherwise; on failure a message is
 * available from gb_cpu_last_error() / gb_cuda_last_error() (thread-local).
 * Pointers named `in` are arrays of n row pointers (one per gradient), never a dense matrix,
 * so inputs can live anywhere (including peer-mapped memory of another GPU).
 * "ref:" comments cite the reference component each entry point is the counterpart of
 * (paths relative to /root/reference/pytorch_impl/libs/native unless absolute).
 */
#ifndef GARFIELD_B200_H
#define GARFIELD_B200_H

#include <stdint.h>

#if defined(__GNUC__)
#define GB_API __attribute__((visibility("default")))
#else
#define GB_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ host library --- */

GB_API char const* gb_cpu_last_error(void);
GB_API int gb_cpu_num_threads(void);                         /* ref: so_threadpool/threadpool.cpp:101-106 */

/* Multi-Krum, native semantics. ref: py_krum/krum.cpp:50-118 */
GB_API int gb_cpu_krum_f32(float const* const* in, int n, int64_t d, int f, int m, float* out);
GB_API int gb_cpu_krum_f64(double const* const* in, int n, int64_t d, int f, int m, double* out);
/* Bulyan over Multi-Krum, native semantics. ref: py_bulyan/bulyan.cpp:53-193 */
GB_API int gb_cpu_bulyan_f32(float const* const* in, int n, int64_t d, int f, int m, float* out);
GB_API int gb_cpu_bulyan_f64(double const* const* in, int n, int64_t d, int f, int m, double* out);
/* Non-finite-resilient coordinate-wise median. ref: py_median/median.cpp:42-77 */
GB_API int gb_cpu_median_f32(float const* const* in, int n, int64_t d, float* out);
GB_API int gb_cpu_median_f64(double const* const* in, int n, int64_t d, double* out);
/* Brute / MDA. ref: py_brute/brute.cpp:47-113 */
GB_API int gb_cpu_brute_f32(float const* const* in, int n, int64_t d, int f, float* out);
GB_API int gb_cpu_brute_f64(double const* const* in, int n, int64_t d, int f, double* out);
/* n×n squared distances (row-major doubles, +inf diagonal). ref: include/operations.hpp:41-55 */
GB_API int gb_cpu_pairwise_sqdist_f32(float const* const* in, int n, int64_t d, double* out);
GB_API int gb_cpu_pairwise_sqdist_f64(double const* const* in, int n, int64_t d, double* out);
/* out = mean of in[sel[0..count)]. ref: include/operations.hpp:64-79 */
GB_API int gb_cpu_selection_average_f32(float const* const* in, int64_t d, int const* sel, int count, float* out);
GB_API int gb_cpu_selection_average_f64(double const* const* in, int64_t d, int const* sel, int count, double* out);
GB_API int gb_cpu_weighted_sum_f32(float const* const* in, int64_t d, int const* sel, float const* w, int count, float* out);

/* k-combinations in lexicographic order. ref: include/combinations.hpp:57-178 */
GB_API uint64_t gb_combinations_count(unsigned n, unsigned k, int* overflow);
GB_API void gb_combinations_seek(unsigned n, unsigned k, uint64_t pos, int* idx);
GB_API int gb_combinations_next(unsigned n, unsigned k, int* idx);   /* 1 = advanced, 0 = was last */

/* Thread-pool coverage probe for tests. ref: include/threadpool.hpp:202-222 */
GB_API int gb_cpu_parallel_for_probe(int64_t total, int64_t grain, int32_t* hits);

/* Multi-replica register. ref: include/multiregister.hpp:232-552.
 * write/read return 0 ok, 1 closed, 2 timed out; timeout_ns < 0 waits forever. */
GB_API void* gb_mreg_create(int replicas, int64_t nbytes, int consume);
GB_API void gb_mreg_close(void* handle);
GB_API void gb_mreg_destroy(void* handle);
GB_API int gb_mreg_write(void* handle, void const* src, int64_t timeout_ns);
GB_API int gb_mreg_read(void* handle, void* dst, uint64_t* seq, int64_t timeout_ns);

#ifdef __cplusplus
}
#endif

#endif /* GARFIELD_B200_H */
