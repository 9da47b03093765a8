// Host-side helpers shared by the .cu translation units: error reporting and device facts.
#pragma once

#include <cuda_runtime.h>

namespace gb {

int cuda_fail(char const* message);                       // records message, returns 1
int cuda_check(cudaError_t err, char const* what);        // 0 when err == cudaSuccess
int device_sm_count();                                    // SMs of the current device (cached)
int pick_tile(int n);
struct RowTable;
bool rows_aligned(RowTable const& rows, int n);

}  // namespace gb
