// Internal launcher of the interval gather/scatter kernels (public C ABI: include/areal_b200.h).
#pragma once
#include "common.cuh"

namespace ab {

int64_t interval_workspace_bytes(int64_t n_intervals);
// slice: flat_in -> packed_out ; set: packed_in -> flat_out.  workspace: interval_workspace_bytes(n).
int launch_interval_copy(bool slice, const void* flat_in, void* flat_out, const void* packed_in,
                         void* packed_out, const int64_t* intervals, int64_t n, int64_t elem_size,
                         int64_t flat_numel, int64_t packed_numel, void* workspace,
                         cudaStream_t stream);

}  // namespace ab
