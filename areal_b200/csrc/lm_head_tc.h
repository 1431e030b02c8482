// tcgen05 LM head with the online-softmax epilogue (internal; the public C ABI is include/areal_b200.h).
#pragma once
#include "common.cuh"

namespace ab {

int64_t lm_head_tc_workspace_bytes(int64_t T, int64_t V);
// hidden [T, H], weight [V, H] bf16 with unit stride in H; outputs f32[T]; entropy / raw_min nullable
int launch_lm_head_tc(const void* hidden, const void* weight, const int64_t* labels, int64_t T, int64_t H, int64_t V,
                      int64_t hidden_row_stride, int64_t weight_row_stride, float inv_temperature, float* logp,
                      float* entropy, float* row_max, float* ln_s, float* raw_min, void* workspace,
                      int64_t workspace_bytes, cudaStream_t stream);

}  // namespace ab
