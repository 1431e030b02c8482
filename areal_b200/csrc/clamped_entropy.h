// Internal launchers of the clamped-entropy kernels (public C ABI: include/areal_b200.h).
#pragma once
#include "common.cuh"

namespace ab {

struct ClampedParams {
  const void* logits;          // [T, V] bf16 / fp16
  int dtype;                   // kBF16 or kFP16
  int64_t T, V, row_stride_bytes;
  const int64_t* labels;       // [T]
  float inv_temperature;
  int64_t k;                   // number of smallest logits masked out per row, 0 <= k < V
  float* logp;                 // [T] out
  float* entropy;              // [T] out (clamped)
  float* saved;                // [T, 8] out (forward) / in (backward)
  int only_flagged;            // forward, internal: redo only the rows the one-pass kernel flagged in saved[row][7]
};

int launch_clamped_entropy_fwd(const ClampedParams& p, cudaStream_t stream);
int launch_clamped_entropy_bwd(const ClampedParams& p, const float* g_logp, const float* g_ent, void* dlogits,
                               int64_t d_row_stride_bytes, cudaStream_t stream);

}  // namespace ab
