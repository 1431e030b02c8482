// Internal launcher of the stats-tracker reduction kernel (public C ABI: include/areal_b200.h).
#pragma once
#include "common.cuh"

namespace ab {

constexpr int kMaxStatKeys = 32;
constexpr int64_t kStatChunk = 2048;   // elements per CTA: 8 per thread, all loads of a thread in flight at once

struct StatsTable {
  const float* val[kMaxStatKeys];      // null: the value of every element is 1 (a denominator's own count)
  const uint8_t* mask[kMaxStatKeys];   // null: every element counts
  double* out[kMaxStatKeys];           // 4 doubles per key: sum, count, min, max
  unsigned int val_u8_bits;            // bit k: val[k] points at u8/bool elements (read as 0.0 / 1.0 ...)
};

int64_t masked_stats_workspace_bytes(int64_t n_keys, int64_t n);
int launch_masked_stats(const StatsTable& tab, int n_keys, int64_t n, int accumulate, void* workspace,
                        cudaStream_t stream);

int launch_segment_masked_mean(const float* v, const uint8_t* mask, const int32_t* cu, int64_t B, double* mean,
                               uint8_t* valid, cudaStream_t stream);

}  // namespace ab
