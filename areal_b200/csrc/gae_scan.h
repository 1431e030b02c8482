// Parameter blocks of the GAE scan kernels (internal; the public C ABI is include/areal_b200.h).
#pragma once
#include "common.cuh"

namespace ab {

struct Gae1dParams {
  const float* rewards;      // [T]
  const float* values;       // [T + B]
  const int32_t* cu;         // [B + 1]
  const uint8_t* bootstrap;  // [B]
  int64_t T, B;
  float g;                   // (float)gamma
  float gl;                  // (float)(gamma * lam), product taken in double like the Python reference
  float* adv;
  float* ret;
  void* workspace;           // look-back descriptors; cleared by the launcher (stream-ordered)
  int vec_ok;                // rewards / adv / ret 16-byte aligned
};

struct GaePaddedParams {
  const float* rewards;      // [B, S]
  const float* values;       // [B, S]
  const float* mask;         // [B, S] already rolled by -1 (actor.py:110); != 0 means "counts"
  const uint8_t* no_eos;     // [B]
  int64_t B, S;
  float g, gl;
  float* adv;
  float* ret;
};

struct Gae2dParams {
  const float* rewards;            // [B, H]
  const float* values;             // [B, H + 1]
  const uint8_t* done;             // [B, H + 1]
  const int32_t* done_y_indices;   // [num_dones], ascending inside each row
  const int32_t* cu_num_dones;     // [B + 1]
  const uint8_t* is_truncate;      // [num_dones]
  int64_t B, H;
  float g, gl;                     // gl = (float)gamma * (float)lam (device fp32 product, gae.cu:84)
  float* adv;
  float* ret;
};

int64_t gae1d_workspace_bytes(int64_t T);
int launch_gae1d(const Gae1dParams& p, cudaStream_t stream);
int launch_gae1d_static(Gae1dParams p, cudaStream_t stream);   // workspace member is ignored
int launch_gae_padded(const GaePaddedParams& p, cudaStream_t stream);
int launch_gae2d_olp(const Gae2dParams& p, cudaStream_t stream);

}  // namespace ab
