// Internal launchers of the normalisation kernels (public C ABI: include/areal_b200.h).
#pragma once
#include "common.cuh"

namespace ab {

enum NormMode : int { kNormBatch = 0, kNormGroup = 1, kNormNone = 2 };

struct NormMeanSpec {
  int mode;               // NormMode
  int leave_one_out;
  int zero_special;       // group_size == 1 and leave-one-out: the group mean is 0
  int64_t group_elems;    // elements per mean group (whole tensor for batch level)
  int64_t n_groups;
  const double* stats;    // [n_groups, 3] = (sum m, sum x*m, sum (x*m)^2), after any all-reduce
};

struct NormStdSpec {
  int mode;
  int unbiased;
  int one_special;        // group_size == 1 and unbiased: the group std is 1
  int64_t group_elems;
  int64_t n_groups;
  const double* stats;    // [n_groups, 2] = (sum m, sum of centred squares)
  double eps;
};

int64_t norm_num_partials(int64_t group_elems);
int launch_moments(const float* x, const float* mask, int64_t n_groups, int64_t group_elems, int64_t P,
                   double* partials, cudaStream_t stream);
int launch_centered_sumsq(const float* x, const float* mask, const NormMeanSpec& ms, int64_t n_std_groups,
                          int64_t std_group_elems, int64_t P, double* partials, cudaStream_t stream);
int launch_norm_apply(const float* x, const float* mask, int64_t N, const NormMeanSpec& ms, const NormStdSpec& ss,
                      const float* total_mask_sum, float* out, cudaStream_t stream);
int launch_masked_normalization_apply(const float* x, const float* mask, int64_t N, const double* stats,
                                      int unbiased, double eps, float* out, cudaStream_t stream);

int launch_rms_update(const double* partials, int64_t P, void* state, int state_is_f64, int mode, double beta,
                      cudaStream_t stream);
int launch_rms_apply(const float* x, int64_t N, const void* state, int state_is_f64, int mode, double eps,
                     int denormalize, float* out, cudaStream_t stream);

}  // namespace ab
