// Masked / grouped normalisation of advantages and rewards (the first "next" row).
//
// Replaces (reference, relative to its root):
//   areal/utils/data.py:1073-1304      Normalization.__call__ / _compute_mean / _compute_std
//                                      (Python loops over groups, `.item()` host syncs, ~10 torch
//                                       kernels per group, up to 4 scalar all-reduces)
//   areal/utils/functional.py:130-168  masked_normalization (== realhf/.../functional.py:251-313)
//
// Three launches whatever the number of groups: (1) fp64 moments per group, as per-slab partials in a
// fixed layout (deterministic); the host layer sums the few partials and, for batch-level statistics
// under data parallelism, all-reduces ONE small tensor; (2) masked centred sum of squares per group,
// with the per-element mean rule (plain, leave-one-out, group special cases); (3) apply.  The
// reference's dtype path is part of its result and is reproduced: batch-level statistics stay fp64
// until the final cast, group-level ones pass through fp32 (`zeros_like(x)`, data.py:1131, 1173).
#include "common.cuh"
#include "masked_norm.h"

namespace ab {
namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// block-wide sum of up to 3 doubles; result valid on thread 0
__device__ __forceinline__ void block_sum3(double& a, double& b, double& c) {
  __shared__ double sm[3][kThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  a = warp_sum_d(a);
  b = warp_sum_d(b);
  c = warp_sum_d(c);
  if (lane == 0) {
    sm[0][warp] = a;
    sm[1][warp] = b;
    sm[2][warp] = c;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    a = b = c = 0.0;
    for (int w = 0; w < kThreads / 32; ++w) {
      a += sm[0][w];
      b += sm[1][w];
      c += sm[2][w];
    }
  }
}

// grid (P, G): slab p of group g -> partials[g][p] = (sum m, sum x*m, sum (x*m)^2)
__global__ void __launch_bounds__(kThreads)
moments_kernel(const float* __restrict__ x, const float* __restrict__ mask, int64_t group_elems,
               double* __restrict__ partials) {
  const int64_t g = blockIdx.y, P = gridDim.x, p = blockIdx.x;
  const int64_t slab = (group_elems + P - 1) / P;
  const int64_t lo = g * group_elems + p * slab;
  int64_t hi = lo + slab;
  const int64_t gend = (g + 1) * group_elems;
  if (hi > gend) hi = gend;
  double c = 0.0, s = 0.0, q = 0.0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kThreads) {
    const double m = mask ? double(mask[i]) : 1.0;
    const double xm = double(x[i]) * m;
    c += m;
    s += xm;
    q += xm * xm;
  }
  block_sum3(c, s, q);
  if (threadIdx.x == 0) {
    double* o = partials + (g * P + p) * 3;
    o[0] = c;
    o[1] = s;
    o[2] = q;
  }
}

// per-element mean following _compute_mean (data.py:1206-1262)
__device__ __forceinline__ double elem_mean(const NormMeanSpec& ms, int64_t i, float xv, float mv, bool has_mask) {
  if (ms.mode == kNormNone) return 0.0;
  const int64_t g = i / ms.group_elems;
  if (g >= ms.n_groups) return 0.0;                              // rows beyond bs // group_size: zeros_like
  if (ms.zero_special) return 0.0;                               // group_size == 1 and leave-one-out (:1138-1143)
  const double F = ms.stats[g * 3 + 0], X = ms.stats[g * 3 + 1];
  double mean;
  if (ms.leave_one_out) {
    if (F <= 1.0) mean = 0.0;                                    // :1235-1236
    else if (!has_mask) mean = (X - double(xv)) / (F - 1.0);     // :1240-1243
    else {
      const double m = double(mv);
      mean = m > 0.0 ? (X - double(xv) * m) / fmax(F - m, 1.0) : X / F;   // :1245-1256
    }
  } else {
    mean = F == 0.0 ? 0.0 : X / F;                               // :1258-1260
  }
  return ms.mode == kNormGroup ? double(float(mean)) : mean;     // group level is stored in fp32 (:1131)
}

// grid (P, Gs): partials[g][p] = (sum m, sum (x*m - mean*m)^2)   (_compute_std, data.py:1264-1304)
__global__ void __launch_bounds__(kThreads)
centered_sumsq_kernel(const float* __restrict__ x, const float* __restrict__ mask, NormMeanSpec ms,
                      int64_t std_group_elems, double* __restrict__ partials) {
  const int64_t g = blockIdx.y, P = gridDim.x, p = blockIdx.x;
  const int64_t slab = (std_group_elems + P - 1) / P;
  const int64_t lo = g * std_group_elems + p * slab;
  int64_t hi = lo + slab;
  const int64_t gend = (g + 1) * std_group_elems;
  if (hi > gend) hi = gend;
  double c = 0.0, q = 0.0, unused = 0.0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kThreads) {
    const float xv = x[i], mv = mask ? mask[i] : 1.f;
    const double mean = elem_mean(ms, i, xv, mv, mask != nullptr);
    const double m = double(mv);
    const double d = mask ? double(xv) * m - mean * m : double(xv) - mean;
    c += m;
    q += d * d;
  }
  block_sum3(c, q, unused);
  if (threadIdx.x == 0) {
    double* o = partials + (g * P + p) * 2;
    o[0] = c;
    o[1] = q;
  }
}

__global__ void __launch_bounds__(kThreads)
norm_apply_kernel(const float* __restrict__ x, const float* __restrict__ mask, int64_t N, NormMeanSpec ms,
                  NormStdSpec ss, const float* __restrict__ total_mask_sum, float* __restrict__ out) {
  const bool passthrough = total_mask_sum != nullptr && total_mask_sum[0] == 0.f;   // data.py:1115-1116
  for (int64_t i = blockIdx.x * int64_t(kThreads) + threadIdx.x; i < N; i += int64_t(gridDim.x) * kThreads) {
    const float xv = x[i];
    if (passthrough) {
      out[i] = xv;
      continue;
    }
    const float mv = mask ? mask[i] : 1.f;
    const double mean = elem_mean(ms, i, xv, mv, mask != nullptr);
    const bool mean64 = ms.mode == kNormBatch;
    // x_centered = x - mean (fp64 when the mean is batch level, fp32 otherwise), then * loss_mask
    double xc;
    float xc32 = 0.f;
    if (mean64) {
      xc = double(xv) - mean;
      if (mask) xc *= double(mv);
    } else {
      xc32 = __fsub_rn(xv, float(mean));
      if (mask) xc32 = __fmul_rn(xc32, mv);
      xc = double(xc32);
    }
    // std per element
    double sd;
    bool std64 = false;
    double eps = ss.eps;
    if (ss.mode == kNormNone) {
      sd = 1.0;
      eps = 0.0;                                                 // :1199-1201
    } else {
      const int64_t g = i / ss.group_elems;
      if (g >= ss.n_groups) sd = 0.0;                            // zeros_like rows beyond bs // group_size
      else if (ss.one_special) sd = 1.0;                         // group_size == 1 and unbiased (:1180-1185)
      else {
        const double F = ss.stats[g * 2 + 0], Q = ss.stats[g * 2 + 1];
        if (ss.unbiased) sd = F <= 1.0 ? 1.0 : sqrt(Q / (F - 1.0));        // :1296-1299
        else sd = F == 0.0 ? 1.0 : sqrt(Q / F);                            // :1301-1303
      }
      std64 = ss.mode == kNormBatch;
      if (!std64) sd = double(float(sd));
    }
    float y;
    if (!mean64 && !std64) {
      y = __fdiv_rn(xc32, __fadd_rn(float(sd), float(eps)));
    } else {
      const double den = std64 ? sd + eps : double(__fadd_rn(float(sd), float(eps)));
      y = float(xc / den);
    }
    out[i] = y;
  }
}

// masked_normalization: ((x*m - mean) / (sqrt(var) + eps)).float(), var = E[(xm)^2] - mean^2
__global__ void __launch_bounds__(kThreads)
masked_normalization_apply_kernel(const float* __restrict__ x, const float* __restrict__ mask, int64_t N,
                                  const double* __restrict__ stats, int unbiased, double eps,
                                  float* __restrict__ out) {
  const double F = stats[0];
  const double mean = stats[1] / F;
  double var = stats[2] / F - mean * mean;
  if (unbiased) var *= F / (F - 1.0);
  const double den = sqrt(var) + eps;
  for (int64_t i = blockIdx.x * int64_t(kThreads) + threadIdx.x; i < N; i += int64_t(gridDim.x) * kThreads) {
    const double xm = mask ? double(x[i]) * double(mask[i]) : double(x[i]);
    out[i] = float((xm - mean) / den);
  }
}

unsigned int grid_1d(int64_t N) {
  const int64_t cap = int64_t(num_sms()) * 8;
  const int64_t need = (N + kThreads - 1) / kThreads;
  return static_cast<unsigned int>(need < 1 ? 1 : (need < cap ? need : cap));
}

// ---- legacy running value statistics (realhf/impl/model/modules/rms.py) ---------------------------
// state (module dtype R = double or float): [0] mean, [1] mean of squares, [2] debiasing term
// (exponential) or accumulated denominator (moving average).  One thread: the update is scalar work;
// what matters is that it stays on the device (no .item()) between the moments kernel and the next use.
template <typename R>
__global__ void rms_update_kernel(const double* __restrict__ partials, int64_t P, R* __restrict__ state, int mode,
                                  double beta) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  double factor = 0.0, x_sum = 0.0, x_sum_sq = 0.0;
  for (int64_t i = 0; i < P; ++i) {                 // fixed order
    factor += partials[3 * i + 0];
    x_sum += partials[3 * i + 1];
    x_sum_sq += partials[3 * i + 2];
  }
  const R f = R(factor), sx = R(x_sum), sq = R(x_sum_sq);
  if (mode == 0) {                                  // rms.py:74-83
    const R b = R(beta), omb = R(1.0 - beta);
    const R batch_mean = sx / f, batch_sq_mean = sq / f;
    state[0] = b * state[0] + batch_mean * omb;
    state[1] = b * state[1] + batch_sq_mean * omb;
    state[2] = b * state[2] + R(1.0) - b;
  } else {                                          // rms.py:164-170
    const R acc = state[2];
    state[0] = (acc * state[0] + sx) / (acc + f);
    state[1] = (acc * state[1] + sq) / (acc + f);
    state[2] = acc + f;
  }
}

template <typename R>
__device__ __forceinline__ void rms_mean_std(const R* state, int mode, R eps, R& mean, R& std_) {
  R m = state[0], msq = state[1];
  if (mode == 0) {                                  // rms.py:86-90
    R d = state[2];
    d = d < eps ? eps : d;                          // clamp(min=eps)
    m = m / d;
    msq = msq / d;
  }
  R var = msq - m * m;                              // :89 / :176
  var = var < R(1e-2) ? R(1e-2) : var;              // clamp(min=1e-2); NaN stays NaN
  mean = m;
  std_ = sqrt(var);
}

template <typename R, bool kDenorm>
__global__ void __launch_bounds__(kThreads)
rms_apply_kernel(const float* __restrict__ x, int64_t N, const R* __restrict__ state, int mode, R eps,
                 float* __restrict__ out) {
  R mean, std_;
  rms_mean_std<R>(state, mode, eps, mean, std_);
  for (int64_t i = blockIdx.x * int64_t(kThreads) + threadIdx.x; i < N; i += int64_t(gridDim.x) * kThreads) {
    const R v = R(x[i]);
    if (kDenorm) {
      out[i] = float(v * std_ + mean);              // rms.py:104-108
    } else {
      R y = (v - mean) / std_;                      // rms.py:95-101
      y = y < R(-5) ? R(-5) : (y > R(5) ? R(5) : y);
      out[i] = float(y);
    }
  }
}

}  // namespace

int64_t norm_num_partials(int64_t group_elems) {
  const int64_t p = (group_elems + 16383) / 16384;
  return p < 1 ? 1 : (p > 128 ? 128 : p);
}

int launch_moments(const float* x, const float* mask, int64_t n_groups, int64_t group_elems, int64_t P,
                   double* partials, cudaStream_t stream) {
  if (n_groups == 0) return kOk;
  moments_kernel<<<dim3(static_cast<unsigned int>(P), static_cast<unsigned int>(n_groups)), kThreads, 0, stream>>>(
      x, mask, group_elems, partials);
  return check_launch("norm moments");
}

int launch_centered_sumsq(const float* x, const float* mask, const NormMeanSpec& ms, int64_t n_std_groups,
                          int64_t std_group_elems, int64_t P, double* partials, cudaStream_t stream) {
  if (n_std_groups == 0) return kOk;
  centered_sumsq_kernel<<<dim3(static_cast<unsigned int>(P), static_cast<unsigned int>(n_std_groups)), kThreads, 0,
                          stream>>>(x, mask, ms, std_group_elems, partials);
  return check_launch("norm centered sumsq");
}

int launch_norm_apply(const float* x, const float* mask, int64_t N, const NormMeanSpec& ms, const NormStdSpec& ss,
                      const float* total_mask_sum, float* out, cudaStream_t stream) {
  if (N == 0) return kOk;
  norm_apply_kernel<<<grid_1d(N), kThreads, 0, stream>>>(x, mask, N, ms, ss, total_mask_sum, out);
  return check_launch("norm apply");
}

int launch_masked_normalization_apply(const float* x, const float* mask, int64_t N, const double* stats,
                                      int unbiased, double eps, float* out, cudaStream_t stream) {
  if (N == 0) return kOk;
  masked_normalization_apply_kernel<<<grid_1d(N), kThreads, 0, stream>>>(x, mask, N, stats, unbiased, eps, out);
  return check_launch("masked_normalization apply");
}

int launch_rms_update(const double* partials, int64_t P, void* state, int state_is_f64, int mode, double beta,
                      cudaStream_t stream) {
  if (state_is_f64) rms_update_kernel<double><<<1, 32, 0, stream>>>(partials, P, static_cast<double*>(state), mode, beta);
  else rms_update_kernel<float><<<1, 32, 0, stream>>>(partials, P, static_cast<float*>(state), mode, beta);
  return check_launch("rms update");
}

int launch_rms_apply(const float* x, int64_t N, const void* state, int state_is_f64, int mode, double eps,
                     int denormalize, float* out, cudaStream_t stream) {
  if (N == 0) return kOk;
  const unsigned int grid = grid_1d(N);
  if (state_is_f64) {
    const double* st = static_cast<const double*>(state);
    if (denormalize) rms_apply_kernel<double, true><<<grid, kThreads, 0, stream>>>(x, N, st, mode, eps, out);
    else rms_apply_kernel<double, false><<<grid, kThreads, 0, stream>>>(x, N, st, mode, eps, out);
  } else {
    const float* st = static_cast<const float*>(state);
    if (denormalize) rms_apply_kernel<float, true><<<grid, kThreads, 0, stream>>>(x, N, st, mode, float(eps), out);
    else rms_apply_kernel<float, false><<<grid, kThreads, 0, stream>>>(x, N, st, mode, float(eps), out);
  }
  return check_launch("rms apply");
}

}  // namespace ab
