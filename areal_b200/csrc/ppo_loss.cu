// Stand-alone PPO actor / critic loss kernels: per-token terms, the 8 logging vectors, the masked
// mean and the per-token gradient coefficient in ONE launch (no host sync for the mask count).
//
// Replaces (reference, relative to its root):
//   areal/utils/functional.py:171-235   ppo_actor_loss_fn   (~25 elementwise torch kernels + a
//                                       host sync at `loss_mask.count_nonzero() or 1`, :190)
//   areal/utils/functional.py:247-311   ppo_critic_loss_fn
//   realhf/impl/model/utils/ppo_functional.py:51-225  actor_loss_fn / critic_loss_fn (legacy)
//
// HBM-bound elementwise pass (~45 B/token), fixed-order reduction: thread-serial -> warp shuffle ->
// CTA -> per-CTA partials in a small workspace -> the last CTA (atomic ticket) sums them in index
// order, so the loss is bit-reproducible run to run.
#include "common.cuh"
#include "ppo_loss.h"

namespace ab {
namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ void block_finalize(float sum, float cnt, void* workspace,
                                               float* loss_out, float* inv_count_out,
                                               bool guard_zero, double fixed_count) {
  __shared__ float ssum[kThreads / 32], scnt[kThreads / 32];
  __shared__ unsigned int last_sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  sum = warp_sum(sum);
  cnt = warp_sum(cnt);
  if (lane == 0) {
    ssum[warp] = sum;
    scnt[warp] = cnt;
  }
  __syncthreads();
  if (tid != 0) return;
  float bs = 0.f, bc = 0.f;
  for (int w = 0; w < kThreads / 32; ++w) {
    bs += ssum[w];
    bc += scnt[w];
  }
  float2* partials = reinterpret_cast<float2*>(reinterpret_cast<char*>(workspace) + 16);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(workspace);
  partials[blockIdx.x] = make_float2(bs, bc);
  __threadfence();
  const unsigned int t = atomicAdd(ticket, 1u);
  if (t != gridDim.x - 1) return;
  __threadfence();
  double s = 0.0, c = 0.0;
  for (unsigned int i = 0; i < gridDim.x; ++i) {
    const volatile float* v = reinterpret_cast<const volatile float*>(&partials[i]);
    s += v[0];
    c += v[1];
  }
  if (fixed_count >= 0.0) c = fixed_count;
  if (guard_zero && c == 0.0) c = 1.0;            // `count_nonzero() or 1`, functional.py:190
  loss_out[0] = static_cast<float>(s / c);        // critic: no guard -> nan like the reference (:306)
  inv_count_out[0] = static_cast<float>(1.0 / c);
  *ticket = 0u;
}

__global__ void __launch_bounds__(kThreads)
actor_loss_kernel(const ActorLossParams p) {
  float sum = 0.f, cnt = 0.f;
  for (int64_t t = blockIdx.x * int64_t(kThreads) + threadIdx.x; t < p.T;
       t += int64_t(gridDim.x) * kThreads) {
    const bool mask = p.mask ? (p.mask[t] != 0) : true;
    const float old = p.old[t];
    const float prox = p.prox ? p.prox[t] : old;
    const ActorTokenOut o = actor_token(p.logp[t], prox, old, p.adv[t], mask, p.cfg);
    p.coef[t] = o.coef;
    if (p.stat_loss) p.stat_loss[t] = o.pg;
    if (p.stat_ratio) p.stat_ratio[t] = o.ratio;
    if (p.stat_kl) p.stat_kl[t] = o.kl;
    if (p.clip_mask) p.clip_mask[t] = o.clip;
    if (p.dual_clip_mask) p.dual_clip_mask[t] = o.dual;
    if (p.behave_w) p.behave_w[t] = o.w;
    if (p.behave_kl) p.behave_kl[t] = o.bk;
    if (p.behave_mask) p.behave_mask[t] = o.bmask;
    if (mask) {
      sum += o.pg;
      cnt += 1.f;
    }
  }
  block_finalize(sum, cnt, p.workspace, p.loss_out, p.inv_count_out, true, -1.0);
}

__global__ void __launch_bounds__(kThreads)
critic_loss_kernel(const CriticLossParams p) {
  float sum = 0.f, cnt = 0.f;
  for (int64_t t = blockIdx.x * int64_t(kThreads) + threadIdx.x; t < p.T;
       t += int64_t(gridDim.x) * kThreads) {
    const bool mask = p.mask ? (p.mask[t] != 0) : true;
    float g;
    bool clip;
    const float l = critic_token(p.value[t], p.old[t], p.target[t], p.eps, p.huber != 0, &g, &clip);
    p.coef[t] = mask ? g : 0.f;
    if (p.stat_loss) p.stat_loss[t] = l;
    if (p.clip_mask) p.clip_mask[t] = clip && mask;
    if (mask) {
      sum += l;
      cnt += 1.f;
    }
  }
  block_finalize(sum, cnt, p.workspace, p.loss_out, p.inv_count_out, false,
                 p.mask ? -1.0 : double(p.T));
}

__global__ void scale_kernel(const float* __restrict__ coef, const float* __restrict__ s0,
                             const float* __restrict__ s1, int64_t T, float* __restrict__ out) {
  const float s = s0[0] * (s1 ? s1[0] : 1.f);
  for (int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; t < T;
       t += int64_t(gridDim.x) * blockDim.x)
    out[t] = coef[t] * s;
}

unsigned int grid_for(int64_t T) {
  const int64_t cap = int64_t(num_sms()) * 4;
  const int64_t need = (T + kThreads - 1) / kThreads;
  return static_cast<unsigned int>(need < 1 ? 1 : (need < cap ? need : cap));
}

}  // namespace

int64_t loss_workspace_bytes() { return 16 + 8 * int64_t(num_sms()) * 8; }

int launch_actor_loss(const ActorLossParams& p, cudaStream_t stream) {
  actor_loss_kernel<<<grid_for(p.T), kThreads, 0, stream>>>(p);
  return check_launch("actor_loss");
}

int launch_critic_loss(const CriticLossParams& p, cudaStream_t stream) {
  critic_loss_kernel<<<grid_for(p.T), kThreads, 0, stream>>>(p);
  return check_launch("critic_loss");
}

int launch_scale(const float* coef, const float* s0, const float* s1, int64_t T, float* out,
                 cudaStream_t stream) {
  if (T == 0) return kOk;
  scale_kernel<<<grid_for(T), kThreads, 0, stream>>>(coef, s0, s1, T, out);
  return check_launch("scale");
}

}  // namespace ab
