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

// get_packed_rewards / get_packed_reward_dense (realhf ppo_functional.py:228-289): KL reward per token,
// plus the clipped task reward on the last token of each sequence (mode 0) or a per-token dense reward
// (mode 1: d[j+1]; mode 2: d[j+1] - d[j], indices in the one-longer-per-sequence layout).
constexpr int kRewItems = 4;
__global__ void __launch_bounds__(kThreads)
packed_rewards_kernel(const float* __restrict__ logp, const float* __restrict__ ref,
                      const float* __restrict__ score, const int32_t* __restrict__ cu,
                      const uint8_t* __restrict__ no_eos, int64_t Ts, int B, float neg_kl_ctl, float clip,
                      int mask_no_eos_with_zero, int mode, float* __restrict__ kl_out,
                      float* __restrict__ tot_out) {
  const int64_t nthreads = (Ts + kRewItems - 1) / kRewItems;
  for (int64_t th = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; th < nthreads;
       th += int64_t(gridDim.x) * blockDim.x) {
    const int64_t t0 = th * kRewItems;
    // sequence of t0: smallest i with cu[i+1] > t0
    int lo = 0, hi = B - 1;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (int64_t(__ldg(cu + mid + 1)) > t0) hi = mid; else lo = mid + 1;
    }
    int seq = lo;
    int64_t end = __ldg(cu + seq + 1);
#pragma unroll
    for (int k = 0; k < kRewItems; ++k) {
      const int64_t t = t0 + k;
      if (t >= Ts) break;
      while (t >= end && seq < B - 1) {
        ++seq;
        end = __ldg(cu + seq + 1);
      }
      const float klr = __fmul_rn(neg_kl_ctl, __fsub_rn(__ldg(logp + t), __ldg(ref + t)));   // :241
      float tot = klr;
      if (mode == 0) {
        // the LAST sequence ending at t+1 adds its score (duplicates: an empty sequence shares the index
        // of its predecessor and, as in the reference's index assignment, the later one wins)
        if (t + 1 == end) {
          int s2 = seq;
          while (s2 + 1 < B && int64_t(__ldg(cu + s2 + 2)) == end) ++s2;
          float sc = clamp_nan(__ldg(score + s2), -clip, clip);                            // :243
          if (mask_no_eos_with_zero && no_eos[s2]) sc = 0.f;                                  // :245-248
          tot = __fadd_rn(tot, sc);
        }
      } else {
        const int64_t j = t + seq;                                                             // offset2 + (t - offset)
        const float d1 = clamp_nan(__ldg(score + j + 1), -clip, clip);
        if (mode == 2) {
          const float d0 = clamp_nan(__ldg(score + j), -clip, clip);
          tot = __fadd_rn(tot, __fsub_rn(d1, d0));                                            // :272-275
        } else {
          tot = __fadd_rn(tot, d1);                                                           // :277-279
        }
      }
      kl_out[t] = klr;
      tot_out[t] = tot;
    }
  }
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

int launch_packed_rewards(const float* logp, const float* ref, const float* score, const int32_t* cu,
                          const uint8_t* no_eos, int64_t Ts, int64_t B, float neg_kl_ctl, float clip,
                          int mask_no_eos_with_zero, int mode, float* kl_out, float* tot_out,
                          cudaStream_t stream) {
  if (Ts == 0) return kOk;
  packed_rewards_kernel<<<grid_for((Ts + kRewItems - 1) / kRewItems), kThreads, 0, stream>>>(
      logp, ref, score, cu, no_eos, Ts, static_cast<int>(B), neg_kl_ctl, clip, mask_no_eos_with_zero, mode, kl_out,
      tot_out);
  return check_launch("packed_rewards");
}

int launch_scale(const float* coef, const float* s0, const float* s1, int64_t T, float* out,
                 cudaStream_t stream) {
  if (T == 0) return kOk;
  scale_kernel<<<grid_for(T), kThreads, 0, stream>>>(coef, s0, s1, T, out);
  return check_launch("scale");
}

}  // namespace ab
