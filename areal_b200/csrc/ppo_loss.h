// Internal launchers of the stand-alone PPO loss kernels (public C ABI: include/areal_b200.h).
#pragma once
#include "common.cuh"
#include "ppo_math.cuh"

namespace ab {

struct ActorLossParams {
  const float *logp, *prox, *old, *adv;   // prox nullable (legacy signature without proximal)
  const uint8_t* mask;                    // nullable == all ones (legacy loss_mask=None)
  int64_t T;
  ActorCfg cfg;
  float *loss_out, *inv_count_out;        // [1] each
  float* coef;                            // [T] count * dloss/dlogp
  float *stat_loss, *stat_ratio, *stat_kl, *behave_w, *behave_kl;   // nullable
  uint8_t *clip_mask, *dual_clip_mask, *behave_mask;                // nullable
  void* workspace;
};

struct CriticLossParams {
  const float *value, *old, *target;
  const uint8_t* mask;                    // nullable: plain mean
  int64_t T;
  float eps;
  int huber;
  float *loss_out, *inv_count_out;
  float* coef;                            // [T] count * dloss/dvalue
  float* stat_loss;                       // nullable
  uint8_t* clip_mask;                     // nullable
  void* workspace;
};

int64_t loss_workspace_bytes();
int launch_actor_loss(const ActorLossParams& p, cudaStream_t stream);
int launch_critic_loss(const CriticLossParams& p, cudaStream_t stream);
// out[t] = coef[t] * s0[0] * s1[0]   (s1 nullable)
int launch_scale(const float* coef, const float* s0, const float* s1, int64_t T, float* out,
                 cudaStream_t stream);

int launch_packed_rewards(const float* logp, const float* ref, const float* score, const int32_t* cu,
                          const uint8_t* no_eos, int64_t Ts, int64_t B, float neg_kl_ctl, float clip,
                          int mask_no_eos_with_zero, int mode, float* kl_out, float* tot_out,
                          cudaStream_t stream);

}  // namespace ab
