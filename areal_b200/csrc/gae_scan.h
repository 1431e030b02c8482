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
  int vec_ok;                // set by the launcher: 128-bit row access allowed
  int64_t B, S;
  float g, gl;
  float* adv;
  float* ret;
};

// Reward construction of PPOActor.compute_advantages fused in front of the padded GAE
// (areal/engine/ppo/actor.py:92-129).  Mask element codes: 0 f32, 1 u8/bool, 2 i32, 3 i64.
struct AdvFusedParams {
  float* score;                // [B] task reward per sequence (data["rewards"]); the overlong penalty is
                               // added IN PLACE like areal/utils/functional.py:394-397
  int overlong;                // 1: reward_overlong_penalty first (actor.py:80-90)
  int64_t overlong_tokens, expected_len;   // expected_len = max_response_length - overlong_tokens
  float overlong_factor;
  const void* loss_mask;       // [B, S] NOT rolled; the kernel reads column (t+1) mod S  (:109-110)
  int loss_mask_code;
  const float* logp;           // [B, S] behaviour / proximal log-probs
  int roll_logp;               // 1: old_logp = roll(logp, -1) (:116); 0: used as is (:114)
  float* ref_logp;             // [B, S] or null (zeros, :120); multiplied by the mask IN PLACE (:121)
  const void* attn_mask;       // [B, S]; seqlens = row sum (:126)
  int attn_code;
  int apply_scaling;           // 1: score = clip((score + bias) * scaling, +-clip) (:98-102)
  float bias, scaling, clip;
  float neg_kl_ctl;            // float(-kl_ctl) (:128)
  int kl_kind;                 // 1, 2, 3 (areal/utils/data.py:1343-1360)
  int kl_clamp;                // clamp to [-10, 10] (:1362)
  int mask_no_eos_with_zero;   // (:133-138)
  float* kl_rewards;           // [B, S] out
  float* tot_rewards;          // [B, S] out
  float* loss_mask_out;        // [B, S] out, rolled, f32
  float* old_logp_out;         // [B, S] out, rolled (if roll_logp) and masked
};

int launch_advantages_padded(const GaePaddedParams& p, const AdvFusedParams& f, cudaStream_t stream);

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
