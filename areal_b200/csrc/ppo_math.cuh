// Per-token PPO math shared by the stand-alone loss kernels (ppo_loss.cu) and the fused
// logits epilogue (logits_fused.cu).  Follows `ppo_actor_loss_fn`
// (reference areal/utils/functional.py:171-235) and `ppo_critic_loss_fn` (:247-311) term by term;
// gradients are the closed forms of SURVEY.md 8a-6 (torch.max / torch.min tie splitting included).
#pragma once
#include "common.cuh"

namespace ab {

struct ActorCfg {
  float lo;        // (float)(1.0 - eps_clip)
  float hi;        // (float)(1.0 + (eps_clip_higher or eps_clip))
  float c_clip;    // valid iff has_c
  float cap;       // valid iff has_cap
  int has_c;
  int has_cap;
  int decoupled;   // proximal_logprobs given (always true for areal's signature)
};

struct ActorTokenOut {
  float pg;        // logging loss (pre-mask), functional.py:220
  float ratio;     // importance_weight
  float kl;        // approx_kl = logp - prox
  float w;         // behave_imp_weight
  float bk;        // behave_approx_kl
  float coef;      // count * dloss/dlogp
  bool clip, dual, bmask;
};

#ifdef __CUDACC__
__device__ __forceinline__ ActorTokenOut actor_token(float lp, float prox, float old, float A,
                                                     bool mask, const ActorCfg& c) {
  ActorTokenOut o;
  const float denorm = c.decoupled ? prox : old;
  o.kl = lp - denorm;
  const float e = expf(o.kl);
  o.ratio = mask ? e : 0.0f;                                  // :191
  const float clipped = clamp_nan(o.ratio, c.lo, c.hi);        // :193-197 (torch.clamp: NaN stays NaN)
  const float l1 = -A * o.ratio;                              // :199
  const float l2 = -A * clipped;                              // :200
  o.clip = (l1 < l2) && mask;                                 // :201, :222
  float pg = max_nan(l1, l2);                                 // :202 (torch.max propagates NaN)
  // backward of torch.max(a, b): each side gets grad, halved on a tie, zeroed only where it is strictly the
  // smaller one -- so with a NaN operand (every comparison false) BOTH sides get the full gradient
  const float w1 = l1 < l2 ? 0.0f : (l1 == l2 ? 0.5f : 1.0f);
  const float w2 = l1 > l2 ? 0.0f : (l1 == l2 ? 0.5f : 1.0f);
  const bool inside = (o.ratio >= c.lo) && (o.ratio <= c.hi);   // clamp passes the gradient inside [lo, hi] only
  float dpg = -A * (w1 + (inside ? w2 : 0.0f));
  o.dual = false;
  if (c.has_c) {
    const float sgn = (A > 0.0f) ? 1.0f : ((A < 0.0f) ? -1.0f : 0.0f);
    const float l3 = (sgn * c.c_clip) * A;                    // :205
    o.dual = (l3 < pg) && mask;                               // :206, :223
    const float wmin = pg > l3 ? 0.0f : (pg == l3 ? 0.5f : 1.0f);   // torch.min backward, NaN-aware as above
    pg = min_nan(pg, l3);                                     // :207
    dpg *= wmin;
  }
  o.w = 0.0f;
  o.bk = 0.0f;
  o.bmask = mask;
  if (c.decoupled) {
    const float bk = prox - old;                              // :210
    const float w = expf(bk);                                 // :211
    o.bmask = c.has_cap ? ((w <= c.cap) && mask) : mask;      // :212-216
    o.bk = o.bmask ? bk : 0.0f;
    o.w = o.bmask ? w : 0.0f;
    pg *= o.w;                                                // :219
    dpg *= o.w;
  }
  o.pg = pg;
  // backward of torch.where(mask, exp(kl), 0): the masked side hands a ZERO gradient to exp, whose own
  // backward multiplies it by exp(kl) -- 0 for any finite value, NaN for a NaN / infinite one, as autograd gives
  o.coef = (mask ? dpg : 0.0f) * e;
  return o;
}

// returns loss element; *g = d loss_el / d value, *clip = clipped loss strictly larger
__device__ __forceinline__ float critic_token(float v, float old, float tgt, float eps, bool huber,
                                              float* g, bool* clip) {
  const float diff = v - old;
  const float vc = old + clamp_nan(diff, -eps, eps);           // :291-293
  float lo_, lc_, dlo, dlc;
  {
    const float d0 = v - tgt, d1 = vc - tgt;
    if (!huber) {
      lo_ = 0.5f * (d0 * d0);                                   // _mse_loss :243-244
      lc_ = 0.5f * (d1 * d1);
      dlo = d0;
      dlc = d1;
    } else {
      const float delta = 10.0f;                                // :280
      const float a0 = fabsf(d0), a1 = fabsf(d1);
      lo_ = a0 < delta ? 0.5f * (a0 * a0) : delta * (a0 - 0.5f * delta);   // _huber_loss :238-240
      lc_ = a1 < delta ? 0.5f * (a1 * a1) : delta * (a1 - 0.5f * delta);
      // sign(NaN) is NaN in autograd's |x| backward: a NaN difference keeps a NaN gradient
      dlo = a0 < delta ? d0 : (d0 != d0 ? d0 : copysignf(delta, d0));
      dlc = a1 < delta ? d1 : (d1 != d1 ? d1 : copysignf(delta, d1));
    }
  }
  const bool inside = (diff >= -eps) && (diff <= eps);
  const float w1 = lo_ < lc_ ? 0.0f : (lo_ == lc_ ? 0.5f : 1.0f);   // torch.max backward (see actor_token)
  const float w2 = lo_ > lc_ ? 0.0f : (lo_ == lc_ ? 0.5f : 1.0f);
  *g = w1 * dlo + (inside ? w2 * dlc : 0.0f);
  *clip = lc_ > lo_;                                          // :300
  return max_nan(lo_, lc_);                                   // :297
}
#endif

}  // namespace ab
