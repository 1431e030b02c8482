// extern "C" entry points of libareal_b200.so (declared in include/areal_b200.h): argument
// validation, parameter-block assembly, launch.  No allocation, no synchronisation.
#include <math.h>
#include <stdarg.h>
#include <string.h>

#include "../../include/areal_b200.h"

// NVTX ranges around every C-ABI call (nsys / ncu launch lists then name the entry point each kernel belongs
// to).  Compile-time switch: AREAL_B200_NVTX=1 python -m areal_b200.build -f  (header-only NVTX v3, no extra library).
#ifdef AREAL_B200_NVTX
#include <nvtx3/nvToolsExt.h>
namespace {
struct NvtxScope {
  explicit NvtxScope(const char* name) { nvtxRangePushA(name); }
  ~NvtxScope() { nvtxRangePop(); }
};
}  // namespace
#define AB_NVTX_RANGE(name) NvtxScope ab_nvtx_scope_(name)
#else
#define AB_NVTX_RANGE(name) do { } while (0)
#endif
#include "common.cuh"
#include "gae_scan.h"
#include "interval_copy.h"
#include "logits_fused.h"
#include "masked_norm.h"
#include "masked_stats.h"
#include "clamped_entropy.h"
#include "lm_head_tc.h"
#include "ppo_loss.h"

namespace ab {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s launch failed: %s", what, cudaGetErrorString(e));
    return kCudaError;
  }
  return kOk;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

static int dtype_size(int dtype) {
  return dtype == kFP32 ? 4 : ((dtype == kBF16 || dtype == kFP16) ? 2 : 0);
}

static int make_cfg(ActorCfg* c, double eps_clip, double eps_clip_higher, double c_clip,
                    double cap, bool decoupled) {
  AB_REQUIRE(!isnan(eps_clip), "eps_clip must be a number");
  c->lo = static_cast<float>(1.0 - eps_clip);
  c->hi = static_cast<float>(1.0 + (isnan(eps_clip_higher) ? eps_clip : eps_clip_higher));
  c->has_c = !isnan(c_clip);
  AB_REQUIRE(!c->has_c || c_clip > 1.0, "c_clip must be > 1.0 (got %f)", c_clip);   // functional.py:204
  c->c_clip = c->has_c ? static_cast<float>(c_clip) : 0.f;
  c->has_cap = !isnan(cap);
  c->cap = c->has_cap ? static_cast<float>(cap) : 0.f;
  c->decoupled = decoupled ? 1 : 0;
  return kOk;
}

static int check_logits(const void* logits, int dtype, int64_t T, int64_t V, int64_t row_stride,
                        const int64_t* labels, double temperature) {
  AB_REQUIRE(T >= 0 && V > 0, "bad shape T=%lld V=%lld", (long long)T, (long long)V);
  AB_REQUIRE(dtype_size(dtype) != 0, "unsupported dtype %d", dtype);
  AB_REQUIRE(T == 0 || (logits && labels), "logits / labels must not be null");
  AB_REQUIRE(row_stride >= V, "row_stride (%lld) < V (%lld)", (long long)row_stride, (long long)V);
  AB_REQUIRE(temperature > 0.0 && isfinite(temperature), "temperature must be finite and > 0");
  return kOk;
}

}  // namespace ab

using namespace ab;

extern "C" {

AREAL_B200_API const char* areal_b200_last_error(void) { return g_err; }
AREAL_B200_API int areal_b200_version(void) { return 100; }

// ---- P1 ------------------------------------------------------------------------------------------
AREAL_B200_API int64_t areal_b200_gae1d_workspace_bytes(int64_t total_tokens) {
  return gae1d_workspace_bytes(total_tokens < 0 ? 0 : total_tokens);
}

AREAL_B200_API int areal_b200_gae1d_nolp_misalign(const float* rewards, const float* values,
                                   const int32_t* cu_seqlens, const uint8_t* bootstrap,
                                   int64_t total_tokens, int64_t batch_size, double gamma,
                                   double lam, float* adv_out, float* ret_out, void* workspace,
                                   int64_t workspace_bytes, void* stream) {
  AB_NVTX_RANGE("areal_b200_gae1d_nolp_misalign");
  AB_REQUIRE(total_tokens >= 0 && batch_size >= 0, "negative size");
  if (total_tokens == 0) return kOk;
  AB_REQUIRE(batch_size > 0, "total_tokens > 0 needs batch_size > 0");
  AB_REQUIRE(rewards && values && cu_seqlens && bootstrap && adv_out && ret_out, "null pointer");
  const int64_t need = gae1d_workspace_bytes(total_tokens);
  if (need > int64_t(sizeof(uint32_t)) * 16) {   // more than the header: look-back needs it
    if (!workspace || workspace_bytes < need) {
      set_error("gae1d workspace too small: need %lld bytes, got %lld", (long long)need,
                (long long)workspace_bytes);
      return kWorkspaceTooSmall;
    }
  }
  Gae1dParams p;
  p.rewards = rewards;
  p.values = values;
  p.cu = cu_seqlens;
  p.bootstrap = bootstrap;
  p.T = total_tokens;
  p.B = batch_size;
  p.g = static_cast<float>(gamma);
  p.gl = static_cast<float>(gamma * lam);
  p.adv = adv_out;
  p.ret = ret_out;
  p.workspace = workspace;
  p.vec_ok = ((reinterpret_cast<uintptr_t>(rewards) | reinterpret_cast<uintptr_t>(adv_out) |
               reinterpret_cast<uintptr_t>(ret_out)) & 15) == 0;
  return launch_gae1d(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_gae1d_nolp_misalign_static(const float* rewards, const float* values,
                                                         const int32_t* cu_seqlens,
                                                         const uint8_t* bootstrap,
                                                         int64_t total_tokens, int64_t batch_size,
                                                         double gamma, double lam, float* adv_out,
                                                         float* ret_out, void* stream) {
  AB_NVTX_RANGE("areal_b200_gae1d_nolp_misalign_static");
  AB_REQUIRE(total_tokens >= 0 && batch_size >= 0, "negative size");
  if (total_tokens == 0) return kOk;
  AB_REQUIRE(batch_size > 0, "total_tokens > 0 needs batch_size > 0");
  AB_REQUIRE(rewards && values && cu_seqlens && bootstrap && adv_out && ret_out, "null pointer");
  Gae1dParams p;
  p.rewards = rewards;
  p.values = values;
  p.cu = cu_seqlens;
  p.bootstrap = bootstrap;
  p.T = total_tokens;
  p.B = batch_size;
  p.g = static_cast<float>(gamma);
  p.gl = static_cast<float>(gamma * lam);
  p.adv = adv_out;
  p.ret = ret_out;
  p.workspace = nullptr;   // filled in by the launcher with the address of the built-in scratch
  p.vec_ok = ((reinterpret_cast<uintptr_t>(rewards) | reinterpret_cast<uintptr_t>(adv_out) |
               reinterpret_cast<uintptr_t>(ret_out)) & 15) == 0;
  return launch_gae1d_static(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_gae_padded_masked(const float* rewards, const float* values,
                                 const float* loss_mask, const uint8_t* seq_no_eos, int64_t B,
                                 int64_t S, double gamma, double lam, float* adv_out,
                                 float* ret_out, void* stream) {
  AB_NVTX_RANGE("areal_b200_gae_padded_masked");
  AB_REQUIRE(B >= 0 && S >= 0, "negative size");
  if (B == 0 || S == 0) return kOk;
  AB_REQUIRE(rewards && values && loss_mask && seq_no_eos && adv_out && ret_out, "null pointer");
  GaePaddedParams p;
  p.rewards = rewards;
  p.values = values;
  p.mask = loss_mask;
  p.no_eos = seq_no_eos;
  p.B = B;
  p.S = S;
  p.g = static_cast<float>(gamma);
  p.gl = static_cast<float>(gamma * lam);
  p.adv = adv_out;
  p.ret = ret_out;
  return launch_gae_padded(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_advantages_padded(
    float* score, int overlong, int64_t overlong_tokens, int64_t max_response_length,
    double overlong_penalty_factor, const void* loss_mask, int loss_mask_code, const float* logp, int roll_logp,
    float* ref_logp, const void* attn_mask, int attn_code, const float* values, int64_t B, int64_t S,
    int apply_scaling, double reward_bias, double reward_scaling, double reward_clip, double kl_ctl,
    int kl_kind, int kl_clamp, int mask_no_eos_with_zero, double gamma, double lam, float* adv_out,
    float* ret_out, float* kl_rewards_out, float* tot_rewards_out, float* loss_mask_out,
    float* old_logp_out, void* stream) {
  AB_NVTX_RANGE("areal_b200_advantages_padded");
  AB_REQUIRE(B >= 0 && S >= 0, "negative size");
  if (B == 0 || S == 0) return kOk;
  AB_REQUIRE(score && loss_mask && logp && attn_mask && adv_out && ret_out && kl_rewards_out &&
                 tot_rewards_out && loss_mask_out && old_logp_out, "null pointer");
  AB_REQUIRE(loss_mask_code >= 0 && loss_mask_code <= 3 && attn_code >= 0 && attn_code <= 3,
             "mask element code must be 0 (f32), 1 (u8/bool), 2 (i32) or 3 (i64)");
  AB_REQUIRE(!overlong || overlong_tokens != 0, "overlong_tokens must be non-zero");
  AB_REQUIRE(kl_kind >= 1 && kl_kind <= 3, "Invalid KL estimator: k%d. Valid choices: k1, k2, k3", kl_kind);
  GaePaddedParams p;
  p.rewards = nullptr;
  p.values = values;
  p.mask = nullptr;
  p.no_eos = nullptr;
  p.B = B;
  p.S = S;
  p.g = static_cast<float>(gamma);
  p.gl = static_cast<float>(gamma * lam);
  p.adv = adv_out;
  p.ret = ret_out;
  AdvFusedParams f;
  f.score = score;
  f.overlong = overlong;
  f.overlong_tokens = overlong_tokens;
  f.expected_len = max_response_length - overlong_tokens;
  f.overlong_factor = static_cast<float>(overlong_penalty_factor);
  f.loss_mask = loss_mask;
  f.loss_mask_code = loss_mask_code;
  f.logp = logp;
  f.roll_logp = roll_logp;
  f.ref_logp = ref_logp;
  f.attn_mask = attn_mask;
  f.attn_code = attn_code;
  f.apply_scaling = apply_scaling;
  f.bias = static_cast<float>(reward_bias);
  f.scaling = static_cast<float>(reward_scaling);
  f.clip = static_cast<float>(reward_clip);
  f.neg_kl_ctl = static_cast<float>(-kl_ctl);
  f.kl_kind = kl_kind;
  f.kl_clamp = kl_clamp;
  f.mask_no_eos_with_zero = mask_no_eos_with_zero;
  f.kl_rewards = kl_rewards_out;
  f.tot_rewards = tot_rewards_out;
  f.loss_mask_out = loss_mask_out;
  f.old_logp_out = old_logp_out;
  return launch_advantages_padded(p, f, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_gae2d_olp(const float* rewards, const float* values, const uint8_t* done,
                         const int32_t* done_y_indices, const int32_t* cu_num_dones,
                         const uint8_t* is_truncate, int64_t B, int64_t H, int64_t num_dones,
                         double gamma, double lam, float* adv_out, float* ret_out, void* stream) {
  AB_NVTX_RANGE("areal_b200_gae2d_olp");
  AB_REQUIRE(B >= 0 && H >= 0 && num_dones >= 0, "negative size");
  if (B == 0 || H == 0) return kOk;
  AB_REQUIRE(rewards && values && done && cu_num_dones && adv_out && ret_out, "null pointer");
  AB_REQUIRE(num_dones == 0 || (done_y_indices && is_truncate), "null segment table");
  Gae2dParams p;
  p.rewards = rewards;
  p.values = values;
  p.done = done;
  p.done_y_indices = done_y_indices;
  p.cu_num_dones = cu_num_dones;
  p.is_truncate = is_truncate;
  p.B = B;
  p.H = H;
  p.g = static_cast<float>(gamma);
  p.gl = static_cast<float>(gamma) * static_cast<float>(lam);
  p.adv = adv_out;
  p.ret = ret_out;
  return launch_gae2d_olp(p, static_cast<cudaStream_t>(stream));
}

// ---- P2 ------------------------------------------------------------------------------------------
static void fill_fwd_common(LogitsFwdParams* p, const void* logits, int dtype, int64_t T, int64_t V,
                            int64_t row_stride, const int64_t* labels, int64_t vocab_start,
                            double temperature) {
  memset(p, 0, sizeof(*p));
  p->logits = logits;
  p->dtype = dtype;
  p->T = T;
  p->V = V;
  p->row_stride_bytes = row_stride * dtype_size(dtype);
  p->labels = labels;
  p->vocab_start = vocab_start;
  p->inv_temperature = static_cast<float>(1.0 / temperature);
}

AREAL_B200_API int areal_b200_logits_fwd(const void* logits, int dtype, int64_t T, int64_t V,
                                         int64_t row_stride, const int64_t* labels,
                                         double temperature, float* logp, float* entropy,
                                         float* row_max, float* ln_sumexp, float* raw_min,
                                         void* stream) {
  AB_NVTX_RANGE("areal_b200_logits_fwd");
  int rc = check_logits(logits, dtype, T, V, row_stride, labels, temperature);
  if (rc) return rc;
  AB_REQUIRE(T == 0 || (logp && row_max && ln_sumexp), "logp, row_max and ln_sumexp outputs are required");
  LogitsFwdParams p;
  fill_fwd_common(&p, logits, dtype, T, V, row_stride, labels, 0, temperature);
  p.mode = kModeStandard;
  p.logp = logp;
  p.entropy = entropy;
  p.row_max = row_max;
  p.ln_s = ln_sumexp;
  p.raw_min = raw_min;
  return launch_logits_fwd(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_logits_partial_fwd(const void* logits, int dtype, int64_t T,
                                                 int64_t V, int64_t row_stride,
                                                 const int64_t* labels, int64_t vocab_start,
                                                 double temperature, float* parts, float* raw_min,
                                                 void* stream) {
  AB_NVTX_RANGE("areal_b200_logits_partial_fwd");
  int rc = check_logits(logits, dtype, T, V, row_stride, labels, temperature);
  if (rc) return rc;
  AB_REQUIRE(T == 0 || parts, "parts output is required");
  LogitsFwdParams p;
  fill_fwd_common(&p, logits, dtype, T, V, row_stride, labels, vocab_start, temperature);
  p.mode = kModePartial;
  p.part_m = parts;
  p.part_s = parts + T;
  p.part_w = parts + 2 * T;
  p.part_xl = parts + 3 * T;
  p.raw_min = raw_min;
  return launch_logits_fwd(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_vocab_combine(const float* parts, int64_t R, int64_t T,
                                            double temperature, float* logp, float* entropy,
                                            float* row_max, float* ln_sumexp, void* stream) {
  AB_NVTX_RANGE("areal_b200_vocab_combine");
  AB_REQUIRE(R > 0 && R <= 4096 && T >= 0, "bad R / T");
  AB_REQUIRE(temperature > 0.0 && isfinite(temperature), "temperature must be finite and > 0");
  AB_REQUIRE(T == 0 || (parts && logp && row_max && ln_sumexp), "null pointer");
  return launch_vocab_combine(parts, static_cast<int>(R), T, static_cast<float>(1.0 / temperature),
                              logp, entropy, row_max, ln_sumexp, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_logits_bwd(const void* logits, int dtype, int64_t T, int64_t V,
                                         int64_t row_stride, const int64_t* labels,
                                         int64_t vocab_start, double temperature,
                                         const float* row_max, const float* ln_sumexp,
                                         const float* entropy, const float* g_logp,
                                         const float* g_entropy, const float* scale0,
                                         const float* scale1, void* dlogits, int64_t d_row_stride,
                                         void* stream) {
  AB_NVTX_RANGE("areal_b200_logits_bwd");
  int rc = check_logits(logits, dtype, T, V, row_stride, labels, temperature);
  if (rc) return rc;
  AB_REQUIRE(T == 0 || (row_max && ln_sumexp && dlogits), "row_max, ln_sumexp and dlogits are required");
  AB_REQUIRE(d_row_stride >= V, "d_row_stride (%lld) < V", (long long)d_row_stride);
  AB_REQUIRE(!g_entropy || entropy, "g_entropy needs the forward's entropy");
  LogitsBwdParams p;
  memset(&p, 0, sizeof(p));
  p.logits = logits;
  p.dlogits = dlogits;
  p.dtype = dtype;
  p.T = T;
  p.V = V;
  p.row_stride_bytes = row_stride * dtype_size(dtype);
  p.d_row_stride_bytes = d_row_stride * dtype_size(dtype);
  p.labels = labels;
  p.vocab_start = vocab_start;
  p.inv_temperature = static_cast<float>(1.0 / temperature);
  p.row_max = row_max;
  p.ln_s = ln_sumexp;
  p.entropy = entropy;
  p.g_logp = g_logp;
  p.g_entropy = g_entropy;
  p.scale0 = scale0;
  p.scale1 = scale1;
  return launch_logits_bwd(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_legacy_entropy(const void* logits, int dtype, int64_t T, int64_t V,
                                             int64_t row_stride, const float* row_max,
                                             const float* ln_sumexp, double eps, float* entropy_out,
                                             void* stream) {
  AB_NVTX_RANGE("areal_b200_legacy_entropy");
  AB_REQUIRE(T >= 0 && V > 0, "bad shape T=%lld V=%lld", (long long)T, (long long)V);
  AB_REQUIRE(dtype_size(dtype) != 0, "unsupported dtype %d", dtype);
  AB_REQUIRE(row_stride >= V, "row_stride < V");
  AB_REQUIRE(T == 0 || (logits && row_max && ln_sumexp && entropy_out), "null pointer");
  return launch_legacy_entropy(logits, dtype, T, V, row_stride * dtype_size(dtype), row_max, ln_sumexp,
                               static_cast<float>(eps), entropy_out, static_cast<cudaStream_t>(stream));
}

// ---- P3 ------------------------------------------------------------------------------------------
AREAL_B200_API int64_t areal_b200_loss_workspace_bytes(void) {
  const int64_t a = loss_workspace_bytes(), b = logits_loss_workspace_bytes();
  return a > b ? a : b;
}

static int check_ws(void* workspace, int64_t workspace_bytes) {
  const int64_t need = areal_b200_loss_workspace_bytes();
  if (!workspace || workspace_bytes < need) {
    set_error("loss workspace too small: need %lld bytes, got %lld", (long long)need,
              (long long)workspace_bytes);
    return kWorkspaceTooSmall;
  }
  return kOk;
}

AREAL_B200_API int areal_b200_ppo_actor_loss(const float* logprobs, const float* proximal_logprobs,
                              const float* old_logprobs, const float* advantages,
                              const uint8_t* loss_mask, int64_t T, double eps_clip,
                              double eps_clip_higher, double c_clip, double behav_imp_weight_cap,
                              float* loss_out, float* inv_count_out, float* coef, float* stat_loss,
                              float* stat_importance_weight, float* stat_approx_kl,
                              uint8_t* stat_clip_mask, uint8_t* stat_dual_clip_mask,
                              float* stat_behave_imp_weight, float* stat_behave_approx_kl,
                              uint8_t* stat_behave_mask, void* workspace, int64_t workspace_bytes,
                              void* stream) {
  AB_NVTX_RANGE("areal_b200_ppo_actor_loss");
  AB_REQUIRE(T >= 0, "negative T");
  AB_REQUIRE(loss_out && inv_count_out, "loss_out / inv_count_out required");
  AB_REQUIRE(T == 0 || (logprobs && old_logprobs && advantages && coef), "null pointer");
  int rc = check_ws(workspace, workspace_bytes);
  if (rc) return rc;
  ActorLossParams p;
  memset(&p, 0, sizeof(p));
  rc = make_cfg(&p.cfg, eps_clip, eps_clip_higher, c_clip, behav_imp_weight_cap,
                proximal_logprobs != nullptr);
  if (rc) return rc;
  p.logp = logprobs;
  p.prox = proximal_logprobs;
  p.old = old_logprobs;
  p.adv = advantages;
  p.mask = loss_mask;
  p.T = T;
  p.loss_out = loss_out;
  p.inv_count_out = inv_count_out;
  p.coef = coef;
  p.stat_loss = stat_loss;
  p.stat_ratio = stat_importance_weight;
  p.stat_kl = stat_approx_kl;
  p.clip_mask = stat_clip_mask;
  p.dual_clip_mask = stat_dual_clip_mask;
  p.behave_w = stat_behave_imp_weight;
  p.behave_kl = stat_behave_approx_kl;
  p.behave_mask = stat_behave_mask;
  p.workspace = workspace;
  return launch_actor_loss(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_ppo_critic_loss(const float* value, const float* old_value,
                               const float* target_value, const uint8_t* loss_mask, int64_t T,
                               double value_eps_clip, int use_huber, float* loss_out,
                               float* inv_count_out, float* coef, float* stat_loss,
                               uint8_t* stat_clip_mask, void* workspace, int64_t workspace_bytes,
                               void* stream) {
  AB_NVTX_RANGE("areal_b200_ppo_critic_loss");
  AB_REQUIRE(T >= 0, "negative T");
  AB_REQUIRE(loss_out && inv_count_out, "loss_out / inv_count_out required");
  AB_REQUIRE(T == 0 || (value && old_value && target_value && coef), "null pointer");
  int rc = check_ws(workspace, workspace_bytes);
  if (rc) return rc;
  CriticLossParams p;
  memset(&p, 0, sizeof(p));
  p.value = value;
  p.old = old_value;
  p.target = target_value;
  p.mask = loss_mask;
  p.T = T;
  p.eps = static_cast<float>(value_eps_clip);
  p.huber = use_huber;
  p.loss_out = loss_out;
  p.inv_count_out = inv_count_out;
  p.coef = coef;
  p.stat_loss = stat_loss;
  p.clip_mask = stat_clip_mask;
  p.workspace = workspace;
  return launch_critic_loss(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_scale_by_scalars(const float* coef, const float* scale0, const float* scale1,
                                int64_t T, float* out, void* stream) {
  AB_NVTX_RANGE("areal_b200_scale_by_scalars");
  AB_REQUIRE(T >= 0, "negative T");
  AB_REQUIRE(T == 0 || (coef && scale0 && out), "null pointer");
  return launch_scale(coef, scale0, scale1, T, out, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_fused_actor_fwd(const void* logits, int dtype, int64_t T, int64_t V,
                               int64_t row_stride, const int64_t* labels, double temperature,
                               const float* proximal_logprobs, const float* old_logprobs,
                               const float* advantages, const uint8_t* loss_mask, double eps_clip,
                               double eps_clip_higher, double c_clip, double behav_imp_weight_cap,
                               float* logp, float* entropy, float* row_max, float* ln_sumexp,
                               float* raw_min, float* loss_out, float* inv_count_out, float* coef,
                               float* stat_loss, float* stat_importance_weight,
                               float* stat_approx_kl, uint8_t* stat_clip_mask,
                               uint8_t* stat_dual_clip_mask, float* stat_behave_imp_weight,
                               float* stat_behave_approx_kl, uint8_t* stat_behave_mask,
                               void* workspace, int64_t workspace_bytes, void* stream) {
  AB_NVTX_RANGE("areal_b200_fused_actor_fwd");
  int rc = check_logits(logits, dtype, T, V, row_stride, labels, temperature);
  if (rc) return rc;
  AB_REQUIRE(T > 0, "fused entry needs T > 0");
  AB_REQUIRE(logp && row_max && ln_sumexp && loss_out && inv_count_out && coef, "required output is null");
  AB_REQUIRE(old_logprobs && advantages && loss_mask, "old_logprobs / advantages / loss_mask required");
  rc = check_ws(workspace, workspace_bytes);
  if (rc) return rc;
  LogitsFwdParams p;
  fill_fwd_common(&p, logits, dtype, T, V, row_stride, labels, 0, temperature);
  rc = make_cfg(&p.cfg, eps_clip, eps_clip_higher, c_clip, behav_imp_weight_cap,
                proximal_logprobs != nullptr);
  if (rc) return rc;
  p.mode = kModeFusedActor;
  p.logp = logp;
  p.entropy = entropy;
  p.row_max = row_max;
  p.ln_s = ln_sumexp;
  p.raw_min = raw_min;
  p.prox = proximal_logprobs;
  p.old = old_logprobs;
  p.adv = advantages;
  p.loss_mask = loss_mask;
  p.coef = coef;
  p.stat_loss = stat_loss;
  p.stat_ratio = stat_importance_weight;
  p.stat_kl = stat_approx_kl;
  p.clip_mask = stat_clip_mask;
  p.dual_clip_mask = stat_dual_clip_mask;
  p.behave_w = stat_behave_imp_weight;
  p.behave_kl = stat_behave_approx_kl;
  p.behave_mask = stat_behave_mask;
  p.loss_out = loss_out;
  p.inv_count_out = inv_count_out;
  p.workspace = workspace;
  return launch_logits_fwd(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_fused_actor_fwd_bwd(const void* logits, int dtype, int64_t T, int64_t V,
                               int64_t row_stride, const int64_t* labels, double temperature,
                               const float* proximal_logprobs, const float* old_logprobs,
                               const float* advantages, const uint8_t* loss_mask, double eps_clip,
                               double eps_clip_higher, double c_clip, double behav_imp_weight_cap,
                               float* logp, float* entropy, float* row_max, float* ln_sumexp,
                               float* raw_min, float* loss_out, float* inv_count_out, float* coef,
                               float* stat_loss, float* stat_importance_weight,
                               float* stat_approx_kl, uint8_t* stat_clip_mask,
                               uint8_t* stat_dual_clip_mask, float* stat_behave_imp_weight,
                               float* stat_behave_approx_kl, uint8_t* stat_behave_mask,
                               const float* loss_scale, void* dlogits, int64_t d_row_stride,
                               int cluster_size, int n_buffers, void* workspace,
                               int64_t workspace_bytes, void* stream) {
  AB_NVTX_RANGE("areal_b200_fused_actor_fwd_bwd");
  int rc = check_logits(logits, dtype, T, V, row_stride, labels, temperature);
  if (rc) return rc;
  AB_REQUIRE(T > 0, "fused entry needs T > 0");
  AB_REQUIRE(dlogits && d_row_stride >= V, "dlogits / d_row_stride invalid");
  AB_REQUIRE(logp && row_max && ln_sumexp && loss_out && inv_count_out && coef, "required output is null");
  AB_REQUIRE(old_logprobs && advantages && loss_mask, "old_logprobs / advantages / loss_mask required");
  rc = check_ws(workspace, workspace_bytes);
  if (rc) return rc;
  LogitsFwdParams p;
  fill_fwd_common(&p, logits, dtype, T, V, row_stride, labels, 0, temperature);
  rc = make_cfg(&p.cfg, eps_clip, eps_clip_higher, c_clip, behav_imp_weight_cap,
                proximal_logprobs != nullptr);
  if (rc) return rc;
  p.mode = kModeFusedActor;
  p.logp = logp;
  p.entropy = entropy;
  p.row_max = row_max;
  p.ln_s = ln_sumexp;
  p.raw_min = raw_min;
  p.prox = proximal_logprobs;
  p.old = old_logprobs;
  p.adv = advantages;
  p.loss_mask = loss_mask;
  p.coef = coef;
  p.stat_loss = stat_loss;
  p.stat_ratio = stat_importance_weight;
  p.stat_kl = stat_approx_kl;
  p.clip_mask = stat_clip_mask;
  p.dual_clip_mask = stat_dual_clip_mask;
  p.behave_w = stat_behave_imp_weight;
  p.behave_kl = stat_behave_approx_kl;
  p.behave_mask = stat_behave_mask;
  p.loss_out = loss_out;
  p.inv_count_out = inv_count_out;
  p.workspace = workspace;
  return launch_fused_rowres(p, dlogits, d_row_stride * dtype_size(dtype), loss_scale, cluster_size, n_buffers,
                             static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_rescale_grad_if_needed(void* dlogits, int64_t T, int64_t V,
                                                     int64_t d_row_stride, const float* upstream,
                                                     const float* assumed, void* stream) {
  AB_NVTX_RANGE("areal_b200_rescale_grad_if_needed");
  AB_REQUIRE(T >= 0 && V > 0 && d_row_stride >= V, "bad shape");
  AB_REQUIRE(T == 0 || (dlogits && upstream), "null pointer");
  return launch_rescale_if_needed(dlogits, T, V, d_row_stride * 2, upstream, assumed,
                                  static_cast<cudaStream_t>(stream));
}

// ---- next row: normalisation -------------------------------------------------------------------------
AREAL_B200_API int64_t areal_b200_norm_num_partials(int64_t group_elems) {
  return norm_num_partials(group_elems < 0 ? 0 : group_elems);
}

AREAL_B200_API int areal_b200_masked_moments(const float* x, const float* mask, int64_t n_groups,
                                             int64_t group_elems, int64_t n_partials, double* partials,
                                             void* stream) {
  AB_NVTX_RANGE("areal_b200_masked_moments");
  AB_REQUIRE(n_groups >= 0 && group_elems >= 0 && n_partials >= 1 && n_partials <= 65535, "bad sizes");
  AB_REQUIRE(n_groups <= 65535, "at most 65535 groups per call");
  if (n_groups == 0 || group_elems == 0) return kOk;
  AB_REQUIRE(x && partials, "null pointer");
  return launch_moments(x, mask, n_groups, group_elems, n_partials, partials, static_cast<cudaStream_t>(stream));
}

static int fill_mean_spec(NormMeanSpec* ms, int mean_mode, int leave_one_out, int mean_zero_special,
                          int64_t mean_group_elems, int64_t n_mean_groups, const double* mean_stats) {
  AB_REQUIRE(mean_mode >= 0 && mean_mode <= 2, "mean_mode must be 0 (batch), 1 (group) or 2 (none)");
  AB_REQUIRE(mean_mode == kNormNone || (mean_group_elems > 0 && mean_stats), "mean statistics required");
  ms->mode = mean_mode;
  ms->leave_one_out = leave_one_out;
  ms->zero_special = mean_zero_special;
  ms->group_elems = mean_group_elems > 0 ? mean_group_elems : 1;
  ms->n_groups = n_mean_groups;
  ms->stats = mean_stats;
  return kOk;
}

AREAL_B200_API int areal_b200_norm_centered_sumsq(const float* x, const float* mask, int mean_mode,
                                                  int leave_one_out, int mean_zero_special,
                                                  int64_t mean_group_elems, int64_t n_mean_groups,
                                                  const double* mean_stats, int64_t n_std_groups,
                                                  int64_t std_group_elems, int64_t n_partials,
                                                  double* partials, void* stream) {
  AB_NVTX_RANGE("areal_b200_norm_centered_sumsq");
  AB_REQUIRE(n_std_groups >= 0 && n_std_groups <= 65535 && std_group_elems >= 0, "bad std grouping");
  AB_REQUIRE(n_partials >= 1 && n_partials <= 65535, "bad n_partials");
  if (n_std_groups == 0 || std_group_elems == 0) return kOk;
  AB_REQUIRE(x && partials, "null pointer");
  NormMeanSpec ms;
  const int rc = fill_mean_spec(&ms, mean_mode, leave_one_out, mean_zero_special, mean_group_elems, n_mean_groups,
                                mean_stats);
  if (rc) return rc;
  return launch_centered_sumsq(x, mask, ms, n_std_groups, std_group_elems, n_partials, partials,
                               static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_norm_apply(const float* x, const float* mask, int64_t N, int mean_mode,
                                         int leave_one_out, int mean_zero_special, int64_t mean_group_elems,
                                         int64_t n_mean_groups, const double* mean_stats, int std_mode,
                                         int std_unbiased, int std_one_special, int64_t std_group_elems,
                                         int64_t n_std_groups, const double* std_stats, double eps,
                                         const float* total_mask_sum, float* out, void* stream) {
  AB_NVTX_RANGE("areal_b200_norm_apply");
  AB_REQUIRE(N >= 0, "negative N");
  if (N == 0) return kOk;
  AB_REQUIRE(x && out, "null pointer");
  AB_REQUIRE(std_mode >= 0 && std_mode <= 2, "std_mode must be 0 (batch), 1 (group) or 2 (none)");
  AB_REQUIRE(std_mode == kNormNone || (std_group_elems > 0 && std_stats), "std statistics required");
  NormMeanSpec ms;
  const int rc = fill_mean_spec(&ms, mean_mode, leave_one_out, mean_zero_special, mean_group_elems, n_mean_groups,
                                mean_stats);
  if (rc) return rc;
  NormStdSpec ss;
  ss.mode = std_mode;
  ss.unbiased = std_unbiased;
  ss.one_special = std_one_special;
  ss.group_elems = std_group_elems > 0 ? std_group_elems : 1;
  ss.n_groups = n_std_groups;
  ss.stats = std_stats;
  ss.eps = eps;
  return launch_norm_apply(x, mask, N, ms, ss, total_mask_sum, out, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_masked_normalization_apply(const float* x, const float* mask, int64_t N,
                                                         const double* stats, int unbiased, double eps,
                                                         float* out, void* stream) {
  AB_NVTX_RANGE("areal_b200_masked_normalization_apply");
  AB_REQUIRE(N >= 0, "negative N");
  if (N == 0) return kOk;
  AB_REQUIRE(x && stats && out, "null pointer");
  return launch_masked_normalization_apply(x, mask, N, stats, unbiased, eps, out, static_cast<cudaStream_t>(stream));
}

// ---- P4 ------------------------------------------------------------------------------------------
AREAL_B200_API int64_t areal_b200_interval_workspace_bytes(int64_t n_intervals) {
  return interval_workspace_bytes(n_intervals < 0 ? 0 : n_intervals);
}

static int interval_common(bool slice, const void* flat_in, void* flat_out, const void* packed_in,
                           void* packed_out, int64_t flat_numel, int64_t packed_numel,
                           const int64_t* intervals, int64_t n, int64_t elem_size, void* workspace,
                           int64_t workspace_bytes, void* stream) {
  AB_REQUIRE(n >= 0 && flat_numel >= 0 && packed_numel >= 0, "negative size");
  AB_REQUIRE(elem_size == 1 || elem_size == 2 || elem_size == 4 || elem_size == 8,
             "elem_size must be 1, 2, 4 or 8 (got %lld)", (long long)elem_size);
  const int64_t need = interval_workspace_bytes(n);
  if (!workspace || workspace_bytes < need) {
    set_error("interval workspace too small: need %lld bytes, got %lld", (long long)need,
              (long long)workspace_bytes);
    return kWorkspaceTooSmall;
  }
  AB_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 7) == 0, "workspace must be 8-byte aligned");
  AB_REQUIRE(n == 0 || intervals, "intervals is null");
  AB_REQUIRE(n == 0 || (reinterpret_cast<uintptr_t>(intervals) & 15) == 0,
             "intervals must be 16-byte aligned");
  AB_REQUIRE(packed_numel == 0 || ((slice ? (flat_in && packed_out) : (packed_in && flat_out))),
             "null buffer");
  return launch_interval_copy(slice, flat_in, flat_out, packed_in, packed_out, intervals, n,
                              elem_size, flat_numel, packed_numel, workspace,
                              static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_slice_intervals(const void* flat, int64_t flat_numel, void* packed,
                               int64_t packed_numel, const int64_t* intervals,
                               int64_t n_intervals, int64_t elem_size, void* workspace,
                               int64_t workspace_bytes, void* stream) {
  AB_NVTX_RANGE("areal_b200_slice_intervals");
  return interval_common(true, flat, nullptr, nullptr, packed, flat_numel, packed_numel, intervals,
                         n_intervals, elem_size, workspace, workspace_bytes, stream);
}

AREAL_B200_API int areal_b200_set_intervals(const void* packed, int64_t packed_numel, void* flat,
                             int64_t flat_numel, const int64_t* intervals, int64_t n_intervals,
                             int64_t elem_size, void* workspace, int64_t workspace_bytes,
                             void* stream) {
  AB_NVTX_RANGE("areal_b200_set_intervals");
  return interval_common(false, nullptr, flat, packed, nullptr, flat_numel, packed_numel, intervals,
                         n_intervals, elem_size, workspace, workspace_bytes, stream);
}

AREAL_B200_API int64_t areal_b200_merge_intervals(const uint64_t* iv, int64_t n, uint64_t* out) {
  if (n < 0 || (n > 0 && (!iv || !out))) return -1;
  int64_t k = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (k > 0 && out[2 * (k - 1) + 1] == iv[2 * i]) {
      out[2 * (k - 1) + 1] = iv[2 * i + 1];
    } else {
      out[2 * k] = iv[2 * i];
      out[2 * k + 1] = iv[2 * i + 1];
      ++k;
    }
  }
  return k;
}

static int clamped_params(ClampedParams& p, const void* logits, int dtype, int64_t T, int64_t V, int64_t row_stride,
                          const int64_t* labels, double temperature) {
  AB_REQUIRE(T >= 0 && V > 0, "bad shape");
  AB_REQUIRE(dtype == kBF16 || dtype == kFP16, "clamped entropy: bf16 or fp16 logits only (16-bit radix select)");
  AB_REQUIRE(row_stride >= V, "row stride smaller than V");
  AB_REQUIRE(V < (int64_t(1) << 24), "clamped entropy: V too large for the 16-bit per-thread histograms");
  AB_REQUIRE(temperature > 0.0, "temperature must be positive");
  AB_REQUIRE(T == 0 || (logits && labels), "null pointer");
  p.logits = logits;
  p.dtype = dtype;
  p.T = T;
  p.V = V;
  p.row_stride_bytes = row_stride * 2;
  p.labels = labels;
  p.inv_temperature = static_cast<float>(1.0 / temperature);
  return kOk;
}

AREAL_B200_API int areal_b200_clamped_entropy_fwd(const void* logits, int dtype, int64_t T, int64_t V,
                                                  int64_t row_stride, const int64_t* labels,
                                                  double temperature, int64_t k_removed, float* logp,
                                                  float* entropy, float* saved, void* stream) {
  AB_NVTX_RANGE("areal_b200_clamped_entropy_fwd");
  ClampedParams p;
  const int rc = clamped_params(p, logits, dtype, T, V, row_stride, labels, temperature);
  if (rc) return rc;
  AB_REQUIRE(k_removed >= 0 && k_removed < V, "clamped entropy: k_removed=%lld must be in [0, V)",
             static_cast<long long>(k_removed));
  AB_REQUIRE(T == 0 || (logp && entropy && saved), "null output");
  p.k = k_removed;
  p.logp = logp;
  p.entropy = entropy;
  p.saved = saved;
  return launch_clamped_entropy_fwd(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int64_t areal_b200_lm_head_workspace_bytes(int64_t T, int64_t V) {
  return lm_head_tc_workspace_bytes(T, V);
}

AREAL_B200_API int areal_b200_lm_head_logprobs(const void* hidden, const void* weight, const int64_t* labels,
                                               int64_t T, int64_t H, int64_t V, int64_t hidden_row_stride,
                                               int64_t weight_row_stride, double temperature, float* logp,
                                               float* entropy, float* row_max, float* ln_sumexp, float* raw_min,
                                               void* workspace, int64_t workspace_bytes, void* stream) {
  AB_NVTX_RANGE("areal_b200_lm_head_logprobs");
  AB_REQUIRE(T >= 0 && H > 0 && V > 0, "lm_head: T=%lld H=%lld V=%lld", (long long)T, (long long)H, (long long)V);
  AB_REQUIRE(temperature > 0.0, "temperature must be positive");
  AB_REQUIRE(T == 0 || (hidden && weight && labels && logp && row_max && ln_sumexp), "lm_head: null pointer");
  return launch_lm_head_tc(hidden, weight, labels, T, H, V, hidden_row_stride, weight_row_stride,
                           static_cast<float>(1.0 / temperature), logp, entropy, row_max, ln_sumexp, raw_min, workspace,
                           workspace_bytes, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_clamped_entropy_bwd(const void* logits, int dtype, int64_t T, int64_t V,
                                                  int64_t row_stride, const int64_t* labels,
                                                  double temperature, const float* saved,
                                                  const float* g_logp, const float* g_entropy,
                                                  void* dlogits, int64_t d_row_stride, void* stream) {
  AB_NVTX_RANGE("areal_b200_clamped_entropy_bwd");
  ClampedParams p;
  const int rc = clamped_params(p, logits, dtype, T, V, row_stride, labels, temperature);
  if (rc) return rc;
  AB_REQUIRE(d_row_stride >= V, "gradient row stride smaller than V");
  AB_REQUIRE(T == 0 || (saved && dlogits), "null pointer");
  p.k = 0;
  p.logp = nullptr;
  p.entropy = nullptr;
  p.saved = const_cast<float*>(saved);
  return launch_clamped_entropy_bwd(p, g_logp, g_entropy, dlogits, d_row_stride * 2, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int64_t areal_b200_vocab_p2p_buffer_bytes(int64_t max_tokens) {
  if (max_tokens < 0) return -1;
  return 256 + 2 * 4 * max_tokens * int64_t(sizeof(float));
}

AREAL_B200_API int areal_b200_vocab_combine_p2p(void* const* peer_bases, int rank, int world_size,
                                                int64_t T, int64_t max_tokens, int64_t epoch,
                                                double temperature, int64_t timeout_cycles, float* logp,
                                                float* entropy, float* row_max, float* ln_sumexp,
                                                int* status, void* stream) {
  AB_NVTX_RANGE("areal_b200_vocab_combine_p2p");
  AB_REQUIRE(world_size >= 1 && world_size <= 16, "vocab_combine_p2p: world size must be 1..16, got %d", world_size);
  AB_REQUIRE(rank >= 0 && rank < world_size, "vocab_combine_p2p: rank %d out of range", rank);
  AB_REQUIRE(T >= 0 && T <= max_tokens, "vocab_combine_p2p: T=%lld exceeds the buffer's %lld tokens",
             static_cast<long long>(T), static_cast<long long>(max_tokens));
  AB_REQUIRE(epoch >= 1, "vocab_combine_p2p: epochs start at 1");
  AB_REQUIRE(temperature > 0.0, "temperature must be positive");
  AB_REQUIRE(peer_bases && status, "null pointer");
  AB_REQUIRE(T == 0 || (logp && row_max && ln_sumexp), "null output");
  for (int r = 0; r < world_size; ++r) AB_REQUIRE(peer_bases[r] != nullptr, "vocab_combine_p2p: null peer buffer %d", r);
  return launch_vocab_combine_p2p(peer_bases, rank, world_size, T, max_tokens, static_cast<int>(epoch & 1),
                                  static_cast<unsigned int>(epoch), static_cast<float>(1.0 / temperature),
                                  static_cast<long long>(timeout_cycles), logp, entropy, row_max, ln_sumexp, status,
                                  static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int64_t areal_b200_vocab_push_buffer_bytes(int64_t max_tokens, int world_size) {
  if (max_tokens < 0 || world_size < 1 || world_size > 16) return -1;
  return vocab_push_buffer_bytes(max_tokens, world_size);
}

AREAL_B200_API int areal_b200_logits_partial_fwd_push(const void* logits, int dtype, int64_t T, int64_t V,
                                                      int64_t row_stride, const int64_t* labels,
                                                      int64_t vocab_start, double temperature,
                                                      void* const* peer_bases, int rank, int world_size,
                                                      int64_t max_tokens, int64_t epoch, int64_t rows_target,
                                                      int64_t timeout_cycles, float* logp, float* entropy,
                                                      float* row_max, float* ln_sumexp, float* raw_min,
                                                      int* status, void* stream) {
  AB_NVTX_RANGE("areal_b200_logits_partial_fwd_push");
  int rc = check_logits(logits, dtype, T, V, row_stride, labels, temperature);
  if (rc) return rc;
  AB_REQUIRE(world_size >= 1 && world_size <= 16, "logits_partial_fwd_push: world size must be 1..16, got %d", world_size);
  AB_REQUIRE(rank >= 0 && rank < world_size, "logits_partial_fwd_push: rank %d out of range", rank);
  AB_REQUIRE(T <= max_tokens, "logits_partial_fwd_push: T=%lld exceeds the buffer's %lld tokens",
             static_cast<long long>(T), static_cast<long long>(max_tokens));
  AB_REQUIRE(epoch >= 1, "logits_partial_fwd_push: epochs start at 1");
  AB_REQUIRE(peer_bases && status, "null pointer");
  AB_REQUIRE(T == 0 || (logp && row_max && ln_sumexp), "null output");
  for (int r = 0; r < world_size; ++r) AB_REQUIRE(peer_bases[r] != nullptr, "logits_partial_fwd_push: null peer buffer %d", r);
  LogitsFwdParams p;
  fill_fwd_common(&p, logits, dtype, T, V, row_stride, labels, vocab_start, temperature);
  p.mode = kModePartial;
  p.raw_min = raw_min;
  p.logp = logp;
  p.entropy = entropy;
  p.row_max = row_max;
  p.ln_s = ln_sumexp;
  p.push_world = world_size;
  p.push_rank = rank;
  p.push_slot = static_cast<int>(epoch & 1);
  p.push_tmax = max_tokens;
  p.push_target = static_cast<unsigned int>(static_cast<uint64_t>(rows_target));
  p.push_timeout = static_cast<long long>(timeout_cycles);
  p.push_status = status;
  for (int r = 0; r < world_size; ++r) p.push_base[r] = static_cast<char*>(peer_bases[r]);
  return launch_logits_partial_push(p, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_rms_update(const double* partials, int64_t n_partials, void* state,
                                         int state_is_f64, int mode, double beta, void* stream) {
  AB_NVTX_RANGE("areal_b200_rms_update");
  AB_REQUIRE(partials && state, "null pointer");
  AB_REQUIRE(n_partials >= 1, "rms_update: need at least one partial row");
  AB_REQUIRE(mode == 0 || mode == 1, "rms_update: mode must be 0 (exponential) or 1 (moving average)");
  return launch_rms_update(partials, n_partials, state, state_is_f64, mode, beta, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_rms_apply(const float* x, int64_t N, const void* state, int state_is_f64,
                                        int mode, double eps, int denormalize, float* out, void* stream) {
  AB_NVTX_RANGE("areal_b200_rms_apply");
  AB_REQUIRE(N >= 0, "negative size");
  if (N == 0) return kOk;
  AB_REQUIRE(x && state && out, "null pointer");
  AB_REQUIRE(mode == 0 || mode == 1, "rms_apply: mode must be 0 (exponential) or 1 (moving average)");
  return launch_rms_apply(x, N, state, state_is_f64, mode, eps, denormalize, out, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_packed_rewards(const float* log_probs, const float* ref_log_probs,
                                             const float* score, const int32_t* short1cu_seqlens,
                                             const uint8_t* seq_no_eos, int64_t total_short_tokens,
                                             int64_t batch_size, double kl_ctl, double clip_reward_value,
                                             int mask_no_eos_with_zero, int mode, float* kl_rewards_out,
                                             float* tot_rewards_out, void* stream) {
  AB_NVTX_RANGE("areal_b200_packed_rewards");
  AB_REQUIRE(total_short_tokens >= 0 && batch_size >= 0, "negative size");
  if (total_short_tokens == 0) return kOk;
  AB_REQUIRE(batch_size > 0, "packed_rewards: tokens without sequences");
  AB_REQUIRE(mode >= 0 && mode <= 2, "packed_rewards: mode must be 0 (sparse), 1 (dense) or 2 (dense delta)");
  AB_REQUIRE(log_probs && ref_log_probs && score && short1cu_seqlens && kl_rewards_out && tot_rewards_out,
             "null pointer");
  AB_REQUIRE(!(mode == 0 && mask_no_eos_with_zero) || seq_no_eos, "packed_rewards: seq_no_eos is required");
  return launch_packed_rewards(log_probs, ref_log_probs, score, short1cu_seqlens, seq_no_eos, total_short_tokens,
                               batch_size, static_cast<float>(-kl_ctl), static_cast<float>(clip_reward_value),
                               mask_no_eos_with_zero, mode, kl_rewards_out, tot_rewards_out,
                               static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int areal_b200_segment_masked_mean(const float* v, const uint8_t* mask,
                                                  const int32_t* cu_seqlens, int64_t batch_size,
                                                  double* mean, uint8_t* valid, void* stream) {
  AB_NVTX_RANGE("areal_b200_segment_masked_mean");
  AB_REQUIRE(batch_size >= 0, "negative size");
  if (batch_size == 0) return kOk;
  AB_REQUIRE(v && cu_seqlens && mean && valid, "null pointer");
  return launch_segment_masked_mean(v, mask, cu_seqlens, batch_size, mean, valid, static_cast<cudaStream_t>(stream));
}

AREAL_B200_API int64_t areal_b200_masked_stats_workspace_bytes(int64_t n_keys, int64_t n) {
  if (n_keys < 0 || n < 0) return -1;
  return masked_stats_workspace_bytes(n_keys, n);
}

AREAL_B200_API int areal_b200_masked_stats(const float* const* values, const uint8_t* const* masks,
                                           double* const* outs, uint32_t value_is_u8_bits,
                                           int64_t n_keys, int64_t n, int accumulate, void* workspace,
                                           int64_t workspace_bytes, void* stream) {
  AB_NVTX_RANGE("areal_b200_masked_stats");
  AB_REQUIRE(n_keys >= 0 && n >= 0, "negative size");
  if (n_keys == 0) return kOk;
  AB_REQUIRE(n_keys <= kMaxStatKeys, "masked_stats: at most %d keys per call, got %lld", kMaxStatKeys,
             static_cast<long long>(n_keys));
  AB_REQUIRE(n > 0, "masked_stats: arrays must be non-empty");   // stats_tracker.py:113-114
  AB_REQUIRE(values && masks && outs && workspace, "null pointer");
  if (workspace_bytes < masked_stats_workspace_bytes(n_keys, n)) {
    set_error("masked_stats: workspace of %lld bytes is too small (need %lld)", static_cast<long long>(workspace_bytes),
              static_cast<long long>(masked_stats_workspace_bytes(n_keys, n)));
    return kWorkspaceTooSmall;
  }
  StatsTable tab;
  tab.val_u8_bits = value_is_u8_bits;
  for (int k = 0; k < kMaxStatKeys; ++k) {
    tab.val[k] = k < n_keys ? values[k] : nullptr;
    tab.mask[k] = k < n_keys ? masks[k] : nullptr;
    tab.out[k] = k < n_keys ? outs[k] : nullptr;
    AB_REQUIRE(k >= n_keys || outs[k] != nullptr, "masked_stats: null output slot for key %d", k);
  }
  return launch_masked_stats(tab, static_cast<int>(n_keys), n, accumulate, workspace, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
