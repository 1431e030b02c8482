// Parameter blocks of the fused logits kernels (internal; the public C ABI is include/areal_b200.h).
#pragma once
#include "common.cuh"
#include "ppo_math.cuh"

namespace ab {

enum LogitsMode : int { kModeStandard = 0, kModePartial = 1, kModeFusedActor = 2 };

struct LogitsFwdParams {
  const void* logits;
  int dtype;
  int mode;
  int64_t T, V;
  int64_t row_stride_bytes;
  const int64_t* labels;
  int64_t vocab_start;
  float inv_temperature;
  // standard / fused outputs
  float* logp;
  float* entropy;   // nullable
  float* row_max;   // max of the raw logits (exact)
  float* ln_s;      // ln sum exp((x - row_max)/temp);  logsumexp(z) = row_max/temp + ln_s
  float* raw_min;   // nullable (enables the min pass)
  // partial-mode outputs
  float *part_m, *part_s, *part_w, *part_xl;
  // fused actor-loss epilogue
  const float *prox, *old, *adv;
  const uint8_t* loss_mask;
  ActorCfg cfg;
  float* coef;
  float *stat_loss, *stat_ratio, *stat_kl, *behave_w, *behave_kl;
  uint8_t *clip_mask, *dual_clip_mask, *behave_mask;
  float *loss_out, *inv_count_out;
  void* workspace;
  // partial mode, push exchange (one launch for partial + exchange + combine): see launch_logits_partial_push
  int push_world;               // 0: off
  int push_rank, push_slot;
  int64_t push_tmax;
  unsigned int push_target;     // cumulative number of rows all ranks will have delivered into a buffer's slot
  long long push_timeout;
  int* push_status;
  char* push_base[16];          // every rank's exchange buffer (symmetric memory), this GPU's mapping
};

struct LogitsBwdParams {
  const void* logits;
  void* dlogits;
  int dtype;
  int64_t T, V;
  int64_t row_stride_bytes, d_row_stride_bytes;
  const int64_t* labels;
  int64_t vocab_start;
  float inv_temperature;
  const float* row_max;     // from the forward (global values on the vocab-sharded path)
  const float* ln_s;
  const float* entropy;     // required iff g_entropy
  const float* g_logp;      // nullable
  const float* g_entropy;   // nullable
  const float* scale0;      // nullable device scalars multiplying both gradients
  const float* scale1;
};

int64_t logits_loss_workspace_bytes();
int launch_logits_fwd(const LogitsFwdParams& p, cudaStream_t stream);
int launch_logits_bwd(const LogitsBwdParams& p, cudaStream_t stream);
// logits_rowres.cu: single-pass forward + gradient with cluster-resident rows (mode must be kModeFusedActor)
int launch_fused_rowres(const LogitsFwdParams& p, void* dlogits, int64_t d_row_stride_bytes,
                        const float* loss_scale, int force_cluster, int force_nbuf, cudaStream_t stream);
int launch_rescale_if_needed(void* dlogits, int64_t T, int64_t V, int64_t row_stride_bytes,
                             const float* upstream, const float* assumed, cudaStream_t stream);
int launch_legacy_entropy(const void* logits, int dtype, int64_t T, int64_t V, int64_t row_stride_bytes,
                          const float* row_max, const float* ln_s, float eps, float* out,
                          cudaStream_t stream);
int launch_vocab_combine(const float* parts, int R, int64_t T, float inv_t, float* logp,
                         float* entropy, float* row_max, float* ln_s, cudaStream_t stream);

// partial statistics of a vocabulary shard PUSHED into every rank's exchange buffer from the row epilogue, then --
// same launch -- combined into the full-vocabulary results once every rank's rows are in (p.push_* set, mode partial,
// p.logp / entropy / row_max / ln_s = the combined outputs)
int64_t vocab_push_buffer_bytes(int64_t max_tokens, int world);
int launch_logits_partial_push(const LogitsFwdParams& p, cudaStream_t stream);

int launch_vocab_combine_p2p(void* const* peer_bases, int rank, int R, int64_t T, int64_t Tmax, int slot,
                             unsigned int epoch, float inv_t, long long timeout_cycles, float* logp, float* entropy,
                             float* row_max, float* ln_s, int* status, cudaStream_t stream);

}  // namespace ab
