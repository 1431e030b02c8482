/* areal_b200 -- C ABI of the B200-native RL post-processing kernels (libareal_b200.so).
 *
 * Drop-in boundary for the reference's native extensions and functional ops on this path.  Every
 * entry point takes plain device pointers, sizes and scalars plus a CUDA stream, launches
 * asynchronously on that stream, never allocates and never synchronises; it returns 0 on success
 * or a non-zero status (1 invalid argument, 2 CUDA error, 3 workspace too small, 4 unsupported)
 * with a message available from areal_b200_last_error() (thread-local).
 *
 * "Reference" citations are relative to the root of JamesKrW/AReaL.
 *
 * Workspaces: entry points that take `workspace` need `*_workspace_bytes()` bytes of device
 * memory.  The loss / fused entry points need it ZEROED ONCE after allocation and leave it
 * zeroed; the GAE and interval entry points initialise what they use themselves (stream-ordered).
 * One workspace can be reused by successive calls on the same stream, never by calls that may
 * overlap on the GPU.
 */
#ifndef AREAL_B200_H_
#define AREAL_B200_H_

#include <stdint.h>

#if defined(__GNUC__)
#define AREAL_B200_API __attribute__((visibility("default")))
#else
#define AREAL_B200_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* element types of logits / dlogits */
#define AREAL_B200_BF16 0
#define AREAL_B200_FP16 1
#define AREAL_B200_FP32 2

AREAL_B200_API const char* areal_b200_last_error(void);
AREAL_B200_API int areal_b200_version(void);

/* ------------------------------------------------------------------------------------------ *
 * P1  GAE
 * ------------------------------------------------------------------------------------------ */

/* Packed 1D "nolp/misalign" GAE.  Replaces `gae_1d_nolp_misalign` (csrc/cugae/gae.cu:30-60,
 * pybind :210), `cugae.cugae1d_nolp_misalign_func` (realhf/impl/model/utils/ppo_functional.py:
 * 359-366) and the CPU path `pygae1d_nolp_misalign` (:292-323).
 *   rewards f32[T], values f32[T+B] (value of token t of sequence i at cu[i]+i+t), cu_seqlens
 *   i32[B+1] (cu[0]=0, cu[B]=T), bootstrap u8/bool[B]; outputs adv, ret f32[T].
 *   gamma, lam are taken as double and combined like the Python reference: (float)gamma and
 *   (float)(gamma*lam).  Inputs are not modified. */
AREAL_B200_API int64_t areal_b200_gae1d_workspace_bytes(int64_t total_tokens);
AREAL_B200_API int areal_b200_gae1d_nolp_misalign(const float* rewards, const float* values,
                                   const int32_t* cu_seqlens, const uint8_t* bootstrap,
                                   int64_t total_tokens, int64_t batch_size, double gamma,
                                   double lam, float* adv_out, float* ret_out, void* workspace,
                                   int64_t workspace_bytes, void* stream);

/* Allocation-free variant of areal_b200_gae1d_nolp_misalign.  It has NO counterpart in the
 * reference; it exists for callers that cannot obtain device scratch themselves: a C/C++ or other
 * non-PyTorch host that binds this ABI directly and so has no caching allocator to draw from, a
 * process whose framework has already reserved all of HBM (cudaMalloc would fail or fragment the
 * pool), and latency-critical threads that must not pay the device-wide synchronisation a
 * cudaMalloc / cudaFree implies.  Same inputs, outputs and results as the entry above, but the
 * look-back descriptors live in a 65,600-byte array inside the library image (each GPU that loads
 * the library keeps it until the process exits) instead of a caller workspace; the call itself
 * performs no cudaMalloc, cudaFree or any other allocating call.
 *   Limit: total_tokens <= AREAL_B200_GAE1D_STATIC_MAX_TOKENS per call; larger batches are
 *   rejected with status 1 and a message (use the workspace entry for those).
 *   Set-up: the first call on each device (made with that device current) resolves the array's
 *   device address and caches it; make that first call before starting any stream capture.
 *   Sharing: all calls on a device use the same array.  It keeps the previous call's contents, so
 *   each call first clears (stream-ordered) the part it will use.  Two calls that can overlap on
 *   the GPU -- different streams, or different host threads -- race on it: serialise them. */
#define AREAL_B200_GAE1D_STATIC_MAX_TOKENS 4194304
AREAL_B200_API int areal_b200_gae1d_nolp_misalign_static(const float* rewards, const float* values,
                                                         const int32_t* cu_seqlens,
                                                         const uint8_t* bootstrap,
                                                         int64_t total_tokens, int64_t batch_size,
                                                         double gamma, double lam, float* adv_out,
                                                         float* ret_out, void* stream);

/* Padded [B,S] masked GAE of areal-lite.  Replaces the Python loop in
 * `PPOActor.compute_advantages` (areal/engine/ppo/actor.py:131-152).
 *   rewards, values, loss_mask f32[B,S] contiguous (loss_mask already rolled by -1, :110; any
 *   non-zero counts), seq_no_eos u8/bool[B]; outputs advantages, returns f32[B,S]. */
AREAL_B200_API int areal_b200_gae_padded_masked(const float* rewards, const float* values,
                                 const float* loss_mask, const uint8_t* seq_no_eos, int64_t B,
                                 int64_t S, double gamma, double lam, float* adv_out,
                                 float* ret_out, void* stream);

/* Reward construction + padded masked GAE in ONE launch: everything `PPOActor.compute_advantages`
 * does between reward scaling and `returns` (areal/engine/ppo/actor.py:96-152; KL estimators k1/k2/k3
 * of areal/utils/data.py:1343-1363).  ~25 torch kernels and the S-step Python loop of the reference.
 *   score f32[B] (data["rewards"]; with overlong=1 the length penalty of
 *   areal/utils/functional.py:376-398 is added to it IN PLACE first); loss_mask [B,S] NOT rolled, element code 0 f32 / 1 u8,bool /
 *   2 i32 / 3 i64; logp f32[B,S] (roll_logp=1: rolled by -1 inside, :116; 0: used as is, :114);
 *   ref_logp f32[B,S] or NULL (zeros, :120) -- multiplied by the rolled mask IN PLACE like :121;
 *   attn_mask [B,S] with its own element code (seqlens = row sums, :126); values f32[B,S] or NULL
 *   (zeros, :141).  apply_scaling: score = clip((score+bias)*scaling, +-clip) (:98-102); pass 0 when
 *   the caller already scaled (e.g. a reward_norm in between, :103).  kl_kind 1|2|3, kl_clamp 0|1.
 *   Outputs f32[B,S]: advantages (before adv_norm), returns, kl_rewards, tot_rewards, loss_mask_out
 *   (rolled, as float), old_logp_out (rolled and masked). */
AREAL_B200_API int areal_b200_advantages_padded(
    float* score, int overlong, int64_t overlong_tokens, int64_t max_response_length,
    double overlong_penalty_factor, const void* loss_mask, int loss_mask_code, const float* logp, int roll_logp,
    float* ref_logp, const void* attn_mask, int attn_code, const float* values, int64_t B, int64_t S,
    int apply_scaling, double reward_bias, double reward_scaling, double reward_clip, double kl_ctl,
    int kl_kind, int kl_clamp, int mask_no_eos_with_zero, double gamma, double lam, float* adv_out,
    float* ret_out, float* kl_rewards_out, float* tot_rewards_out, float* loss_mask_out,
    float* old_logp_out, void* stream);

/* Segment-table 2D GAE.  Replaces `gae_2d_olp` AND `gae_2d_nolp` (csrc/cugae/gae.cu:91-134,
 * 163-206, pybind :213-215): the reference's `_nolp` host wrapper launches the `_olp` kernel
 * (:199), so both exports have olp semantics.
 *   rewards f32[B,H], values f32[B,H+1], done u8[B,H+1], done_y_indices i32[num_dones] (ascending
 *   per row), cu_num_dones i32[B+1], is_truncate u8[num_dones]; outputs f32[B,H]. */
AREAL_B200_API int areal_b200_gae2d_olp(const float* rewards, const float* values, const uint8_t* done,
                         const int32_t* done_y_indices, const int32_t* cu_num_dones,
                         const uint8_t* is_truncate, int64_t B, int64_t H, int64_t num_dones,
                         double gamma, double lam, float* adv_out, float* ret_out, void* stream);

/* ------------------------------------------------------------------------------------------ *
 * P2  log-prob / entropy from [T,V] logits, and the dlogits backward
 * ------------------------------------------------------------------------------------------ */

/* Forward.  Replaces `gather_logprobs` / `gather_logprobs_entropy`
 * (areal/utils/functional.py:43-127; inner :20-34) and the two extra [T,V] reads
 * `logits.min(-1)`, `logits.max(-1)` (areal/engine/ppo/actor.py:375-376).
 *   logits [T,V] of `dtype`, unit stride in V, `row_stride` ELEMENTS between rows; labels i64[T];
 *   z = float(logits)/temperature.  Outputs f32[T]: logp (required); row_max = max_v logits (raw
 *   units, exact) and ln_sumexp = ln sum_v exp((logits - row_max)/temperature) (both required and
 *   saved for the backward: logsumexp(z) = row_max/temperature + ln_sumexp, kept as two terms so
 *   that 1 - p_label survives fp32 on peaked rows); entropy, raw_min nullable.
 *   A label outside [0,V) yields logp = NaN for that row (the reference's gather would raise). */
AREAL_B200_API int areal_b200_logits_fwd(const void* logits, int dtype, int64_t T, int64_t V,
                                         int64_t row_stride, const int64_t* labels,
                                         double temperature, float* logp, float* entropy,
                                         float* row_max, float* ln_sumexp, float* raw_min,
                                         void* stream);

/* Backward: dlogits = d/dlogits of sum_t g_logp[t]*logp[t] + g_entropy[t]*entropy[t], times the
 * optional device scalars *scale0 * *scale1, rounded once to `dtype`.  Replaces autograd through
 * the reference's log_softmax path (functional.py:20-34) and `_VocabParallelEntropy.backward`
 * (areal/utils/mcore/functional.py:41-52).  g_entropy nullable (then entropy may be null); rows
 * whose gradients are all zero are written as zeros without reading the logits.  On the
 * vocab-sharded path pass the shard's logits, its vocab_start and the GLOBAL row_max / ln_sumexp. */
AREAL_B200_API int areal_b200_logits_bwd(const void* logits, int dtype, int64_t T, int64_t V,
                                         int64_t row_stride, const int64_t* labels,
                                         int64_t vocab_start, double temperature,
                                         const float* row_max, const float* ln_sumexp,
                                         const float* entropy, const float* g_logp,
                                         const float* g_entropy, const float* scale0,
                                         const float* scale1, void* dlogits, int64_t d_row_stride,
                                         void* stream);

/* Vocab-sharded forward, per shard: the partials that the reference obtains with three
 * all-reduces each in `_VocabParallelEntropy.forward` (areal/utils/mcore/functional.py:15-34) and
 * Megatron's `vocab_parallel_cross_entropy` (call sites areal/utils/functional.py:55-57, 95-97;
 * vendored copy realhf/impl/model/parallelism/tensor_parallel/modules.py:1060-1118).
 *   parts f32[4,T] = rows (m = shard max of the raw logits, S = sum 2^t, W = sum 2^t * t with
 *   t = (x - m) * log2(e)/temperature, raw label logit or 0 if the label is not in the shard);
 *   labels are GLOBAL vocabulary ids, this shard holds [vocab_start, vocab_start+V). */
AREAL_B200_API int areal_b200_logits_partial_fwd(const void* logits, int dtype, int64_t T,
                                                 int64_t V, int64_t row_stride,
                                                 const int64_t* labels, int64_t vocab_start,
                                                 double temperature, float* parts, float* raw_min,
                                                 void* stream);

/* Combine R shards' partials (parts f32[R,4,T], e.g. the result of one all-gather) into logp,
 * entropy (nullable) and the global row_max / ln_sumexp, f32[T] each. */
AREAL_B200_API int areal_b200_vocab_combine(const float* parts, int64_t R, int64_t T,
                                            double temperature, float* logp, float* entropy,
                                            float* row_max, float* ln_sumexp, void* stream);

/* Legacy entropy statistic -sum_v p_v * log(p_v + eps), p = softmax(logits) (temperature already
 * applied by the caller).  Replaces `calc_entropy` (realhf/impl/model/interface/ppo_interface.py:
 * 52-58, eps = 1e-7), a no-grad logging value of the legacy actor interface; second read-only pass
 * that takes row_max / ln_sumexp from areal_b200_logits_fwd (called with temperature 1). */
AREAL_B200_API int areal_b200_legacy_entropy(const void* logits, int dtype, int64_t T, int64_t V,
                                             int64_t row_stride, const float* row_max,
                                             const float* ln_sumexp, double eps, float* entropy_out,
                                             void* stream);

/* ------------------------------------------------------------------------------------------ *
 * P3  PPO losses
 * ------------------------------------------------------------------------------------------ */

/* Shared by actor / critic / fused entry points: 16 + 8 * (#SM * 8) bytes. */
AREAL_B200_API int64_t areal_b200_loss_workspace_bytes(void);

/* Decoupled PPO actor loss.  Replaces `ppo_actor_loss_fn` (areal/utils/functional.py:171-235) and
 * the legacy `actor_loss_fn` (realhf/impl/model/utils/ppo_functional.py:51-149; pass prox = NULL
 * for its proximal_logprobs=None case, loss_mask = NULL for loss_mask=None).
 *   eps_clip_higher, c_clip, behav_imp_weight_cap: NaN means "None".
 *   loss_out[1] = sum(mask ? L : 0) / max(count(mask), 1); inv_count_out[1] = 1/max(count,1);
 *   coef[T] = count * dloss/dlogprobs (closed form incl. torch.max / torch.min tie splitting), so
 *   grad_logprobs = upstream * inv_count * coef.  The eight logging vectors are nullable. */
AREAL_B200_API int areal_b200_ppo_actor_loss(const float* logprobs, const float* proximal_logprobs,
                              const float* old_logprobs, const float* advantages,
                              const uint8_t* loss_mask, int64_t T, double eps_clip,
                              double eps_clip_higher, double c_clip, double behav_imp_weight_cap,
                              float* loss_out, float* inv_count_out, float* coef, float* stat_loss,
                              float* stat_importance_weight, float* stat_approx_kl,
                              uint8_t* stat_clip_mask, uint8_t* stat_dual_clip_mask,
                              float* stat_behave_imp_weight, float* stat_behave_approx_kl,
                              uint8_t* stat_behave_mask, void* workspace, int64_t workspace_bytes,
                              void* stream);

/* PPO critic loss.  Replaces `ppo_critic_loss_fn` (areal/utils/functional.py:247-311) and legacy
 * `critic_loss_fn` (realhf/impl/model/utils/ppo_functional.py:161-225).  loss_mask nullable (plain
 * mean).  No `or 1` on the count, like the reference (:306): an all-false mask gives NaN. */
AREAL_B200_API int areal_b200_ppo_critic_loss(const float* value, const float* old_value,
                               const float* target_value, const uint8_t* loss_mask, int64_t T,
                               double value_eps_clip, int use_huber, float* loss_out,
                               float* inv_count_out, float* coef, float* stat_loss,
                               uint8_t* stat_clip_mask, void* workspace, int64_t workspace_bytes,
                               void* stream);

/* out[t] = coef[t] * (*scale0) * (*scale1)  (scale1 nullable): grad of a masked-mean loss from its
 * saved coefficient, the saved 1/count and the upstream scalar, all on the device. */
AREAL_B200_API int areal_b200_scale_by_scalars(const float* coef, const float* scale0, const float* scale1,
                                int64_t T, float* out, void* stream);

/* Fused GRPO/PPO actor step: ONE pass over the logits producing everything `grpo_loss_fn`
 * (areal/engine/ppo/actor.py:313-391) computes from them: logp, entropy, row_max, ln_sumexp, raw_min, the
 * actor loss, its logging vectors and the gradient coefficient (row_max doubles as logits.max(-1)).  Same argument meaning as
 * areal_b200_logits_fwd + areal_b200_ppo_actor_loss.  Backward = areal_b200_logits_bwd with
 * g_logp = coef, scale0 = inv_count, scale1 = upstream. */
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
                               void* workspace, int64_t workspace_bytes, void* stream);

/* Single-pass variant of the fused actor step: the same outputs as areal_b200_fused_actor_fwd AND
 * dlogits, from ONE read of the logits (4*T*V bytes instead of 6*T*V for forward + backward).  Rows
 * are kept on chip by a thread-block cluster between the statistics pass and the gradient pass.
 * For callers that know, when the forward runs, the scalar that will multiply the loss before
 * .backward() (the engine's loss_scale, areal/engine/fsdp_engine.py:578-586):
 *   dlogits = (*loss_scale) * d loss / d logits   (loss_scale nullable = 1), loss_out is NOT scaled.
 * bf16 logits / dlogits, 16-byte aligned rows, V % 8 == 0; status 4 (unsupported) otherwise -- use
 * the two-pass entry points then.  cluster_size / n_buffers: 0 = choose automatically. */
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
                               int64_t workspace_bytes, void* stream);

/* Companion of the entry above for autograd: multiplies dlogits (bf16, in place) by
 * *upstream / *assumed unless the two device scalars are equal -- the expected case, in which the
 * kernel returns at once. */
AREAL_B200_API int areal_b200_rescale_grad_if_needed(void* dlogits, int64_t T, int64_t V,
                                                     int64_t d_row_stride, const float* upstream,
                                                     const float* assumed, void* stream);

/* ------------------------------------------------------------------------------------------ *
 * Next row: advantage / reward normalisation
 * ------------------------------------------------------------------------------------------ */

/* Replaces `Normalization` (areal/utils/data.py:1073-1304: Python loops over groups, `.item()`
 * syncs, up to 4 scalar all-reduces) and `masked_normalization` (areal/utils/functional.py:130-168)
 * with three launches whatever the number of groups.  x, mask f32[N] (mask nullable = all ones;
 * a float weight, as in the reference); a "group" is `group_elems` consecutive elements.
 *   1. areal_b200_masked_moments: partials f64[n_groups, n_partials, 3] = per-slab
 *      (sum m, sum x*m, sum (x*m)^2); the caller sums over n_partials (fixed order) and, for
 *      batch-level statistics under data parallelism, all-reduces that one small tensor.
 *   2. areal_b200_norm_centered_sumsq: partials f64[n_std_groups, n_partials, 2] =
 *      (sum m, sum (x*m - mean*m)^2) with the per-element mean of `_compute_mean` (:1206-1262).
 *   3. areal_b200_norm_apply: out = ((x - mean) * m) / (std + eps) with `_compute_std`'s rules
 *      (:1264-1304) and the reference's dtype path (batch-level statistics fp64, group-level fp32).
 *   modes: 0 = batch, 1 = group, 2 = none.  *_special flags: the group_size == 1 cases
 *   (:1138-1143, :1180-1185).  total_mask_sum (device scalar, nullable): when it is 0 the input is
 *   returned unchanged (:1115-1116) without a host sync. */
AREAL_B200_API int64_t areal_b200_norm_num_partials(int64_t group_elems);
AREAL_B200_API int areal_b200_masked_moments(const float* x, const float* mask, int64_t n_groups,
                                             int64_t group_elems, int64_t n_partials, double* partials,
                                             void* stream);
AREAL_B200_API int areal_b200_norm_centered_sumsq(const float* x, const float* mask, int mean_mode,
                                                  int leave_one_out, int mean_zero_special,
                                                  int64_t mean_group_elems, int64_t n_mean_groups,
                                                  const double* mean_stats, int64_t n_std_groups,
                                                  int64_t std_group_elems, int64_t n_partials,
                                                  double* partials, void* stream);
AREAL_B200_API int areal_b200_norm_apply(const float* x, const float* mask, int64_t N, int mean_mode,
                                         int leave_one_out, int mean_zero_special, int64_t mean_group_elems,
                                         int64_t n_mean_groups, const double* mean_stats, int std_mode,
                                         int std_unbiased, int std_one_special, int64_t std_group_elems,
                                         int64_t n_std_groups, const double* std_stats, double eps,
                                         const float* total_mask_sum, float* out, void* stream);
/* masked_normalization's apply: stats f64[3] = (sum m, sum x*m, sum (x*m)^2);
 * out = ((x*m - mean) / (sqrt(E[(xm)^2] - mean^2 [* n/(n-1)]) + eps)).float() */
AREAL_B200_API int areal_b200_masked_normalization_apply(const float* x, const float* mask, int64_t N,
                                                         const double* stats, int unbiased, double eps,
                                                         float* out, void* stream);

/* ------------------------------------------------------------------------------------------ *
 * P4  interval ops (parameter reallocation)
 * ------------------------------------------------------------------------------------------ */

/* Scratch for the packed offsets and their scan: (n_intervals + 8 + ceil(n_intervals / 4096)) * 8 bytes; need not be
 * zeroed.  After the call, workspace[0] (int64) holds 0
 * (ok), 1 (an interval is negative / reversed / beyond the flat buffer) or 2 (sum of sizes !=
 * packed_numel); on a non-zero status no byte of the destination is written. */
AREAL_B200_API int64_t areal_b200_interval_workspace_bytes(int64_t n_intervals);

/* packed[cumsum offsets] = flat[s:e) for each interval in the given order.  Replaces
 * `slice_intervals_{fp32,fp16,bf16}` (csrc/interval_op/interval_op.cu:68-104, pybind :110-113) and
 * `_slice_intervals_py` (realhf/impl/model/nn/flatten_param.py:65-75).  intervals i64[n,2] on the
 * device; elem_size in bytes (1, 2, 4 or 8); bit-exact copy for any alignment. */
AREAL_B200_API int areal_b200_slice_intervals(const void* flat, int64_t flat_numel, void* packed,
                               int64_t packed_numel, const int64_t* intervals,
                               int64_t n_intervals, int64_t elem_size, void* workspace,
                               int64_t workspace_bytes, void* stream);

/* flat[s:e) = packed[cumsum offsets] (in place on `flat`).  Replaces
 * `set_intervals_{fp32,fp16,bf16}` (interval_op.cu:28-66, pybind :106-109) and `_set_intervals_py`
 * (flatten_param.py:78-90). */
AREAL_B200_API int areal_b200_set_intervals(const void* packed, int64_t packed_numel, void* flat,
                             int64_t flat_numel, const int64_t* intervals, int64_t n_intervals,
                             int64_t elem_size, void* workspace, int64_t workspace_bytes,
                             void* stream);

/* Host-side.  Replaces `merge_intervals` (csrc/interval_op/interval_op.cpp:4-25): merges (a,b),(b,c)
 * -> (a,c) for neighbours in input order.  in/out are u64[n,2] host arrays (out capacity n);
 * returns the number of merged intervals, or -1 on a null pointer. */
AREAL_B200_API int64_t areal_b200_merge_intervals(const uint64_t* intervals, int64_t n, uint64_t* out);

/* Per-sequence masked mean of a packed vector: mean[i] = sum over cu[i] <= t < cu[i+1] of (mask[t] ? v[t] : 0)
 * divided by the number of selected tokens (0 and valid[i] = 0 when there is none).  Replaces the
 * Python loop over sequences with one `.item()` each in `compute_packed_sft_loss`
 * (areal/engine/sft/lm_engine.py:80-91).  v f32[T], mask u8/bool[T] or NULL, cu_seqlens i32[B+1];
 * mean f64[B], valid u8[B]. */
AREAL_B200_API int areal_b200_segment_masked_mean(const float* v, const uint8_t* mask,
                                                  const int32_t* cu_seqlens, int64_t batch_size,
                                                  double* mean, uint8_t* valid, void* stream);

/* Vocab-sharded combine WITHOUT the all-gather: the partial sets are read straight out of the peers'
 * memory over NVLink (the reference does six all-reduces here: areal/utils/mcore/functional.py:16-31 +
 * Megatron's vocab_parallel_cross_entropy).  No counterpart as a single reference function.
 *   Every rank owns a peer-mapped ("symmetric") buffer of areal_b200_vocab_p2p_buffer_bytes(Tmax)
 *   bytes, ZEROED once: u32 flag[16] (256 B) | parts[2][4][Tmax] f32.  peer_bases: HOST array of R
 *   device pointers, peer_bases[r] = rank r's buffer as mapped into THIS process (peer_bases[rank] is
 *   the local one).  Protocol per step e = 1, 2, 3, ... (the same e on every rank):
 *     1. areal_b200_logits_partial_fwd writes this rank's parts [4,T] at
 *        (char*)local + 256 + (e & 1) * 16 * Tmax   (same stream);
 *     2. this call: release-stores e into flag[rank] of every peer, waits (at most timeout_cycles SM
 *        clocks) until all R flags here are >= e, then combines in rank order -- bit-identical to
 *        areal_b200_vocab_combine on the gathered parts.
 *   *status (device int, zeroed by the caller) is set to 1 if a peer did not arrive in time; the
 *   outputs are then left untouched.  T may be 0 (the rank still takes part in the handshake). */
AREAL_B200_API int64_t areal_b200_vocab_p2p_buffer_bytes(int64_t max_tokens);
AREAL_B200_API int areal_b200_vocab_combine_p2p(void* const* peer_bases, int rank, int world_size,
                                                int64_t T, int64_t max_tokens, int64_t epoch,
                                                double temperature, int64_t timeout_cycles, float* logp,
                                                float* entropy, float* row_max, float* ln_sumexp,
                                                int* status, void* stream);

/* Vocab-sharded forward as ONE launch: partial statistics of this rank's shard, their exchange over NVLink and the
 * combine (replaces, like the pair above, the six all-reduces of areal/utils/mcore/functional.py:16-31 + Megatron's
 * vocab_parallel_cross_entropy; arguments of the shard as in areal_b200_logits_partial_fwd).
 *   Every rank owns a peer-mapped ("symmetric") buffer of areal_b200_vocab_push_buffer_bytes(Tmax, R) bytes, ZEROED
 *   once: u32 rows_in[2] | u32 pad[2] | u32 ticket[2] | pad to 256 B | float4 parts[2][R][Tmax].  peer_bases as above.
 *   The row epilogue of the streaming kernel stores a row's (max, S, W, x_label) into every rank's buffer (plain
 *   16-byte stores, slot e & 1); a CTA that has finished its rows fences and takes a local ticket, the rank's last CTA
 *   adds T to rows_in[e & 1] of every rank (system-scope release), and every CTA
 *   waits (at most timeout_cycles SM clocks) until its own counter has reached rows_target -- the CUMULATIVE number
 *   of rows delivered to that slot over all steps so far, i.e. the caller adds R*T for every step that used the
 *   slot (mod 2^32) -- and then combines its share of the rows from local memory, in rank order: bit-identical to
 *   areal_b200_vocab_combine on the gathered parts.  All ranks call with the same T and e = 1, 2, 3, ...
 *   raw_min (f32[T], may be NULL): row minimum of THIS shard, from the same pass.
 *   *status (device int, zeroed by the caller) is set to 1 if some rank's rows did not arrive in time; the outputs
 *   are then left untouched.  Rows must be 16-byte aligned (base, row width, row stride). */
AREAL_B200_API int64_t areal_b200_vocab_push_buffer_bytes(int64_t max_tokens, int world_size);
AREAL_B200_API int areal_b200_logits_partial_fwd_push(const void* logits, int dtype, int64_t T, int64_t V,
                                                      int64_t row_stride, const int64_t* labels,
                                                      int64_t vocab_start, double temperature,
                                                      void* const* peer_bases, int rank, int world_size,
                                                      int64_t max_tokens, int64_t epoch, int64_t rows_target,
                                                      int64_t timeout_cycles, float* logp, float* entropy,
                                                      float* row_max, float* ln_sumexp, float* raw_min,
                                                      int* status, void* stream);

/* Token log-prob + CLAMPED softmax entropy of the AEnt recipe, and their gradient.  Replaces
 * `gather_logprobs_clamped_entropy` / `clamped_softmax_entropy` (recipe/AEnt/functional.py:15-80): the
 * k = int(V * entropy_clamp) smallest logits of every row are left out of the entropy's softmax.  The
 * reference finds them with a top-k on a CPU copy of the logits (:18-27).
 *   logits [T,V] bf16 (dtype 0) or fp16 (dtype 1), rows `row_stride` elements apart; labels i64[T];
 *   k_removed in [0, V).  Outputs logp f32[T] (full softmax, as gather_logprobs), entropy f32[T]
 *   (clamped), saved f32[T,8] (row max, ln S, ln S_kept, H, threshold key, tie weight, row min, 0) for
 *   the backward and the vocab_min / vocab_max statistics.
 *   Logits equal to the threshold value are interchangeable: each carries the weight
 *   (n_tie - removed_tie) / n_tie in the sums and in the gradient (the reference's top-k removes some
 *   `removed_tie` of them, which ones is unspecified).
 *   backward: dlogits (same dtype / shape) = d(sum g_logp*logp + g_entropy*entropy)/dlogits;
 *   g_logp / g_entropy f32[T], either may be NULL. */
AREAL_B200_API int areal_b200_clamped_entropy_fwd(const void* logits, int dtype, int64_t T, int64_t V,
                                                  int64_t row_stride, const int64_t* labels,
                                                  double temperature, int64_t k_removed, float* logp,
                                                  float* entropy, float* saved, void* stream);
AREAL_B200_API int areal_b200_clamped_entropy_bwd(const void* logits, int dtype, int64_t T, int64_t V,
                                                  int64_t row_stride, const int64_t* labels,
                                                  double temperature, const float* saved,
                                                  const float* g_logp, const float* g_entropy,
                                                  void* dlogits, int64_t d_row_stride, void* stream);

/* LM head + log-prob / entropy WITHOUT the [T,V] logits tensor, forward only: a tcgen05 GEMM (TMA tensor-map
 * loads, accumulators in tensor memory, cta_group::2 pairs) whose epilogue is the online softmax.  Replaces
 * `outputs.logits` followed by `gather_logprobs(logits, labels, temperature)` in the actor's compute_logp
 * hook (areal/engine/ppo/actor.py:52-70; the engine side areal/engine/fsdp_engine.py:764-777) and
 * `gather_logprobs_entropy` on the same logits (areal/utils/functional.py:84-127).
 *   hidden bf16[T,H] (last hidden state), weight bf16[V,H] (lm_head.weight), unit stride in H, row strides
 *   in ELEMENTS (multiples of 8), 16-byte aligned bases; labels i64[T].
 *   logits = bf16(hidden @ weight^T) as F.linear returns them; z = float(logits)/temperature.
 *   Outputs f32[T] with the meaning of areal_b200_logits_fwd: logp, row_max, ln_sumexp required; entropy and
 *   raw_min nullable.  workspace: areal_b200_lm_head_workspace_bytes(T, V) bytes, need not be zeroed. */
AREAL_B200_API int64_t areal_b200_lm_head_workspace_bytes(int64_t T, int64_t V);
AREAL_B200_API int areal_b200_lm_head_logprobs(const void* hidden, const void* weight, const int64_t* labels,
                                               int64_t T, int64_t H, int64_t V, int64_t hidden_row_stride,
                                               int64_t weight_row_stride, double temperature, float* logp,
                                               float* entropy, float* row_max, float* ln_sumexp, float* raw_min,
                                               void* workspace, int64_t workspace_bytes, void* stream);

/* Legacy running value statistics.  Replaces the arithmetic of `ExponentialRunningMeanStd` /
 * `MovingAverageRunningMeanStd` (realhf/impl/model/modules/rms.py:19-183): update = three reductions
 * + ~10 scalar kernels there; here areal_b200_masked_moments (one group) followed by
 * areal_b200_rms_update, and one elementwise launch for normalize / denormalize.
 *   state: 3 values of the module dtype (state_is_f64 = 1: double, 0: float): mean, mean of squares,
 *   and the debiasing term (mode 0, exponential) or the accumulated denominator (mode 1, moving
 *   average).  partials f64[n_partials,3] as written by areal_b200_masked_moments (or one already
 *   all-reduced row with n_partials = 1).
 *   rms_apply: out = clip((x - mean) / std, -5, 5) or x * std + mean (denormalize = 1), computed in
 *   the module dtype and rounded to f32; std = sqrt(max(mean_sq - mean^2, 1e-2)). */
AREAL_B200_API int areal_b200_rms_update(const double* partials, int64_t n_partials, void* state,
                                         int state_is_f64, int mode, double beta, void* stream);
AREAL_B200_API int areal_b200_rms_apply(const float* x, int64_t N, const void* state, int state_is_f64,
                                        int mode, double eps, int denormalize, float* out, void* stream);

/* KL-regularised token rewards of the legacy packed path in ONE launch.  Replaces
 * `get_packed_rewards` (mode 0) and `get_packed_reward_dense` (mode 1; mode 2 = reward_delta)
 * (realhf/impl/model/utils/ppo_functional.py:228-289; 6 torch kernels, or a Python loop over the
 * sequences with 2 kernels each for the dense variant).
 *   log_probs, ref_log_probs f32[Ts] (one step shorter per sequence than the packed ids);
 *   short1cu_seqlens i32[B+1]; score f32[B] (mode 0) or f32[Ts+B] dense (modes 1, 2);
 *   seq_no_eos u8/bool[B] (mode 0 with mask_no_eos_with_zero; may be NULL otherwise).
 *   Outputs kl_rewards, tot_rewards f32[Ts]. */
AREAL_B200_API int areal_b200_packed_rewards(const float* log_probs, const float* ref_log_probs,
                                             const float* score, const int32_t* short1cu_seqlens,
                                             const uint8_t* seq_no_eos, int64_t total_short_tokens,
                                             int64_t batch_size, double kl_ctl, double clip_reward_value,
                                             int mask_no_eos_with_zero, int mode, float* kl_rewards_out,
                                             float* tot_rewards_out, void* stream);

/* Masked sum / count / min / max of n_keys (<= 32) float arrays of n elements in ONE launch: the
 * reductions `DistributedStatsTracker` does per key at export time
 * (areal/utils/stats_tracker.py:210-289: torch.where(d, v, fill) + sum/min/max per key and per
 * recorded micro-batch) for the keys `grpo_loss_fn` records (areal/engine/ppo/actor.py:352-391).
 *   values[k] f32[n] device pointer or NULL (every element is 1: a denominator's own count); bit k
 *   of value_is_u8_bits says values[k] points at u8/bool elements instead (`mask.float()`, :367-368);
 *   masks[k] u8/bool[n] device pointer or NULL (every element counts).  `values` / `masks` are HOST
 *   arrays of n_keys device pointers, and so is `outs`: outs[k] -> f64[4] = (sum, count, min, max) of
 *   key k (distinct per key); min/max are +inf/-inf when nothing is selected.  accumulate=1 folds the
 *   result into what outs[k] already holds (running totals over micro-batches; initialise a slot to
 *   0, 0, +inf, -inf).  Sums are fp64, in a fixed order.
 *   workspace: areal_b200_masked_stats_workspace_bytes(n_keys, n) bytes, zeroed before first use
 *   (the kernel leaves it zeroed). */
AREAL_B200_API int64_t areal_b200_masked_stats_workspace_bytes(int64_t n_keys, int64_t n);
AREAL_B200_API int areal_b200_masked_stats(const float* const* values, const uint8_t* const* masks,
                                           double* const* outs, uint32_t value_is_u8_bits,
                                           int64_t n_keys, int64_t n, int accumulate, void* workspace,
                                           int64_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AREAL_B200_H_ */
