"""Host-side mirror of `PPOActor.compute_advantages` (areal/engine/ppo/actor.py:72-164).

The reference builds the KL-regularised token rewards with ~25 small torch kernels on `[B,S]`
tensors and then runs the GAE recursion as a Python loop of S steps (5 kernels per step).  Here the
whole block between "Reward Scaling" (:96) and `data["returns"]` (:152) is ONE launch of
`areal_b200_advantages_padded`; the optional reward / advantage `Normalization` stay the three-launch
CUDA path of `areal_b200.data`.  Same keys written into `data`, same in-place side effects
(`data["ref_logp"]` is masked in place, `data["rewards"]` takes the overlong penalty in place).
CUDA tensors only, no fallback.
"""
from __future__ import annotations

from typing import Any, Dict

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr

_MASK_CODE = {torch.float32: 0, torch.uint8: 1, torch.bool: 1, torch.int32: 2, torch.int64: 3}


def _mask_arg(t: torch.Tensor):
    """(tensor kept alive, element code) for a [B,S] mask of any dtype."""
    if t.dtype not in _MASK_CODE:
        t = t.float()
    t = t.contiguous()
    return t, _MASK_CODE[t.dtype]


def _kl_kind(est) -> tuple:
    name = getattr(est, "kl_estimator", est)
    if name not in ("k1", "k2", "k3"):
        raise ValueError(f"Invalid KL estimator: {name}. Valid choices: k1, k2, k3")
    return int(name[1]), bool(getattr(est, "apply_clamp", True))


@torch.no_grad()
def compute_advantages(self, data: Dict[str, Any]) -> None:
    """Drop-in for `PPOActor.compute_advantages(self, data)`; `self` is the reference's PPOActor (or
    anything with the same attributes: config, reward_bias/scaling/clip, reward_norm, kl_ctl,
    kl_estimator, adv_norm, discount, gae_lambda, mask_no_eos_with_zero)."""
    cfg = self.config
    input_ids = data["input_ids"]
    dev = _lib.require_cuda(input_ids, data["rewards"], data["loss_mask"], data["logprobs"], data["attention_mask"])
    B, S = input_ids.shape

    score = data["rewards"]
    overlong = bool(getattr(cfg, "overlong_reward_penalty", False))
    in_kernel_scaling = self.reward_norm is None
    if overlong and not (score.dtype == torch.float32 and score.is_contiguous() and in_kernel_scaling):
        # rare combinations (a reward_norm between scaling and GAE, or non-f32 scores): the [B]-sized
        # prologue runs as the reference's own expressions, the [B,S] work stays in the kernel
        resp = data["loss_mask"].sum(dim=-1).long()
        pen = -(resp - (cfg.max_new_tokens - cfg.overlong_tokens)) / cfg.overlong_tokens * cfg.overlong_penalty_factor
        score += torch.clamp(pen, max=0).to(score.dtype)
        overlong = False
    if in_kernel_scaling:
        score_arg = score if score.dtype == torch.float32 and score.is_contiguous() else score.float().contiguous()
    else:
        rs = (score + self.reward_bias) * self.reward_scaling
        rs = torch.clip(rs, max=self.reward_clip, min=-self.reward_clip)
        score_arg = self.reward_norm(rs).float().contiguous()

    lm, lm_code = _mask_arg(data["loss_mask"])
    am, am_code = _mask_arg(data["attention_mask"])
    decoupled = bool(cfg.use_decoupled_loss)
    if not decoupled and cfg.recompute_logprob:
        logp, roll = data["prox_logp"], 0          # actor.py:112-114: overwritten, not rolled
    else:
        logp, roll = data["logprobs"], 1           # :116
    logp_c = logp.float().contiguous()
    ref = data.get("ref_logp", None)
    ref_c = None
    if ref is not None:
        ref_c = ref if (ref.dtype == torch.float32 and ref.is_contiguous()) else ref.float().contiguous()
    values = data.get("values", None)
    val_c = None if values is None else values.float().contiguous()
    kind, clamp = _kl_kind(self.kl_estimator)

    f32 = dict(dtype=torch.float32, device=dev)
    adv, ret, klr, tot, lm_out, old_out = (torch.empty(B, S, **f32) for _ in range(6))
    with torch.cuda.device(dev):
        call("areal_b200_advantages_padded", ptr(score_arg), int(overlong),
             int(getattr(cfg, "overlong_tokens", 1) or 1) if overlong else 1,
             int(getattr(cfg, "max_new_tokens", 0) or 0) if overlong else 0,
             float(getattr(cfg, "overlong_penalty_factor", 0.0) or 0.0) if overlong else 0.0,
             ptr(lm), lm_code, ptr(logp_c), roll, ptr(ref_c) if ref_c is not None else None, ptr(am), am_code,
             ptr(val_c) if val_c is not None else None, B, S, int(in_kernel_scaling), float(self.reward_bias),
             float(self.reward_scaling), float(self.reward_clip), float(self.kl_ctl), kind, int(clamp),
             int(bool(self.mask_no_eos_with_zero)), float(self.discount), float(self.gae_lambda), ptr(adv), ptr(ret),
             ptr(klr), ptr(tot), ptr(lm_out), ptr(old_out), stream_ptr(dev))
    if ref is not None and ref_c is not ref:
        ref.copy_(ref_c)                            # keep the in-place semantics of :121

    if not decoupled:
        data["prox_logp"] = old_out                 # :114 / :118-119
    data["returns"] = ret
    if self.adv_norm is not None:
        adv = self.adv_norm(adv, lm_out)
    data["advantages"] = adv
    data["kl_rewards"] = klr
    data["tot_rewards"] = tot
    data["loss_mask"] = lm_out
    data["logprobs"] = old_out


_ONES = {}


def _all_tokens(n: int, device) -> torch.Tensor:
    """The `n_tokens` denominator (actor.py:353): built once per (n, device), never written."""
    key = (int(n), str(device))
    t = _ONES.get(key)
    if t is None:
        t = _ONES[key] = torch.ones(int(n), dtype=torch.bool, device=device)
    return t


def grpo_loss_fn(logits: torch.Tensor, input_data: Dict, temperature: float, eps_clip: float,
                 eps_clip_higher: float | None, c_clip: float | None, behav_imp_weight_cap: float | None,
                 tracker=None, loss_scale=None):
    """Drop-in for `grpo_loss_fn` (areal/engine/ppo/actor.py:313-391): same arguments, same keys
    recorded under the same denominators, same returned loss.  The reference reads `[T,V]` four times
    (log-softmax, entropy, min, max), builds ~15 `[T]` temporaries and clones each of them into the
    tracker; here ONE pass over the logits gives every per-token quantity
    (`areal_b200.functional.fused_actor_loss`) and ONE more launch reduces all ~20 keys into the
    tracker's running totals (`stats_tracker.record`).

    `tracker`: an `areal_b200.stats_tracker.DistributedStatsTracker` (default: its DEFAULT_TRACKER).
    `loss_scale`: see fused_actor_loss (single-pass forward+backward when the caller's scale is known).
    When it is None, the scale announced by `areal_b200.engine_hooks.install_loss_scale_hook` for the
    `train_batch` call in progress is used, if any -- that is how the unmodified engine reaches the
    single pass."""
    from . import functional as F
    from . import stats_tracker as ST
    from .engine_hooks import announced_loss_scale
    tracker = ST.DEFAULT_TRACKER if tracker is None else tracker
    if loss_scale is None:
        loss_scale = announced_loss_scale()
    labels = input_data.get("rolled_input_ids", None)
    if labels is None:
        labels = torch.roll(input_data["input_ids"], shifts=-1, dims=-1)
    old_logp = input_data["logprobs"]
    advantages = input_data["advantages"]
    loss_mask = input_data.get("full_loss_mask", input_data["loss_mask"]).bool()
    prox_logp = input_data["prox_logp"]

    if F._tp_group() is not None or F._sp_group() is not None:
        # vocabulary shards (Megatron TP) or token slices (Ulysses): the log-probs cross ranks inside
        # gather_logprobs_entropy, so the loss runs on its result as upstream (:329-347); min / max are of
        # the local logits, as upstream (:375-376), and come out of the same pass over them
        logprobs, entropy, vmin, vmax = F.gather_logprobs_entropy_minmax(logits, labels, temperature)
        loss, stat = F.ppo_actor_loss_fn(logprobs, prox_logp, old_logp, advantages, eps_clip, loss_mask,
                                         eps_clip_higher=eps_clip_higher, c_clip=c_clip,
                                         behav_imp_weight_cap=behav_imp_weight_cap)
        stat = dict(stat, logprobs=logprobs.detach(), entropy=entropy.detach().float(),
                    vocab_min_logits=vmin, vocab_max_logits=vmax)
    else:
        loss, stat = F.fused_actor_loss(logits, labels, temperature, prox_logp, old_logp, advantages, loss_mask,
                                        eps_clip, eps_clip_higher=eps_clip_higher, c_clip=c_clip,
                                        behav_imp_weight_cap=behav_imp_weight_cap, with_logits_minmax=True,
                                        loss_scale=loss_scale)
    _record_actor_stats(tracker, stat, old_logp, loss_mask, logits.shape[0], logits.device)
    return loss


def _record_actor_stats(tracker, stat, old_logp, loss_mask, n_rows, device) -> None:
    """The tracker calls of grpo_loss_fn (actor.py:349-390) as ONE record(): same keys, same denominators."""
    old_f = old_logp if old_logp.dtype == torch.float32 else old_logp.float()
    denominators = dict(n_tokens=_all_tokens(n_rows, device), n_valid_tokens=loss_mask,
                        clipped_tokens=stat["clip_mask"], dual_clipped_tokens=stat["dual_clip_mask"])
    stats = [("n_valid_tokens", dict(importance_weight=stat["importance_weight"], approx_kl=stat["approx_kl"],
                                     new_logp=stat["logprobs"], old_logp=old_f, entropy=stat["entropy"],
                                     actor_loss=stat["loss"], clip_ratio=stat["clip_mask"],
                                     dual_clip_ratio=stat["dual_clip_mask"]))]
    if "behave_imp_weight" in stat:
        denominators["unclipped_behave_tokens"] = stat["behave_mask"]
        stats.append(("unclipped_behave_tokens", dict(behave_imp_weight=stat["behave_imp_weight"],
                                                      behave_approx_kl=stat["behave_approx_kl"])))
    stats.append(("n_tokens", dict(vocab_min_logits=stat["vocab_min_logits"], vocab_max_logits=stat["vocab_max_logits"])))
    # torch.where(clip_mask, x, 0.0) under the `clipped_tokens` denominator (:383-390) selects exactly
    # the clipped entries of x: no temporaries needed
    stats.append(("clipped_tokens", dict(clipped_new_logp=stat["logprobs"], clipped_old_logp=old_f)))
    tracker.record(denominators, stats)


def grpo_loss_fn_from_hidden(hidden: torch.Tensor, lm_head_weight: torch.Tensor, input_data: Dict, temperature: float,
                             eps_clip: float, eps_clip_higher: float | None, c_clip: float | None,
                             behav_imp_weight_cap: float | None, tracker=None, loss_scale=1.0,
                             chunk_rows: int = 4096):
    """`grpo_loss_fn` for an engine that hands over the last hidden state `[T, H]` and the LM-head weight
    `[V, H]` instead of `[T, V]` logits (areal_b200/lm_head.py): same micro-batch keys read, same tracker
    keys recorded, same returned loss; `(loss * loss_scale).backward()` delivers the gradients of the hidden
    state and of the weight, which were computed block by block in the forward."""
    from . import stats_tracker as ST
    from .lm_head import fused_lm_head_actor_loss
    tracker = ST.DEFAULT_TRACKER if tracker is None else tracker
    labels = input_data.get("rolled_input_ids", None)
    if labels is None:
        labels = torch.roll(input_data["input_ids"], shifts=-1, dims=-1)
    old_logp = input_data["logprobs"]
    loss_mask = input_data.get("full_loss_mask", input_data["loss_mask"]).bool()
    loss, stat = fused_lm_head_actor_loss(hidden, lm_head_weight, labels, temperature, input_data["prox_logp"],
                                          old_logp, input_data["advantages"], loss_mask, eps_clip,
                                          eps_clip_higher=eps_clip_higher, c_clip=c_clip,
                                          behav_imp_weight_cap=behav_imp_weight_cap, with_logits_minmax=True,
                                          loss_scale=loss_scale, chunk_rows=chunk_rows)
    _record_actor_stats(tracker, stat, old_logp, loss_mask, hidden.shape[0], hidden.device)
    return loss
