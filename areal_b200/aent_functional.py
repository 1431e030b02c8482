"""Host-side mirror of the AEnt recipe's clamped-entropy entry points, `recipe/AEnt/functional.py`
(`gather_logprobs_clamped_entropy` :54-86, `clamped_softmax_entropy` :15-36).

The reference computes the clamped entropy per 1024-row chunk by copying the chunk's logits to the CPU,
running a top-k there to find the `int(V * entropy_clamp)` smallest logits of each row, and scattering
a `[chunk, V]` boolean mask back to the GPU.  Here one kernel reads each row three times (the second
and third time out of L2), finds the threshold with a 16-bit radix select in shared memory and returns
the full-softmax log-prob of the label and the clamped entropy; the backward is one streaming pass.
CUDA tensors only, bf16 / fp16 logits, no fallback.

Ties: logits equal to the threshold value are interchangeable for the entropy; where the reference's
top-k removes an unspecified subset of them, every one of them here carries the same fractional weight
(kept count / tie count) in the entropy and in its gradient -- the value is identical, the gradient is
the average over the possible tie-breaks.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr
from .functional import _f32, _rows


class _ClampedEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, entropy_clamp, temperature):
        if logits.dtype not in (torch.bfloat16, torch.float16):
            raise TypeError(f"clamped entropy needs bf16 or fp16 logits, got {logits.dtype}")
        dev, x, lab, lead, V = _rows(logits, labels)
        T = x.shape[0]
        k = int(V * float(entropy_clamp))                 # recipe/AEnt/functional.py:21
        if not 0 <= k < V:
            raise ValueError(f"entropy_clamp={entropy_clamp} removes {k} of {V} logits; need 0 <= k < V")
        f32 = dict(dtype=torch.float32, device=dev)
        logp, ent = torch.empty(T, **f32), torch.empty(T, **f32)
        saved = torch.empty(T, 8, **f32)
        with torch.cuda.device(dev):
            call("areal_b200_clamped_entropy_fwd", ptr(x), _lib.DTYPE_CODE[x.dtype], T, V, x.stride(0) if T > 1 else V,
                 ptr(lab), float(temperature), k, ptr(logp), ptr(ent), ptr(saved), stream_ptr(dev))
        ctx.save_for_backward(x, lab, saved)
        ctx.temperature = float(temperature)
        ctx.in_shape = logits.shape
        ctx.set_materialize_grads(False)
        row_max, row_min = saved[:, 0].view(lead), saved[:, 6].view(lead)
        ctx.mark_non_differentiable(row_max, row_min)
        return logp.view(lead), ent.view(lead), row_max, row_min

    @staticmethod
    def backward(ctx, g_logp, g_ent, *unused):
        x, lab, saved = ctx.saved_tensors
        if g_logp is None and g_ent is None:
            return None, None, None, None
        T, V = x.shape
        dev = x.device
        gl = None if g_logp is None else _f32(g_logp.reshape(-1))
        gh = None if g_ent is None else _f32(g_ent.reshape(-1))
        d = torch.empty((T, V), dtype=x.dtype, device=dev)
        with torch.cuda.device(dev):
            call("areal_b200_clamped_entropy_bwd", ptr(x), _lib.DTYPE_CODE[x.dtype], T, V, x.stride(0) if T > 1 else V,
                 ptr(lab), ctx.temperature, ptr(saved), ptr(gl), ptr(gh), ptr(d), V, stream_ptr(dev))
        return d.view(ctx.in_shape), None, None, None


def gather_logprobs_clamped_entropy(logits: torch.Tensor, labels: torch.Tensor, entropy_clamp: float,
                                    temperature: float = 1.0, chunk_size: int = 1024):
    """Drop-in for `gather_logprobs_clamped_entropy` (recipe/AEnt/functional.py:54-86): (log-prob of
    `labels` under the full softmax, entropy of the softmax with the `int(V * entropy_clamp)` smallest
    logits of each row masked out), both f32 with the shape of `labels`, differentiable w.r.t. `logits`.
    `chunk_size` is accepted for signature compatibility (rows are streamed; no chunking needed)."""
    return _ClampedEntropy.apply(logits, labels, entropy_clamp, temperature)[:2]


def clamped_softmax_entropy(logits: torch.Tensor, entropy_clamp: float) -> torch.Tensor:
    """Drop-in for `clamped_softmax_entropy` (recipe/AEnt/functional.py:15-36); logits already divided
    by the temperature, as the reference calls it (:48-50)."""
    labels = torch.zeros(logits.shape[:-1], dtype=torch.long, device=logits.device)
    return _ClampedEntropy.apply(logits, labels, entropy_clamp, 1.0)[1]


def clamped_entropy_loss_fn(clamped_entropy: torch.Tensor, loss_mask: torch.Tensor) -> torch.Tensor:
    """Drop-in for `clamped_entropy_loss_fn` (recipe/AEnt/actor.py:273-278): masked mean of the clamped
    entropy; `loss_mask.count_nonzero() or 1` without the host sync."""
    m = loss_mask.bool()
    count = torch.clamp(m.sum(), min=1)
    return torch.where(m, clamped_entropy, 0.0).sum() / count


def aent_grpo_loss_fn(logits: torch.Tensor, input_data: dict, temperature: float, eps_clip: float,
                      entropy_coeff: float, entropy_clamp: float, eps_clip_higher: float | None,
                      c_clip: float | None, behav_imp_weight_cap: float | None, tracker=None):
    """Drop-in for `aent_grpo_loss_fn` (recipe/AEnt/actor.py:185-270): PPO actor loss minus
    `entropy_coeff` times the masked mean clamped entropy, the same keys recorded in the stats tracker.
    The logits are read by the clamped-entropy kernel only (its row max / min feed the vocab statistics);
    with `entropy_clamp == 0` the fused forward of `areal_b200.functional` is used, as upstream does (:207-212)."""
    from . import functional as F
    from . import stats_tracker as ST
    from .actor import _all_tokens
    tracker = ST.DEFAULT_TRACKER if tracker is None else tracker
    labels = input_data.get("rolled_input_ids", None)
    if labels is None:
        labels = torch.roll(input_data["input_ids"], shifts=-1, dims=-1)
    old_logp = input_data["logprobs"]
    advantages = input_data["advantages"]
    loss_mask = input_data.get("full_loss_mask", input_data["loss_mask"]).bool()
    prox_logp = input_data["prox_logp"]
    if entropy_clamp > 0:
        logprobs, ent, vmax, vmin = _ClampedEntropy.apply(logits, labels, entropy_clamp, temperature)
    else:
        st = F.logits_stats(logits.detach(), labels, temperature)
        logprobs, ent = F.gather_logprobs_entropy(logits, labels, temperature)
        vmax, vmin = st["vocab_max_logits"], st["vocab_min_logits"]
    ppo_loss, stat = F.ppo_actor_loss_fn(logprobs=logprobs, old_logprobs=old_logp, advantages=advantages,
                                         eps_clip=eps_clip, eps_clip_higher=eps_clip_higher, loss_mask=loss_mask,
                                         c_clip=c_clip, proximal_logprobs=prox_logp,
                                         behav_imp_weight_cap=behav_imp_weight_cap)
    loss = ppo_loss - entropy_coeff * clamped_entropy_loss_fn(ent, loss_mask)
    old_f = old_logp if old_logp.dtype == torch.float32 else old_logp.float()
    lp_d = logprobs.detach()
    denominators = dict(n_tokens=_all_tokens(logits.shape[0], logits.device), n_valid_tokens=loss_mask,
                        clipped_tokens=stat["clip_mask"], dual_clipped_tokens=stat["dual_clip_mask"])
    stats = [("n_valid_tokens", dict(importance_weight=stat["importance_weight"], approx_kl=stat["approx_kl"],
                                     new_logp=lp_d, old_logp=old_f, entropy=ent.detach().float(),
                                     actor_loss=stat["loss"], clip_ratio=stat["clip_mask"],
                                     dual_clip_ratio=stat["dual_clip_mask"]))]
    if "behave_imp_weight" in stat:
        denominators["unclipped_behave_tokens"] = stat["behave_mask"]
        stats.append(("unclipped_behave_tokens", dict(behave_imp_weight=stat["behave_imp_weight"],
                                                      behave_approx_kl=stat["behave_approx_kl"])))
    stats.append(("n_tokens", dict(vocab_min_logits=vmin.float(), vocab_max_logits=vmax.float())))
    stats.append(("clipped_tokens", dict(clipped_new_logp=lp_d, clipped_old_logp=old_f)))
    tracker.record(denominators, stats)
    return loss
