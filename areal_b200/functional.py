"""Host-side mirror of `areal/utils/functional.py` for the RL post-processing hot path.

Same names, argument meaning, return types and error behaviour as the reference functions, but every
op is one launch of a hand-written sm_100a kernel behind the C ABI (include/areal_b200.h):

  gather_logprobs / gather_logprobs_entropy   reference areal/utils/functional.py:43-127
  ppo_actor_loss_fn                           reference areal/utils/functional.py:171-235
  ppo_critic_loss_fn                          reference areal/utils/functional.py:247-311
  fused_actor_loss                            everything `grpo_loss_fn` derives from the logits
                                              (areal/engine/ppo/actor.py:313-391) in ONE pass

CUDA tensors only; there is no CPU / eager fallback.
"""
from __future__ import annotations

import math
import sys
from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr

_NAN = float("nan")


def _opt(x: Optional[float]) -> float:
    return _NAN if x is None else float(x)


def _rows(logits: torch.Tensor, labels: torch.Tensor):
    if logits.dtype not in _lib.DTYPE_CODE:
        raise TypeError(f"logits dtype {logits.dtype} not supported (bf16, fp16, fp32)")
    if logits.shape[:-1] != labels.shape:
        raise RuntimeError(f"labels shape {tuple(labels.shape)} does not match logits {tuple(logits.shape)}")
    dev = _lib.require_cuda(logits, labels)
    lead = logits.shape[:-1]
    V = logits.shape[-1]
    x = logits if logits.dim() == 2 else logits.reshape(-1, V)
    if x.stride(-1) != 1 or (x.shape[0] > 1 and x.stride(0) < V):
        x = x.contiguous()
    lab = labels.reshape(-1)
    if lab.dtype != torch.int64:
        lab = lab.long()
    lab = lab.contiguous()
    return dev, x, lab, lead, V


def _f32(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _mask_u8(m: torch.Tensor) -> torch.Tensor:
    if m.dtype != torch.bool:
        m = m.bool()
    return m.contiguous().view(torch.uint8)


# --------------------------------------------------------------------------------------------------
# log-prob / entropy
# --------------------------------------------------------------------------------------------------
def _logits_forward(x, lab, temperature, want_entropy, want_min=False):
    T, V = x.shape
    dev = x.device
    f32 = dict(dtype=torch.float32, device=dev)
    logp = torch.empty(T, **f32)
    row_max = torch.empty(T, **f32)
    ln_s = torch.empty(T, **f32)
    ent = torch.empty(T, **f32) if want_entropy else None
    mn = torch.empty(T, **f32) if want_min else None
    with torch.cuda.device(dev):
        call("areal_b200_logits_fwd", ptr(x), _lib.DTYPE_CODE[x.dtype], T, V, x.stride(0) if T > 1 else V,
             ptr(lab), float(temperature), ptr(logp), ptr(ent), ptr(row_max), ptr(ln_s), ptr(mn),
             stream_ptr(dev))
    return logp, ent, row_max, ln_s, mn


def _logits_backward(x, lab, temperature, row_max, ln_s, ent, g_logp, g_ent, scale0=None, scale1=None,
                     vocab_start=0):
    T, V = x.shape
    dev = x.device
    dlogits = torch.empty((T, V), dtype=x.dtype, device=dev)
    with torch.cuda.device(dev):
        call("areal_b200_logits_bwd", ptr(x), _lib.DTYPE_CODE[x.dtype], T, V, x.stride(0) if T > 1 else V,
             ptr(lab), int(vocab_start), float(temperature), ptr(row_max), ptr(ln_s), ptr(ent), ptr(g_logp),
             ptr(g_ent), ptr(scale0), ptr(scale1), ptr(dlogits), V, stream_ptr(dev))
    return dlogits


class _GatherLogprobsEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, temperature, want_entropy, want_minmax=False):
        dev, x, lab, lead, V = _rows(logits, labels)
        logp, ent, row_max, ln_s, mn = _logits_forward(x, lab, temperature, want_entropy, want_minmax)
        ctx.save_for_backward(x, lab, row_max, ln_s, ent if want_entropy else None)
        ctx.temperature = float(temperature)
        ctx.in_shape = logits.shape
        ctx.set_materialize_grads(False)
        logp = logp.view(lead)
        if want_minmax:                       # row min / max of the raw logits out of the same pass
            mn, mx = mn.view(lead), row_max.clone().view(lead)
            ctx.mark_non_differentiable(mn, mx)
            return logp, ent.view(lead), mn, mx
        if want_entropy:
            return logp, ent.view(lead)
        return logp

    @staticmethod
    def backward(ctx, g_logp, g_ent=None, *unused):
        x, lab, row_max, ln_s, ent = ctx.saved_tensors
        gl = None if g_logp is None else _f32(g_logp.reshape(-1))
        gh = None if (g_ent is None or ent is None) else _f32(g_ent.reshape(-1))
        if gl is None and gh is None:
            return None, None, None, None, None
        d = _logits_backward(x, lab, ctx.temperature, row_max, ln_s, ent, gl, gh)
        return d.view(ctx.in_shape), None, None, None, None


# The two distributed branches INSIDE the reference functions (areal/utils/functional.py:50-58 and 91-98:
# Megatron tensor parallelism, logits sharded over the vocabulary; :76-79 and :120-125: Ulysses sequence
# parallelism, every rank holds a slice of the tokens and the results are all-gathered).
_GROUPS = {"explicit": False, "tp": None, "sp": None}


def set_parallel_groups(tp_group=None, sp_group=None) -> None:
    """Tell gather_logprobs / gather_logprobs_entropy which process groups shard the vocabulary (tp) and
    the sequence (sp).  Without this call the groups are looked up where the reference looks them up --
    `megatron.core.parallel_state` and `areal.utils.ulysses` -- if the host program has imported them.
    `set_parallel_groups()` with no arguments pins both branches off."""
    _GROUPS.update(explicit=True, tp=tp_group, sp=sp_group)


def reset_parallel_groups() -> None:
    _GROUPS.update(explicit=False, tp=None, sp=None)


def _world(group) -> int:
    import torch.distributed as dist
    return dist.get_world_size(group) if (group is not None and dist.is_initialized()) else 1


def _tp_group():
    if _GROUPS["explicit"]:
        g = _GROUPS["tp"]
        return g if _world(g) > 1 else None
    mpu = sys.modules.get("megatron.core.parallel_state")
    try:
        if mpu is not None and mpu.is_initialized() and mpu.get_tensor_model_parallel_world_size() > 1:
            return mpu.get_tensor_model_parallel_group()
    except Exception:
        pass
    return None


def _sp_group():
    if _GROUPS["explicit"]:
        g = _GROUPS["sp"]
        return g if _world(g) > 1 else None
    uly = sys.modules.get("areal.utils.ulysses")
    try:
        if uly is not None and uly.get_ulysses_sequence_parallel_world_size() > 1:
            return uly.get_ulysses_sequence_parallel_group()
    except Exception:
        pass
    return None


def _sp_all_gather(x: torch.Tensor, group) -> torch.Tensor:
    """functional.py:76-79: differentiable all-gather over the sequence-parallel group, concatenated on
    the last (token) dimension in rank order."""
    import torch.distributed.nn.functional as dist_F
    return torch.cat(dist_F.all_gather(x, group=group), dim=-1)


def _vocab_parallel(logits, labels, temperature, group):
    from .dist import vocab_parallel_logprobs_entropy
    lp, ent = vocab_parallel_logprobs_entropy(logits, labels, temperature, group=group)
    return lp.view(labels.shape), ent.view(labels.shape)


def gather_logprobs(logits: torch.Tensor, labels: torch.Tensor, temperature: float = 1.0,
                    chunk_size: int = 1024) -> torch.Tensor:
    """log_softmax(logits.float() / temperature)[..., labels]  (reference functional.py:43-81).
    `chunk_size` is accepted for signature compatibility; the kernel streams rows and needs no
    chunking.  Differentiable w.r.t. `logits` (gradient in the logits' dtype).  Under tensor
    parallelism `logits` is this rank's vocabulary shard and `labels` hold global ids (:50-58); under
    Ulysses the per-rank results are all-gathered (:76-79)."""
    tp = _tp_group()
    if tp is not None:
        return _vocab_parallel(logits, labels, temperature, tp)[0]
    logp = _GatherLogprobsEntropy.apply(logits, labels, temperature, False)
    sp = _sp_group()
    if sp is not None:
        logp = _sp_all_gather(logp, sp)
    return logp


def gather_logprobs_entropy(logits: torch.Tensor, labels: torch.Tensor, temperature: float = 1.0,
                            chunk_size: int = 1024) -> Tuple[torch.Tensor, torch.Tensor]:
    """(log-prob of labels, entropy of softmax(logits/temperature))  (reference functional.py:84-127),
    with the same two distributed branches (:91-98, :120-125)."""
    tp = _tp_group()
    if tp is not None:
        return _vocab_parallel(logits, labels, temperature, tp)
    logp, ent = _GatherLogprobsEntropy.apply(logits, labels, temperature, True)
    sp = _sp_group()
    if sp is not None:
        logp, ent = _sp_all_gather(logp, sp), _sp_all_gather(ent, sp)
    return logp, ent


def gather_logprobs_entropy_minmax(logits: torch.Tensor, labels: torch.Tensor, temperature: float = 1.0):
    """`gather_logprobs_entropy` plus the row minimum / maximum of the LOCAL logits -- what `grpo_loss_fn` logs as
    vocab_min_logits / vocab_max_logits (areal/engine/ppo/actor.py:375-376: `logits.detach().min(-1)` / `.max(-1)`, two
    more reads of the rank's logits) -- out of the same pass, on every branch: vocabulary shards (the partial kernel's
    shard maximum and raw minimum), Ulysses token slices and the plain case (the forward kernel's)."""
    tp = _tp_group()
    if tp is not None:
        from .dist import vocab_parallel_logprobs_entropy
        lp, ent, mn, mx = vocab_parallel_logprobs_entropy(logits, labels, temperature, group=tp, with_minmax=True)
        return lp.view(labels.shape), ent.view(labels.shape), mn.view(labels.shape), mx.view(labels.shape)
    logp, ent, mn, mx = _GatherLogprobsEntropy.apply(logits, labels, temperature, True, True)
    sp = _sp_group()
    if sp is not None:
        logp, ent = _sp_all_gather(logp, sp), _sp_all_gather(ent, sp)
    return logp, ent, mn, mx


@torch.no_grad()
def logits_stats(logits: torch.Tensor, labels: torch.Tensor, temperature: float = 1.0):
    """One pass returning dict(logp, entropy, lse, vocab_min_logits, vocab_max_logits): the values
    `grpo_loss_fn` gets from three separate reads of the logits (actor.py:340, 375-376)."""
    dev, x, lab, lead, V = _rows(logits, labels)
    logp, ent, row_max, ln_s, mn = _logits_forward(x, lab, temperature, True, True)
    return dict(logp=logp.view(lead), entropy=ent.view(lead), lse=(row_max / temperature + ln_s).view(lead),
                vocab_min_logits=mn.view(lead), vocab_max_logits=row_max.view(lead))


# --------------------------------------------------------------------------------------------------
# PPO actor loss
# --------------------------------------------------------------------------------------------------
_STAT_F32 = ("loss", "importance_weight", "approx_kl", "behave_imp_weight", "behave_approx_kl")
_STAT_BOOL = ("clip_mask", "dual_clip_mask", "behave_mask")


def _alloc_stats(T, dev, decoupled):
    f = {k: torch.empty(T, dtype=torch.float32, device=dev) for k in _STAT_F32[:3]}
    b = {k: torch.empty(T, dtype=torch.bool, device=dev) for k in _STAT_BOOL[:2]}
    if decoupled:
        f.update({k: torch.empty(T, dtype=torch.float32, device=dev) for k in _STAT_F32[3:]})
        b["behave_mask"] = torch.empty(T, dtype=torch.bool, device=dev)
    return f, b


def _u8(t):
    return None if t is None else t.view(torch.uint8)


class _ActorLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logprobs, prox, old, adv, mask, eps_clip, eps_clip_higher, c_clip, cap):
        dev = _lib.require_cuda(logprobs, prox, old, adv, mask)
        shape = logprobs.shape
        lp = _f32(logprobs.reshape(-1))
        T = lp.numel()
        px = None if prox is None else _f32(prox.reshape(-1))
        od = _f32(old.reshape(-1))
        av = _f32(adv.reshape(-1))
        mk = None if mask is None else _mask_u8(mask.reshape(-1))
        loss = torch.empty((), dtype=torch.float32, device=dev)
        inv = torch.empty(1, dtype=torch.float32, device=dev)      # 1 / max(count, 1)
        coef = torch.empty(T, dtype=torch.float32, device=dev)
        f, b = _alloc_stats(T, dev, px is not None)
        with torch.cuda.device(dev):
            nbytes = _lib.lib().areal_b200_loss_workspace_bytes()
            ws = _lib.workspace("loss", dev, nbytes)
            call("areal_b200_ppo_actor_loss", ptr(lp), ptr(px), ptr(od), ptr(av), ptr(mk), T,
                 float(eps_clip), _opt(eps_clip_higher), _opt(c_clip), _opt(cap),
                 ptr(loss), ptr(inv), ptr(coef), ptr(f["loss"]), ptr(f["importance_weight"]),
                 ptr(f["approx_kl"]), ptr(_u8(b["clip_mask"])), ptr(_u8(b["dual_clip_mask"])),
                 ptr(f.get("behave_imp_weight")), ptr(f.get("behave_approx_kl")),
                 ptr(_u8(b.get("behave_mask"))), ptr(ws), ws.numel(), stream_ptr(dev))
        ctx.save_for_backward(coef, inv)
        ctx.shape = shape
        ctx.set_materialize_grads(False)
        stats = [f["loss"], f["importance_weight"], f["approx_kl"], b["clip_mask"], b["dual_clip_mask"]]
        if px is not None:
            stats += [f["behave_imp_weight"], f["behave_approx_kl"], b["behave_mask"]]
        stats = [s.view(shape) for s in stats]
        ctx.mark_non_differentiable(*stats)
        return (loss, *stats)

    @staticmethod
    def backward(ctx, g_loss, *unused):
        if g_loss is None:
            return (None,) * 9
        coef, inv = ctx.saved_tensors
        dev = coef.device
        g = _f32(g_loss.reshape(1))
        grad = torch.empty_like(coef)
        with torch.cuda.device(dev):
            call("areal_b200_scale_by_scalars", ptr(coef), ptr(inv), ptr(g), coef.numel(), ptr(grad),
                 stream_ptr(dev))
        return grad.view(ctx.shape), None, None, None, None, None, None, None, None


def ppo_actor_loss_fn(logprobs: torch.Tensor, proximal_logprobs: torch.Tensor, old_logprobs: torch.Tensor,
                      advantages: torch.Tensor, eps_clip: float, loss_mask: torch.Tensor,
                      eps_clip_higher: Optional[float] = None, c_clip: Optional[float] = None,
                      behav_imp_weight_cap: Optional[float] = None) -> Tuple[torch.Tensor, Dict]:
    """Decoupled PPO clipped actor loss (reference functional.py:171-235): same signature, same
    (scalar loss with grad, dict of 8 per-token tensors) return; one kernel, no host sync."""
    if c_clip is not None:
        assert c_clip > 1.0, c_clip
    res = _ActorLoss.apply(logprobs, proximal_logprobs, old_logprobs, advantages, loss_mask, eps_clip,
                           eps_clip_higher, c_clip, behav_imp_weight_cap)
    loss = res[0]
    stat = dict(loss=res[1], importance_weight=res[2], approx_kl=res[3], clip_mask=res[4],
                dual_clip_mask=res[5])
    if proximal_logprobs is not None:
        stat["behave_imp_weight"] = res[6]
        stat["behave_approx_kl"] = res[7]
        stat["behave_mask"] = res[8]
    return loss, stat


# --------------------------------------------------------------------------------------------------
# PPO critic loss
# --------------------------------------------------------------------------------------------------
class _CriticLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, value, old_value, target_value, eps, mask, huber):
        dev = _lib.require_cuda(value, old_value, target_value, mask)
        shape = value.shape
        v = value.reshape(-1).contiguous()
        ov = old_value.reshape(-1).contiguous()
        tv = target_value.reshape(-1).contiguous()
        mk = None if mask is None else _mask_u8(mask.reshape(-1))
        T = v.numel()
        loss = torch.empty((), dtype=torch.float32, device=dev)
        inv = torch.empty(1, dtype=torch.float32, device=dev)
        coef = torch.empty(T, dtype=torch.float32, device=dev)
        stat_loss = torch.empty(T, dtype=torch.float32, device=dev)
        clip = torch.empty(T, dtype=torch.bool, device=dev)
        with torch.cuda.device(dev):
            nbytes = _lib.lib().areal_b200_loss_workspace_bytes()
            ws = _lib.workspace("loss", dev, nbytes)
            call("areal_b200_ppo_critic_loss", ptr(v), ptr(ov), ptr(tv), ptr(mk), T, float(eps), int(huber),
                 ptr(loss), ptr(inv), ptr(coef), ptr(stat_loss), ptr(_u8(clip)), ptr(ws),
                 ws.numel(), stream_ptr(dev))
        ctx.save_for_backward(coef, inv)
        ctx.shape = shape
        ctx.set_materialize_grads(False)
        stat_loss = stat_loss.view(shape)
        clip = clip.view(shape)
        ctx.mark_non_differentiable(stat_loss, clip)
        return loss, stat_loss, clip

    @staticmethod
    def backward(ctx, g_loss, *unused):
        if g_loss is None:
            return (None,) * 6
        coef, inv = ctx.saved_tensors
        dev = coef.device
        g = _f32(g_loss.reshape(1))
        grad = torch.empty_like(coef)
        with torch.cuda.device(dev):
            call("areal_b200_scale_by_scalars", ptr(coef), ptr(inv), ptr(g), coef.numel(), ptr(grad),
                 stream_ptr(dev))
        return grad.view(ctx.shape), None, None, None, None, None


def ppo_critic_loss_fn(value: torch.Tensor, old_value: torch.Tensor, target_value: torch.Tensor,
                       value_eps_clip: float, loss_mask: Optional[torch.Tensor] = None,
                       loss_fn_type: str = "mse") -> Tuple[torch.Tensor, Dict]:
    """PPO clipped value loss (reference functional.py:247-311), same asserts and return."""
    assert value.dtype == torch.float32
    assert old_value.dtype == torch.float32
    assert target_value.dtype == torch.float32
    if loss_fn_type not in ("mse", "huber"):
        raise NotImplementedError(f"Unknown loss fn type: {loss_fn_type}")
    loss, stat_loss, clip = _CriticLoss.apply(value, old_value, target_value, value_eps_clip, loss_mask,
                                              loss_fn_type == "huber")
    return loss, dict(clip_mask=clip, loss=stat_loss)


# --------------------------------------------------------------------------------------------------
# fused: logits -> logp, entropy, min/max, actor loss (+ dlogits in backward)
# --------------------------------------------------------------------------------------------------
class _FusedActor(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, temperature, prox, old, adv, mask, eps_clip, eps_clip_higher, c_clip,
                cap, want_minmax):
        dev, x, lab, lead, V = _rows(logits, labels)
        _lib.require_cuda(x, prox, old, adv, mask)
        T = x.shape[0]
        px = None if prox is None else _f32(prox.reshape(-1))
        od = _f32(old.reshape(-1))
        av = _f32(adv.reshape(-1))
        mk = _mask_u8(mask.reshape(-1))
        f32 = dict(dtype=torch.float32, device=dev)
        logp = torch.empty(T, **f32)
        ent = torch.empty(T, **f32)
        row_max = torch.empty(T, **f32)
        ln_s = torch.empty(T, **f32)
        mn = torch.empty(T, **f32) if want_minmax else None
        loss = torch.empty((), **f32)
        inv = torch.empty(1, **f32)
        coef = torch.empty(T, **f32)
        f, b = _alloc_stats(T, dev, px is not None)
        with torch.cuda.device(dev):
            nbytes = _lib.lib().areal_b200_loss_workspace_bytes()
            ws = _lib.workspace("loss", dev, nbytes)
            call("areal_b200_fused_actor_fwd", ptr(x), _lib.DTYPE_CODE[x.dtype], T, V,
                 x.stride(0) if T > 1 else V, ptr(lab), float(temperature), ptr(px), ptr(od), ptr(av), ptr(mk),
                 float(eps_clip), _opt(eps_clip_higher), _opt(c_clip), _opt(cap),
                 ptr(logp), ptr(ent), ptr(row_max), ptr(ln_s), ptr(mn), ptr(loss), ptr(inv), ptr(coef),
                 ptr(f["loss"]), ptr(f["importance_weight"]), ptr(f["approx_kl"]), ptr(_u8(b["clip_mask"])),
                 ptr(_u8(b["dual_clip_mask"])), ptr(f.get("behave_imp_weight")), ptr(f.get("behave_approx_kl")),
                 ptr(_u8(b.get("behave_mask"))), ptr(ws), ws.numel(), stream_ptr(dev))
        ctx.save_for_backward(x, lab, row_max, ln_s, coef, inv)
        ctx.temperature = float(temperature)
        ctx.in_shape = logits.shape
        ctx.set_materialize_grads(False)
        extras = [logp, ent, f["loss"], f["importance_weight"], f["approx_kl"], b["clip_mask"],
                  b["dual_clip_mask"]]
        if px is not None:
            extras += [f["behave_imp_weight"], f["behave_approx_kl"], b["behave_mask"]]
        if want_minmax:
            extras += [mn, row_max.clone()]
        extras = [e.view(lead) for e in extras]
        ctx.mark_non_differentiable(*extras)
        return (loss, *extras)

    @staticmethod
    def backward(ctx, g_loss, *unused):
        if g_loss is None:
            return (None,) * 12
        x, lab, row_max, ln_s, coef, inv = ctx.saved_tensors
        g = _f32(g_loss.reshape(1))
        d = _logits_backward(x, lab, ctx.temperature, row_max, ln_s, None, coef, None, inv, g)
        return (d.view(ctx.in_shape),) + (None,) * 11


class _FusedActorRowRes(torch.autograd.Function):
    """Single-pass variant: forward also writes dlogits (scaled by `loss_scale`); backward hands it out."""

    @staticmethod
    def forward(ctx, logits, labels, temperature, prox, old, adv, mask, eps_clip, eps_clip_higher, c_clip,
                cap, want_minmax, loss_scale, cluster_size, n_buffers):
        dev, x, lab, lead, V = _rows(logits, labels)
        _lib.require_cuda(x, prox, old, adv, mask, loss_scale)
        T = x.shape[0]
        px = None if prox is None else _f32(prox.reshape(-1))
        od = _f32(old.reshape(-1))
        av = _f32(adv.reshape(-1))
        mk = _mask_u8(mask.reshape(-1))
        f32 = dict(dtype=torch.float32, device=dev)
        logp = torch.empty(T, **f32)
        ent = torch.empty(T, **f32)
        row_max = torch.empty(T, **f32)
        ln_s = torch.empty(T, **f32)
        mn = torch.empty(T, **f32) if want_minmax else None
        loss = torch.empty((), **f32)
        inv = torch.empty(1, **f32)
        coef = torch.empty(T, **f32)
        f, b = _alloc_stats(T, dev, px is not None)
        dlogits = torch.empty((T, V), dtype=x.dtype, device=dev)
        scale = _f32(loss_scale.reshape(1))
        with torch.cuda.device(dev):
            nbytes = _lib.lib().areal_b200_loss_workspace_bytes()
            ws = _lib.workspace("loss", dev, nbytes)
            call("areal_b200_fused_actor_fwd_bwd", ptr(x), _lib.DTYPE_CODE[x.dtype], T, V,
                 x.stride(0) if T > 1 else V, ptr(lab), float(temperature), ptr(px), ptr(od), ptr(av), ptr(mk),
                 float(eps_clip), _opt(eps_clip_higher), _opt(c_clip), _opt(cap),
                 ptr(logp), ptr(ent), ptr(row_max), ptr(ln_s), ptr(mn), ptr(loss), ptr(inv), ptr(coef),
                 ptr(f["loss"]), ptr(f["importance_weight"]), ptr(f["approx_kl"]), ptr(_u8(b["clip_mask"])),
                 ptr(_u8(b["dual_clip_mask"])), ptr(f.get("behave_imp_weight")), ptr(f.get("behave_approx_kl")),
                 ptr(_u8(b.get("behave_mask"))), ptr(scale), ptr(dlogits), V, int(cluster_size), int(n_buffers),
                 ptr(ws), ws.numel(), stream_ptr(dev))
        ctx.save_for_backward(dlogits, scale)
        ctx.in_shape = logits.shape
        ctx.set_materialize_grads(False)
        extras = [logp, ent, f["loss"], f["importance_weight"], f["approx_kl"], b["clip_mask"],
                  b["dual_clip_mask"]]
        if px is not None:
            extras += [f["behave_imp_weight"], f["behave_approx_kl"], b["behave_mask"]]
        if want_minmax:
            extras += [mn, row_max]
        extras = [e.view(lead) for e in extras]
        ctx.mark_non_differentiable(*extras)
        return (loss, *extras)

    @staticmethod
    def backward(ctx, g_loss, *unused):
        if g_loss is None:
            return (None,) * 15
        if getattr(ctx, "handed_out", False):
            # the stored gradient now belongs to the caller (autograd may have taken it as .grad and
            # accumulated into it); a second backward through this node cannot be served from it
            raise RuntimeError("fused_actor_loss(loss_scale=...) keeps ONE gradient buffer and has already handed it "
                               "out; call it with loss_scale=None to back-propagate more than once (retain_graph)")
        ctx.handed_out = True
        dlogits, scale = ctx.saved_tensors
        g = _f32(g_loss.reshape(1))
        T, V = dlogits.shape
        with torch.cuda.device(dlogits.device):
            # no-op when the upstream scalar equals the loss_scale given to the forward (the contract)
            call("areal_b200_rescale_grad_if_needed", ptr(dlogits), T, V, V, ptr(g), ptr(scale),
                 stream_ptr(dlogits.device))
        return (dlogits.view(ctx.in_shape),) + (None,) * 14


def _single_pass_eligible(logits: torch.Tensor) -> bool:
    """What areal_b200_fused_actor_fwd_bwd accepts: bf16 rows, unit stride in V, V % 8 == 0, rows 16-byte
    aligned.  Anything else (fp16 / fp32 logits, odd row strides) takes the two-pass entries."""
    if logits.dtype != torch.bfloat16 or logits.dim() < 1:
        return False
    V = logits.shape[-1]
    if V % 8 != 0 or logits.stride(-1) != 1 or logits.data_ptr() % 16 != 0:
        return False
    if logits.dim() == 2 and logits.shape[0] > 1 and logits.stride(0) % 8 != 0:
        return False
    return True


def fused_actor_loss(logits: torch.Tensor, labels: torch.Tensor, temperature: float,
                     proximal_logprobs: Optional[torch.Tensor], old_logprobs: torch.Tensor,
                     advantages: torch.Tensor, loss_mask: torch.Tensor, eps_clip: float,
                     eps_clip_higher: Optional[float] = None, c_clip: Optional[float] = None,
                     behav_imp_weight_cap: Optional[float] = None,
                     with_logits_minmax: bool = True, loss_scale=None,
                     cluster_size: int = 0, n_buffers: int = 0) -> Tuple[torch.Tensor, Dict]:
    """ONE streaming pass over `logits` giving what `grpo_loss_fn` (areal/engine/ppo/actor.py:313-391)
    computes with gather_logprobs_entropy + ppo_actor_loss_fn + logits.min/max: returns
    (loss, stat) where stat has the 8 `ppo_actor_loss_fn` entries plus `logprobs`, `entropy`
    (detached, as grpo_loss_fn uses them) and `vocab_min_logits` / `vocab_max_logits`.
    loss.backward() writes dlogits with one more pass (rows with a zero coefficient are not read).

    `loss_scale` (float or 0-dim/1-element tensor): the scalar the caller will multiply the loss by
    before `.backward()` -- what the engines do (`loss *= loss_scale; loss.backward()`,
    areal/engine/fsdp_engine.py:578-586).  When given (bf16, 16-byte aligned rows), ONE kernel reads the
    logits once and writes dlogits = loss_scale * dloss/dlogits in the same pass (rows stay in a
    thread-block cluster's shared memory between the two phases); the returned loss is still
    UNSCALED, and `(loss * loss_scale).backward()` just hands the stored gradient out (it is rescaled,
    correctly but at the cost of an extra pass, if the upstream scalar turns out to differ).  That
    single pass is used only where it applies -- bf16 logits, 16-byte aligned rows, V % 8 == 0, a row
    that fits a cluster of 8 CTAs; every other input silently takes the two passes, with the same
    results.  The single-pass gradient can be back-propagated once (no retain_graph re-use)."""
    if c_clip is not None:
        assert c_clip > 1.0, c_clip
    res = None
    if loss_scale is not None and logits.requires_grad and _single_pass_eligible(logits):
        if not torch.is_tensor(loss_scale):
            loss_scale = torch.full((1,), float(loss_scale), dtype=torch.float32, device=logits.device)
        try:
            res = _FusedActorRowRes.apply(logits, labels, temperature, proximal_logprobs, old_logprobs, advantages,
                                          loss_mask, eps_clip, eps_clip_higher, c_clip, behav_imp_weight_cap,
                                          with_logits_minmax, loss_scale, cluster_size, n_buffers)
        except _lib.ArealB200Error as e:
            # status 4 = the rows cannot be kept on chip (e.g. a vocabulary too wide for a cluster of 8):
            # same result from the two-pass entries
            if "(status 4)" not in str(e):
                raise
    if res is None:
        res = _FusedActor.apply(logits, labels, temperature, proximal_logprobs, old_logprobs, advantages,
                                loss_mask, eps_clip, eps_clip_higher, c_clip, behav_imp_weight_cap,
                                with_logits_minmax)
    it = iter(res)
    loss = next(it)
    stat = dict(logprobs=next(it), entropy=next(it), loss=next(it), importance_weight=next(it),
                approx_kl=next(it), clip_mask=next(it), dual_clip_mask=next(it))
    if proximal_logprobs is not None:
        stat["behave_imp_weight"] = next(it)
        stat["behave_approx_kl"] = next(it)
        stat["behave_mask"] = next(it)
    if with_logits_minmax:
        stat["vocab_min_logits"] = next(it)
        stat["vocab_max_logits"] = next(it)
    return loss, stat
