"""Host-side mirror of the GAE / legacy-loss entry points of
`realhf/impl/model/utils/ppo_functional.py` and of the native module `realhf._C.cugae`
(csrc/cugae/gae.cu), plus the padded masked GAE that `PPOActor.compute_advantages`
(areal/engine/ppo/actor.py:131-152) runs as a Python loop.  CUDA tensors only, no fallback."""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import _lib
from . import functional as F
from ._lib import call, ptr, stream_ptr


def _check(cond, msg):
    if not cond:
        raise RuntimeError(msg)   # TORCH_CHECK in the reference raises RuntimeError (gae.cu:36-50)


def gae_1d_nolp_misalign(rewards: torch.Tensor, values: torch.Tensor, cu_seqlens: torch.Tensor,
                         bootstrap: torch.Tensor, gamma: float, lmbda: float):
    """`realhf._C.cugae.gae_1d_nolp_misalign` (csrc/cugae/gae.cu:30-60, pybind :210) and
    `cugae.cugae1d_nolp_misalign_func`: same checks except the two `.item()` host syncs on
    cu_seqlens[0] / cu_seqlens[-1] (gae.cu:48-49), which the Python caller already asserts
    (ppo_functional.py:365).  Returns fresh (advantages, returns); inputs are not modified."""
    for name, t in (("rewards", rewards), ("values", values), ("cu_seqlens", cu_seqlens), ("bootstrap", bootstrap)):
        _check(t.is_cuda, f"{name} must be on CUDA")
        _check(t.is_contiguous(), f"{name} must be contiguous")
    bs = cu_seqlens.numel() - 1
    T = rewards.size(0)
    _check(tuple(bootstrap.shape) == (bs,), "bootstrap must have shape (batch_size)")
    _check(tuple(values.shape) == (T + bs,), "values must have shape (total_seqlen + batch_size)")
    _check(bootstrap.dtype == torch.bool, "bootstrap must be bool")
    _check(cu_seqlens.dtype == torch.int32, "cu_seqlens must be int32")
    _check(rewards.dtype == values.dtype, "rewards and values must have the same dtype")
    _check(rewards.dtype == torch.float32, "rewards must be float32")
    dev = rewards.device
    adv = torch.empty_like(rewards)
    ret = torch.empty_like(rewards)
    with torch.cuda.device(dev):
        nbytes = _lib.lib().areal_b200_gae1d_workspace_bytes(T)
        ws = _lib.workspace("gae1d", dev, nbytes)
        call("areal_b200_gae1d_nolp_misalign", ptr(rewards), ptr(values), ptr(cu_seqlens),
             ptr(bootstrap.view(torch.uint8)), T, bs, float(gamma), float(lmbda), ptr(adv), ptr(ret),
             ptr(ws), ws.numel(), stream_ptr(dev))
    return adv, ret


def cugae1d_nolp_misalign_func(rewards: torch.Tensor, values: torch.Tensor, cu_seqlens: torch.Tensor,
                               truncate: torch.Tensor, gamma: float, lam: float):
    """Drop-in for `cugae1d_nolp_misalign_func` (ppo_functional.py:326-366)."""
    assert len(rewards.shape) == len(values.shape) == len(cu_seqlens.shape) == 1
    return gae_1d_nolp_misalign(rewards, values, cu_seqlens, truncate, gamma, lam)


@torch.no_grad()
def get_packed_advantages_and_returns(gamma: float, lam: float, values: torch.Tensor, rewards: torch.Tensor,
                                      short1cu_seqlens: torch.Tensor, seq_no_eos_mask: torch.Tensor):
    """Drop-in for `get_packed_advantages_and_returns` (ppo_functional.py:369-394); CUDA only --
    the reference's CPU branch (pygae1d_nolp_misalign) is deliberately not reproduced here."""
    if rewards.get_device() == -1:
        raise _lib.ArealB200Error("areal_b200 has no CPU path; move the batch to the GPU")
    return cugae1d_nolp_misalign_func(rewards.contiguous(), values.contiguous(),
                                      short1cu_seqlens.int().contiguous(),
                                      seq_no_eos_mask.bool().contiguous(), gamma, lam)


def _packed_rewards(mode, kl_ctl, clip_reward_value, log_probs, ref_log_probs, score, short1cu_seqlens,
                    seq_no_eos_mask, mask_no_eos_with_zero):
    dev = _lib.require_cuda(log_probs, ref_log_probs, score, short1cu_seqlens)
    lp = log_probs.float().contiguous()
    rf = ref_log_probs.float().contiguous()
    _check(lp.dim() == 1 and lp.shape == rf.shape, "log_probs / ref_log_probs must be 1D with equal shapes")
    cu = short1cu_seqlens.int().contiguous()
    B, Ts = cu.numel() - 1, lp.numel()
    sc = score.float().contiguous()
    _check(sc.numel() == (B if mode == 0 else Ts + B), "reward tensor does not match the sequence table")
    ne = None
    if seq_no_eos_mask is not None and mode == 0:
        ne = seq_no_eos_mask.bool().contiguous().view(torch.uint8)
    kl = torch.empty_like(lp)
    tot = torch.empty_like(lp)
    with torch.cuda.device(dev):
        call("areal_b200_packed_rewards", ptr(lp), ptr(rf), ptr(sc), ptr(cu), ptr(ne), Ts, B, float(kl_ctl),
             float(clip_reward_value), int(bool(mask_no_eos_with_zero)), mode, ptr(kl), ptr(tot), stream_ptr(dev))
    return kl, tot


@torch.no_grad()
def get_packed_rewards(kl_ctl: float, clip_reward_value: float, log_probs: torch.Tensor, ref_log_probs: torch.Tensor,
                       reward_score: torch.Tensor, short1cu_seqlens: torch.Tensor, seq_no_eos_mask: torch.Tensor,
                       mask_no_eos_with_zero: bool = False) -> Tuple[torch.Tensor, torch.Tensor]:
    """Drop-in for `get_packed_rewards` (ppo_functional.py:228-253): (kl_rewards, tot_rewards), one launch."""
    return _packed_rewards(0, kl_ctl, clip_reward_value, log_probs, ref_log_probs, reward_score, short1cu_seqlens,
                           seq_no_eos_mask, mask_no_eos_with_zero)


@torch.no_grad()
def get_packed_reward_dense(kl_ctl: float, clip_reward_value: float, log_probs: torch.Tensor,
                            ref_log_probs: torch.Tensor, dense_reward_score: torch.Tensor,
                            short1cu_seqlens: torch.Tensor, seq_no_eos_mask: torch.Tensor, reward_delta: bool):
    """Drop-in for `get_packed_reward_dense` (ppo_functional.py:256-289): the reference walks the
    sequences in a Python loop (a host sync per sequence); here one launch."""
    return _packed_rewards(2 if reward_delta == True else 1, kl_ctl, clip_reward_value, log_probs,  # noqa: E712
                           ref_log_probs, dense_reward_score, short1cu_seqlens, None, False)


@torch.no_grad()
def gae_padded_masked(rewards: torch.Tensor, values: torch.Tensor, loss_mask: torch.Tensor,
                      seq_no_eos_mask: torch.Tensor, discount: float, gae_lambda: float):
    """The GAE block of `PPOActor.compute_advantages` (areal/engine/ppo/actor.py:131-152) as one
    launch: rewards / values / loss_mask are [B,S] (loss_mask float, already rolled by -1),
    seq_no_eos_mask bool [B].  Returns (advantages, returns) -- `returns` is what the caller stores
    as data["returns"] (:152)."""
    dev = _lib.require_cuda(rewards, values, loss_mask, seq_no_eos_mask)
    _check(rewards.dim() == 2 and rewards.shape == values.shape == loss_mask.shape, "expected [B,S] inputs")
    B, S = rewards.shape
    r = rewards.float().contiguous()
    v = values.float().contiguous()
    m = loss_mask.float().contiguous()
    ne = seq_no_eos_mask.bool().contiguous().view(torch.uint8)
    adv = torch.empty_like(r)
    ret = torch.empty_like(r)
    with torch.cuda.device(dev):
        call("areal_b200_gae_padded_masked", ptr(r), ptr(v), ptr(m), ptr(ne), B, S, float(discount),
             float(gae_lambda), ptr(adv), ptr(ret), stream_ptr(dev))
    return adv, ret


def gae_2d_olp(rewards, values, done, done_y_indices, cu_num_dones, max_num_dones, is_truncate, gamma, lmbda):
    """`realhf._C.cugae.gae_2d_olp` (csrc/cugae/gae.cu:91-134, pybind :213).  `max_num_dones` only
    sized the reference's launch grid and is ignored."""
    for name, t in (("rewards", rewards), ("values", values), ("done", done), ("done_y_indices", done_y_indices),
                    ("cu_num_dones", cu_num_dones), ("is_truncate", is_truncate)):
        _check(t.is_cuda, f"{name} must be on CUDA")
        _check(t.is_contiguous(), f"{name} must be contiguous")
    B, H = rewards.shape
    nd = done_y_indices.numel()
    _check(tuple(values.shape) == (B, H + 1), "values must have shape (batch_size, horizon + 1)")
    _check(tuple(done.shape) == (B, H + 1), "done must have shape (batch_size, horizon + 1)")
    _check(tuple(cu_num_dones.shape) == (B + 1,), "cu_num_dones must have shape (batch_size + 1)")
    _check(tuple(is_truncate.shape) == (nd,), "is_truncate must have shape (num_dones)")
    _check(done_y_indices.dtype == torch.int32, "done_y_indices must be int32")
    _check(cu_num_dones.dtype == torch.int32, "cu_num_dones must be int32")
    _check(is_truncate.dtype == torch.bool, "is_truncate must be bool")
    _check(done.dtype == torch.bool, "done must be bool")
    _check(rewards.dtype == values.dtype == torch.float32, "rewards and values must be float32")
    dev = rewards.device
    adv = torch.empty_like(rewards)
    ret = torch.empty_like(rewards)
    with torch.cuda.device(dev):
        call("areal_b200_gae2d_olp", ptr(rewards), ptr(values), ptr(done.view(torch.uint8)), ptr(done_y_indices),
             ptr(cu_num_dones), ptr(is_truncate.view(torch.uint8)), B, H, nd, float(gamma), float(lmbda),
             ptr(adv), ptr(ret), stream_ptr(dev))
    return adv, ret


def gae_2d_nolp(rewards, values, on_reset, on_reset_y_indices, cu_num_resets, max_num_resets, bootstrap,
                gamma, lmbda):
    """`realhf._C.cugae.gae_2d_nolp` (gae.cu:163-206, pybind :215).  The reference's host wrapper
    launches `gae_kernel_2d_olp` with these arguments (gae.cu:199), so this does exactly that."""
    return gae_2d_olp(rewards, values, on_reset, on_reset_y_indices, cu_num_resets, max_num_resets, bootstrap,
                      gamma, lmbda)


def actor_loss_fn(logprobs: torch.Tensor, old_logprobs: torch.Tensor, advantages: torch.Tensor, eps_clip: float,
                  loss_mask: Optional[torch.Tensor] = None, c_clip: Optional[float] = None,
                  proximal_logprobs: Optional[torch.Tensor] = None,
                  behav_imp_weight_cap: Optional[float] = None) -> Tuple[torch.Tensor, Dict]:
    """Legacy `actor_loss_fn` (ppo_functional.py:51-149): same math as the areal one, without
    eps_clip_higher; proximal_logprobs / loss_mask optional."""
    assert logprobs.dtype == torch.float32
    assert old_logprobs.dtype == torch.float32
    assert advantages.dtype == torch.float32
    if proximal_logprobs is not None:
        assert proximal_logprobs.dtype == torch.float32
    if c_clip is not None:
        assert c_clip > 1.0, c_clip
    res = F._ActorLoss.apply(logprobs, proximal_logprobs, old_logprobs, advantages, loss_mask, eps_clip, None,
                             c_clip, behav_imp_weight_cap if proximal_logprobs is not None else None)
    stat = dict(loss=res[1], importance_weight=res[2], approx_kl=res[3], clip_mask=res[4], dual_clip_mask=res[5])
    if proximal_logprobs is not None:
        stat["behave_imp_weight"] = res[6]
        stat["behave_approx_kl"] = res[7]
        stat["behave_mask"] = res[8]
    return res[0], stat


def critic_loss_fn(value, old_value, target_value, value_eps_clip: float, loss_mask=None,
                   loss_fn_type: str = "mse"):
    """Legacy `critic_loss_fn` (ppo_functional.py:161-225) == the areal one."""
    return F.ppo_critic_loss_fn(value, old_value, target_value, value_eps_clip, loss_mask, loss_fn_type)
