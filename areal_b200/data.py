"""Host-side mirror of the normalisation entry points next to the hot path:
`Normalization` / `NormConfig` (areal/utils/data.py:1073-1304, areal/api/cli_args.py:24-59) and
`masked_normalization` (areal/utils/functional.py:130-168).  Same constructor checks, call signature and
results; three kernel launches instead of a Python loop over groups, no `.item()` host syncs, and ONE
all-reduce per statistic under data parallelism.  CUDA tensors only."""
from __future__ import annotations

import dataclasses
from typing import Optional

import torch
import torch.distributed as dist

from . import _lib
from ._lib import call, ptr, stream_ptr

_MODE = {"batch": 0, "group": 1, None: 2}


@dataclasses.dataclass
class NormConfig:
    """Field names and defaults of the reference's NormConfig (areal/api/cli_args.py:24-59)."""
    mean_level: Optional[str] = "batch"
    mean_leave1out: bool = False
    std_level: Optional[str] = "batch"
    std_unbiased: bool = True
    eps: float = 1e-5
    group_size: int = 1


def _f32c(t: torch.Tensor) -> torch.Tensor:
    return t.float().contiguous()


def _moments(x, mask, n_groups, group_elems):
    dev = x.device
    P = _lib.lib().areal_b200_norm_num_partials(group_elems)
    partials = torch.empty(n_groups, P, 3, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        call("areal_b200_masked_moments", ptr(x), ptr(mask), n_groups, group_elems, P, ptr(partials), stream_ptr(dev))
    return partials.sum(1)


class Normalization:
    """Adaptive normalisation with independent levels for mean and std ("batch", "group", None);
    drop-in for the reference class (areal/utils/data.py:1073-1204)."""

    def __init__(self, config):
        if config.mean_level not in {"batch", "group", None}:
            raise ValueError(f"mean_level must be 'batch', 'group' or None, got {config.mean_level}")
        if config.std_level not in {"batch", "group", None}:
            raise ValueError(f"std_level must be 'batch', 'group', or None, got {config.std_level}")
        if (config.mean_level == "group" or config.std_level == "group") and config.group_size is None:
            raise ValueError("group_size must be provided if using group normalization")
        self.mean_level = config.mean_level
        self.mean_leave1out = config.mean_leave1out
        self.std_level = config.std_level
        self.std_unbiased = config.std_unbiased
        self.group_size = config.group_size
        self.eps = config.eps

    @torch.no_grad()
    def __call__(self, x: torch.Tensor, loss_mask: Optional[torch.Tensor] = None, high_precision: bool = True,
                 reduce_group=None) -> torch.Tensor:
        """`high_precision` is accepted for compatibility; statistics are always accumulated in fp64."""
        dev = _lib.require_cuda(x, loss_mask)
        shape = x.shape
        bs = x.size(0)
        xf = _f32c(x).reshape(-1)
        mf = None if loss_mask is None else _f32c(loss_mask).reshape(-1)
        N = xf.numel()
        row = N // bs if bs else 0
        total_mask = None if mf is None else mf.sum().reshape(1)            # device scalar, no sync
        world = dist.get_world_size(reduce_group) if dist.is_initialized() else 1

        def grouping(level):
            if level == "batch":
                return 1, N
            if level == "group":
                return bs // self.group_size, self.group_size * row
            return 0, 0

        # ---- mean statistics
        gm, gem = grouping(self.mean_level)
        mean_stats = None
        if self.mean_level is not None and gm > 0:
            mean_stats = _moments(xf, mf, gm, gem)
            if self.mean_level == "batch" and world > 1:
                dist.all_reduce(mean_stats, group=reduce_group)
        mean_zero_special = int(self.mean_level == "group" and self.group_size == 1 and self.mean_leave1out)
        mean_args = (_MODE[self.mean_level] if gm > 0 or self.mean_level is None else 2,
                     int(self.mean_leave1out), mean_zero_special, gem, gm, ptr(mean_stats))
        if self.mean_level == "group" and gm == 0:
            mean_args = (2, 0, 0, 0, 0, None)          # no complete group: zeros_like, as upstream

        # ---- std statistics
        gs, ges = grouping(self.std_level)
        std_stats = None
        if self.std_level is not None and gs > 0:
            P = _lib.lib().areal_b200_norm_num_partials(ges)
            partials = torch.empty(gs, P, 2, dtype=torch.float64, device=dev)
            with torch.cuda.device(dev):
                call("areal_b200_norm_centered_sumsq", ptr(xf), ptr(mf), *mean_args, gs, ges, P, ptr(partials),
                     stream_ptr(dev))
            std_stats = partials.sum(1)
            if self.std_level == "batch" and world > 1:
                dist.all_reduce(std_stats, group=reduce_group)
        std_one_special = int(self.std_level == "group" and self.group_size == 1 and self.std_unbiased)
        std_mode = _MODE[self.std_level]
        if self.std_level == "group" and gs == 0:
            std_stats = torch.zeros(1, 2, dtype=torch.float64, device=dev)     # std stays zeros_like
            ges, gs = max(N, 1), 0

        out = torch.empty(N, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            call("areal_b200_norm_apply", ptr(xf), ptr(mf), N, *mean_args, std_mode, int(self.std_unbiased),
                 std_one_special, ges, gs, ptr(std_stats), float(self.eps), ptr(total_mask), ptr(out), stream_ptr(dev))
        return out.view(shape)


@torch.no_grad()
def masked_normalization(x: torch.Tensor, mask: Optional[torch.Tensor] = None, dim=None, unbiased: bool = False,
                         eps: float = 1e-5, high_precision: bool = True, all_reduce: bool = True, reduce_group=None):
    """Drop-in for `masked_normalization` (areal/utils/functional.py:130-168) for `dim=None`: fp64
    moments of x*mask (one all-reduce of a 3-vector instead of three), output in fp32."""
    if dim is not None and tuple(dim) != tuple(range(x.dim())):
        raise NotImplementedError("areal_b200.masked_normalization supports dim=None (all dimensions) only")
    dev = _lib.require_cuda(x, mask)
    shape = x.shape
    xf = _f32c(x).reshape(-1)
    mf = None if mask is None else _f32c(mask).reshape(-1)
    N = xf.numel()
    stats = _moments(xf, mf, 1, N).reshape(3)
    if dist.is_initialized() and all_reduce and dist.get_world_size(reduce_group) > 1:
        dist.all_reduce(stats, group=reduce_group)
    out = torch.empty(N, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        call("areal_b200_masked_normalization_apply", ptr(xf), ptr(mf), N, ptr(stats), int(unbiased), float(eps),
             ptr(out), stream_ptr(dev))
    return out.view(shape)


# --------------------------------------------------------------------------------------------------
# packing glue: padded [B, S, ...] <-> packed [total, ...] through the interval kernels
# --------------------------------------------------------------------------------------------------
def _row_intervals(lens: torch.Tensor, S: int, inner: int) -> torch.Tensor:
    starts = torch.arange(lens.numel(), device=lens.device, dtype=torch.int64) * (S * inner)
    return torch.stack([starts, starts + lens.to(torch.int64) * inner], dim=1).contiguous()


def pack_tensor_dict(data: dict) -> dict:
    """Drop-in for `pack_tensor_dict` (areal/utils/data.py:266-315): every tensor of shape [B, S, ...]
    becomes [total_length, ...] (the first `len_i` positions of row i, in order), "attention_mask" is
    replaced by "cu_seqlens" (+ "max_seqlen"), other entries pass through.  One gather launch per
    tensor and ONE host sync in total (the reference loops over sequences in Python with two `.item()`
    calls per sequence and per key)."""
    from . import flatten_param as fp
    assert "attention_mask" in data, "Input data must contain 'attention_mask' key."
    attention_mask = data["attention_mask"]
    assert attention_mask.ndim == 2, "Attention mask must be a 2D tensor."
    _lib.require_cuda(attention_mask)
    bs, seq_len = attention_mask.shape
    lens = attention_mask.sum(dim=1, dtype=torch.int32)
    cu_seqlens = torch.nn.functional.pad(torch.cumsum(lens, dim=0, dtype=torch.int32), (1, 0), value=0)
    max_seqlen, total_length = (int(v) for v in torch.stack([lens.max(), cu_seqlens[-1]]).tolist()) if bs else (0, 0)
    packed = {}
    iv_cache = {}
    for key, value in data.items():
        if key == "attention_mask":
            packed["cu_seqlens"] = cu_seqlens
            packed["max_seqlen"] = max_seqlen
            continue
        if torch.is_tensor(value) and value.ndim >= 2 and value.shape[0] == bs and value.shape[1] == seq_len:
            inner = 1
            for d in value.shape[2:]:
                inner *= int(d)
            if inner not in iv_cache:
                iv_cache[inner] = _row_intervals(lens, seq_len, inner)
            flat = value.contiguous().view(-1)
            as_u8 = flat.dtype == torch.bool
            out = fp.slice_intervals_any(flat.view(torch.uint8) if as_u8 else flat, iv_cache[inner], total_length * inner)
            if as_u8:
                out = out.view(torch.bool)
            packed[key] = out.view(total_length, *value.shape[2:])
        else:
            packed[key] = value
    return packed


def unpack_to_padded(packed: torch.Tensor, cu_seqlens: torch.Tensor, seq_len: int, fill=0) -> torch.Tensor:
    """Inverse of the packing above for one tensor: [total, ...] -> [B, seq_len, ...], `fill` elsewhere
    (what `pad_and_stack_tensors_along_first_dim(unpack_sequence(...))` produces, data.py:318-330)."""
    from . import flatten_param as fp
    _lib.require_cuda(packed, cu_seqlens)
    lens = (cu_seqlens[1:] - cu_seqlens[:-1])
    bs = lens.numel()
    inner = 1
    for d in packed.shape[1:]:
        inner *= int(d)
    out = torch.full((bs, seq_len, *packed.shape[1:]), fill, dtype=packed.dtype, device=packed.device)
    iv = _row_intervals(lens, seq_len, inner)
    src = packed.contiguous().view(-1)
    dst = out.view(-1)
    if packed.dtype == torch.bool:
        src, dst = src.view(torch.uint8), dst.view(torch.uint8)
    fp.set_intervals_any(src, dst, iv)
    return out


# --------------------------------------------------------------------------------------------------
# dynamic sampling (areal/utils/functional.py:314-373): drop the groups whose rewards are all equal
# --------------------------------------------------------------------------------------------------
def dynamic_sampling(data: dict, group_size: int):
    """Drop-in for `dynamic_sampling` (areal/utils/functional.py:314-373): samples of a group are adjacent;
    groups whose rewards are all equal are removed from every tensor whose first dimension is the batch.
    Same return value (filtered dict, stats dict) and the same early exits.  The reference indexes each
    tensor with a boolean mask (a `nonzero` + host sync per key, plus two `.item()`s); here ONE host sync
    reads the kept groups and each tensor is gathered by one launch of the interval kernels (kept groups
    are contiguous row ranges)."""
    import warnings

    from . import flatten_param as fp
    rewards = data["rewards"]
    if not torch.is_tensor(rewards):
        raise TypeError("data['rewards'] must be a torch.Tensor")
    batch_size = rewards.shape[0]
    if group_size <= 0:
        warnings.warn("group_size <= 0; returning original data")
        return data, dict(n_group_kept=0, n_group_filtered=0)
    if batch_size % group_size != 0:
        warnings.warn("The group size is not divisible by the batch size. Return the original data")
        return data, dict(n_group_kept=batch_size // max(group_size, 1), n_group_filtered=0)
    _lib.require_cuda(rewards)
    num_groups = batch_size // group_size
    r = rewards.view(num_groups, group_size)
    valid = ~(r == r[:, 0:1]).all(dim=1)
    kept = torch.nonzero(valid).flatten().tolist()          # the one host sync
    if not kept:
        return data, dict(n_group_kept=0, n_group_filtered=num_groups)
    n_kept = len(kept)
    starts = torch.tensor(kept, dtype=torch.int64, device=rewards.device) * group_size     # first row of each kept group
    filtered = {}
    iv_cache = {}
    for k, v in data.items():
        if torch.is_tensor(v) and v.shape[:1] == (batch_size,):
            if n_kept == num_groups:
                filtered[k] = v.clone()                      # v[mask] with an all-true mask copies
                continue
            inner = 1
            for d in v.shape[1:]:
                inner *= int(d)
            if inner not in iv_cache:
                iv_cache[inner] = torch.stack([starts * inner, (starts + group_size) * inner], dim=1).contiguous()
            flat = v.contiguous().view(-1)
            as_u8 = flat.dtype == torch.bool
            out = fp.slice_intervals_any(flat.view(torch.uint8) if as_u8 else flat, iv_cache[inner],
                                         n_kept * group_size * inner)
            if as_u8:
                out = out.view(torch.bool)
            filtered[k] = out.view(n_kept * group_size, *v.shape[1:])
        else:
            filtered[k] = v
    return filtered, dict(n_group_kept=n_kept, n_group_filtered=int(num_groups - n_kept))
