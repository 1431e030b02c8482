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
        saved = torch.empty(T, 6, **f32)
        with torch.cuda.device(dev):
            call("areal_b200_clamped_entropy_fwd", ptr(x), _lib.DTYPE_CODE[x.dtype], T, V, x.stride(0) if T > 1 else V,
                 ptr(lab), float(temperature), k, ptr(logp), ptr(ent), ptr(saved), stream_ptr(dev))
        ctx.save_for_backward(x, lab, saved)
        ctx.temperature = float(temperature)
        ctx.in_shape = logits.shape
        ctx.set_materialize_grads(False)
        return logp.view(lead), ent.view(lead)

    @staticmethod
    def backward(ctx, g_logp, g_ent):
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
    return _ClampedEntropy.apply(logits, labels, entropy_clamp, temperature)


def clamped_softmax_entropy(logits: torch.Tensor, entropy_clamp: float) -> torch.Tensor:
    """Drop-in for `clamped_softmax_entropy` (recipe/AEnt/functional.py:15-36); logits already divided
    by the temperature, as the reference calls it (:48-50)."""
    labels = torch.zeros(logits.shape[:-1], dtype=torch.long, device=logits.device)
    return _ClampedEntropy.apply(logits, labels, entropy_clamp, 1.0)[1]
