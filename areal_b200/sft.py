"""Host-side mirror of `compute_packed_sft_loss` (areal/engine/sft/lm_engine.py:58-110).

The reference gathers log-probs (log-softmax over `[T,V]`), then loops over the sequences in Python with
one `.item()` host sync each for the per-sequence perplexity, then reads the logits twice more for their
min / max.  Here: one streaming pass (log-prob + row min / max), one launch for the per-sequence masked
means, one launch for the tracker keys; the backward is the usual `dlogits` pass.  CUDA only."""
from __future__ import annotations

from typing import Any, Dict

import torch

from . import _lib
from . import functional as F
from ._lib import call, ptr, stream_ptr


class _LogpMinMax(torch.autograd.Function):
    """log-prob of the labels (differentiable) + row max / min of the raw logits from the same pass."""

    @staticmethod
    def forward(ctx, logits, labels):
        dev, x, lab, lead, V = F._rows(logits, labels)
        logp, _, row_max, ln_s, mn = F._logits_forward(x, lab, 1.0, False, True)
        ctx.save_for_backward(x, lab, row_max, ln_s)
        ctx.in_shape = logits.shape
        ctx.set_materialize_grads(False)
        row_max_v, mn_v = row_max.view(lead), mn.view(lead)
        ctx.mark_non_differentiable(row_max_v, mn_v)
        return logp.view(lead), row_max_v, mn_v

    @staticmethod
    def backward(ctx, g_logp, *unused):
        if g_logp is None:
            return None, None
        x, lab, row_max, ln_s = ctx.saved_tensors
        d = F._logits_backward(x, lab, 1.0, row_max, ln_s, None, F._f32(g_logp.reshape(-1)), None)
        return d.view(ctx.in_shape), None


def segment_masked_mean(v: torch.Tensor, mask: torch.Tensor, cu_seqlens: torch.Tensor):
    """(mean f64[B], valid bool[B]): masked mean of `v` over each packed sequence, no host sync."""
    dev = _lib.require_cuda(v, mask, cu_seqlens)
    B = cu_seqlens.numel() - 1
    vf = F._f32(v.reshape(-1))
    mk = mask.reshape(-1).contiguous().view(torch.uint8)
    cu = cu_seqlens.int().contiguous()
    mean = torch.empty(B, dtype=torch.float64, device=dev)
    valid = torch.empty(B, dtype=torch.bool, device=dev)
    with torch.cuda.device(dev):
        call("areal_b200_segment_masked_mean", ptr(vf), ptr(mk), ptr(cu), B, ptr(mean), ptr(valid.view(torch.uint8)),
             stream_ptr(dev))
    return mean, valid


def compute_packed_sft_loss(logits: torch.Tensor, input_: Dict[str, Any], tracker=None) -> torch.Tensor:
    """Drop-in for `compute_packed_sft_loss`: same loss (masked mean negative log-likelihood of the rolled
    labels), same tracker keys (`ppl` per sequence, `loss` per valid token, vocab min / max logits)."""
    from . import stats_tracker as ST
    from .actor import _all_tokens
    tracker = ST.DEFAULT_TRACKER if tracker is None else tracker
    labels = input_.get("rolled_input_ids", None)
    if labels is None:
        labels = torch.roll(input_["input_ids"], shifts=-1, dims=-1)
    cu_seqlens = input_["cu_seqlens"]
    loss_mask = input_.get("full_loss_mask", input_["loss_mask"]).bool()
    logprobs, vmax, vmin = _LogpMinMax.apply(logits, labels)
    loss_mask = torch.roll(loss_mask, shifts=-1, dims=-1)                       # lm_engine.py:71
    logprobs = torch.where(loss_mask, logprobs, 0)                              # :72
    loss = -logprobs.sum() / loss_mask.count_nonzero()                          # :75 (no host sync: stays a tensor)
    with torch.no_grad():
        lp_d = logprobs.detach()
        seqlogp, n_seqs = segment_masked_mean(lp_d, loss_mask, cu_seqlens)      # :77-91 without the loop
        tracker.record(
            dict(n_seqs=n_seqs, n_tokens=_all_tokens(logits.shape[0], logits.device), n_valid_tokens=loss_mask,
                 prompt_tokens=loss_mask.logical_not()),
            [("n_seqs", dict(ppl=(-seqlogp).exp().float())), ("n_valid_tokens", dict(loss=-lp_d)),
             ("n_tokens", dict(vocab_min_logits=vmin.float(), vocab_max_logits=vmax.float()))])
    return loss
