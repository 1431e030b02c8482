"""Host-side mirror of the packed log-prob helpers of `realhf/impl/model/utils/functional.py:91-234`
and of `calc_entropy` (`realhf/impl/model/interface/ppo_interface.py:52-58`) -- the legacy (realhf)
spelling of P2.  Same names / arguments / shapes; the index builders are O(T) torch ops (the
reference's are O(T*B) broadcasts, functional.py:118-126, 162-165), the log-softmax is the fused fp32
streaming kernel (the reference runs log_softmax in the logits' own dtype and returns that dtype:
results are cast back to it, so they agree to one ulp of that dtype)."""
from __future__ import annotations

import torch

from . import _lib
from . import functional as F
from ._lib import call, ptr, stream_ptr


def build_shift_one_indices(x: torch.Tensor, cu_seqlens: torch.Tensor) -> torch.Tensor:
    """Indices of every token except the FIRST of each sequence (functional.py:91-126)."""
    total = x.shape[0]
    keep = torch.ones(total, dtype=torch.bool, device=cu_seqlens.device)
    keep[cu_seqlens[:-1].long()] = False
    return keep.nonzero(as_tuple=True)[0]


def build_leave_one_indices(x: torch.Tensor, cu_seqlens: torch.Tensor) -> torch.Tensor:
    """Indices of every token except the LAST of each sequence (functional.py:129-166)."""
    total = x.shape[0]
    keep = torch.ones(total, dtype=torch.bool, device=cu_seqlens.device)
    keep[cu_seqlens[1:].long() - 1] = False
    return keep.nonzero(as_tuple=True)[0]


def gather_logprobs(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """log_softmax(logits)[labels] in the logits' dtype (functional.py:169-186)."""
    return F.gather_logprobs(logits, labels, 1.0).to(logits.dtype)


def gather_packed_shifted_log_probs(logits: torch.Tensor, cu_seqlens: torch.Tensor,
                                    labels: torch.Tensor) -> torch.Tensor:
    """Log-prob of token t+1 under the logits at t, for every t but the last of each sequence
    (functional.py:189-234); shape [tot_seqlen - #seqs], dtype of `logits`, differentiable."""
    labels = torch.nn.functional.pad(labels[1:], (0, 1), value=0)
    leave_one_indices = build_leave_one_indices(logits, cu_seqlens)
    log_probs_labels = gather_logprobs(logits, labels)[leave_one_indices]
    assert log_probs_labels.shape[0] == logits.shape[0] - cu_seqlens.shape[0] + 1
    return log_probs_labels


@torch.no_grad()
def calc_entropy(logits: torch.Tensor, cu_seqlens: torch.Tensor) -> torch.Tensor:
    """-sum softmax * log(softmax + 1e-7) per token, last token of each sequence dropped
    (ppo_interface.py:52-58), in the logits' dtype."""
    dev, x, lab, lead, V = F._rows(logits.detach(), torch.zeros(logits.shape[:-1], dtype=torch.long, device=logits.device))
    _, _, row_max, ln_s, _ = F._logits_forward(x, lab, 1.0, False)
    T = x.shape[0]
    out = torch.empty(T, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        call("areal_b200_legacy_entropy", ptr(x), _lib.DTYPE_CODE[x.dtype], T, V, x.stride(0) if T > 1 else V,
             ptr(row_max), ptr(ln_s), 1e-7, ptr(out), stream_ptr(dev))
    return out[build_leave_one_indices(logits, cu_seqlens)].to(logits.dtype)


def masked_normalization(x: torch.Tensor, mask=None, dim=None, inplace: bool = False, unbiased: bool = False,
                         eps: float = 1e-5, high_precision: bool = True, all_reduce: bool = True, group=None):
    """Drop-in for the legacy `masked_normalization` (realhf/impl/model/utils/functional.py:252-319), `dim=None`
    (how the PPO interface calls it): fp64 moments, one all-reduce of three numbers instead of three
    all-reduces, fp32 result.  `inplace` is accepted; the reference returns a new fp32 tensor either way
    (`.float()` of its float64 working copy)."""
    from . import data as _D
    return _D.masked_normalization(x, mask, dim=dim, unbiased=unbiased, eps=eps, high_precision=high_precision,
                                   all_reduce=all_reduce, reduce_group=group)


def gather_shifted_log_probs(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """Drop-in for the padded `gather_shifted_log_probs` (realhf/impl/model/utils/functional.py:70-88):
    log-prob of `labels[:, 1:]` under `logits[:, :-1]`, shape [bs, seqlen - 1], in the logits' dtype.
    The logits are streamed where they are (rows [b, s] of the full tensor with the labels rolled by one);
    the last position of every sequence is computed and dropped instead of copying `logits[:, :-1]`."""
    if logits.dim() != 3 or labels.shape != logits.shape[:2]:
        raise RuntimeError(f"expected logits [bs, seqlen, V] and labels [bs, seqlen], got {tuple(logits.shape)} / {tuple(labels.shape)}")
    shifted = torch.roll(labels, shifts=-1, dims=1)
    lp = F.gather_logprobs(logits, shifted, 1.0)
    return lp[:, :-1].to(logits.dtype)
