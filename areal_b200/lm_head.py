"""LM head + actor loss without the `[T, V]` logits tensor (the step before the path).

The reference materialises the full logits (`outputs = self.model(**inputs)`; `logits = outputs.logits`,
areal/engine/fsdp_engine.py:571-578: 4.98 GB of bf16 at T=16,384, V=151,936, and as much again for their
gradient) and hands them to `loss_fn(logits, mb)` (areal/engine/ppo/actor.py:313-391).  Here the projection
runs over blocks of `chunk_rows` tokens:

    logits_blk  = hidden_blk @ weight.T                         library GEMM (cuBLAS through torch.mm)
    areal_b200_fused_actor_fwd_bwd(logits_blk) -> statistics, loss terms, dlogits_blk    (our single pass)
    d_hidden_blk = dlogits_blk @ weight                         library GEMM
    d_weight    += dlogits_blk.T @ hidden_blk                   library GEMM, fp32 accumulator

so only two `[chunk_rows, V]` buffers ever exist.  The masked mean runs over ALL tokens: block i's kernel
normalises by its own mask count n_i, so it is given the scale `loss_scale * n_i / n` (a device scalar, no
host sync) and the block losses are recombined as `sum_i loss_i * n_i / n`.

This changes the engine boundary (the model has to return the last hidden state instead of logits), which
is why it sits after every row of the path proper; INTEGRATION.md shows the call.
CUDA tensors only; there is no CPU / eager fallback.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr
from .functional import _alloc_stats, _f32, _mask_u8, _opt, _u8

_ADDMM_F32_OUT = None  # does torch.addmm accept bf16 operands with an fp32 accumulator here? (probed once)


def _accumulate_wgrad(acc: torch.Tensor, a_t: torch.Tensor, b: torch.Tensor, first: bool) -> None:
    """acc (+)= a_t @ b with acc possibly fp32 while a_t, b are bf16/fp16."""
    global _ADDMM_F32_OUT
    if acc.dtype == a_t.dtype:
        if first:
            torch.mm(a_t, b, out=acc)
        else:
            acc.addmm_(a_t, b)
        return
    if first:
        torch.mm(a_t, b, out_dtype=acc.dtype, out=acc)
        return
    if _ADDMM_F32_OUT is None:
        try:
            torch.addmm(acc, a_t, b, out_dtype=acc.dtype, out=acc)
            _ADDMM_F32_OUT = True
            return
        except (RuntimeError, TypeError):
            _ADDMM_F32_OUT = False
    if _ADDMM_F32_OUT:
        torch.addmm(acc, a_t, b, out_dtype=acc.dtype, out=acc)
    else:
        acc.add_(torch.mm(a_t, b, out_dtype=acc.dtype))


class _LMHeadActor(torch.autograd.Function):
    @staticmethod
    def forward(ctx, hidden, weight, labels, temperature, prox, old, adv, mask, eps_clip, eps_clip_higher,
                c_clip, cap, want_minmax, loss_scale, chunk_rows, wgrad_dtype):
        dev = _lib.require_cuda(hidden, weight, labels, old, adv, mask, loss_scale)
        if hidden.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16:
            raise TypeError("fused_lm_head_actor_loss: bf16 hidden states and weight only")
        lead = hidden.shape[:-1]
        H = hidden.shape[-1]
        V = weight.shape[0]
        if weight.dim() != 2 or weight.shape[1] != H:
            raise RuntimeError(f"weight {tuple(weight.shape)} does not match hidden size {H}")
        if V % 8 != 0 or H % 8 != 0:
            raise RuntimeError(f"vocabulary size {V} and hidden size {H} must be multiples of 8 (16-byte rows)")
        if labels.shape != lead:
            raise RuntimeError(f"labels shape {tuple(labels.shape)} does not match hidden {tuple(hidden.shape)}")
        h = hidden.reshape(-1, H).contiguous()
        w = weight.contiguous()
        T = h.shape[0]
        lab = labels.reshape(-1).long().contiguous()
        px = None if prox is None else _f32(prox.reshape(-1))
        od = _f32(old.reshape(-1))
        av = _f32(adv.reshape(-1))
        mk = _mask_u8(mask.reshape(-1))
        R = max(1, min(int(chunk_rows), T))
        n_blk = (T + R - 1) // R
        f32 = dict(dtype=torch.float32, device=dev)
        logp = torch.empty(T, **f32)
        ent = torch.empty(T, **f32)
        row_max = torch.empty(T, **f32)
        ln_s = torch.empty(T, **f32)
        mn = torch.empty(T, **f32) if want_minmax else None
        coef = torch.empty(T, **f32)
        f, b = _alloc_stats(T, dev, px is not None)
        blk_loss = torch.empty(n_blk, **f32)
        blk_inv = torch.empty(n_blk, **f32)
        # per-block mask counts n_i and the whole-batch count n, on the device
        m32 = mk.to(torch.float32)
        if T % R:
            m32 = torch.nn.functional.pad(m32, (0, n_blk * R - T))
        blk_cnt = m32.view(n_blk, R).sum(1)
        share = blk_cnt / blk_cnt.sum().clamp_min(1.0)
        blk_scale = (share * _f32(loss_scale.reshape(1))).contiguous()

        need_dh, need_dw = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        d_h = torch.empty_like(h) if need_dh else None
        d_w = None
        if need_dw:     # written by the first block's GEMM; an empty batch has no block
            d_w = (torch.empty if n_blk else torch.zeros)((V, H), dtype=wgrad_dtype, device=dev)
        logits = torch.empty((R, V), dtype=torch.bfloat16, device=dev)
        dlogits = torch.empty((R, V), dtype=torch.bfloat16, device=dev)
        w_t = w.t()

        def sl(t, s, e):
            return None if t is None else t[s:e]

        with torch.cuda.device(dev):
            ws = _lib.workspace("loss", dev, _lib.lib().areal_b200_loss_workspace_bytes())
            for i in range(n_blk):
                s, e = i * R, min((i + 1) * R, T)
                n = e - s
                x, d = logits[:n], dlogits[:n]
                torch.mm(h[s:e], w_t, out=x)
                call("areal_b200_fused_actor_fwd_bwd", ptr(x), _lib.DTYPE_CODE[x.dtype], n, V, V,
                     ptr(lab[s:e]), float(temperature), ptr(sl(px, s, e)), ptr(od[s:e]), ptr(av[s:e]),
                     ptr(mk[s:e]), float(eps_clip), _opt(eps_clip_higher), _opt(c_clip), _opt(cap),
                     ptr(logp[s:e]), ptr(ent[s:e]), ptr(row_max[s:e]), ptr(ln_s[s:e]), ptr(sl(mn, s, e)),
                     ptr(blk_loss[i:i + 1]), ptr(blk_inv[i:i + 1]), ptr(coef[s:e]),
                     ptr(f["loss"][s:e]), ptr(f["importance_weight"][s:e]), ptr(f["approx_kl"][s:e]),
                     ptr(_u8(b["clip_mask"])[s:e]), ptr(_u8(b["dual_clip_mask"])[s:e]),
                     ptr(sl(f.get("behave_imp_weight"), s, e)), ptr(sl(f.get("behave_approx_kl"), s, e)),
                     ptr(sl(_u8(b.get("behave_mask")), s, e)), ptr(blk_scale[i:i + 1]), ptr(d), V, 0, 0,
                     ptr(ws), ws.numel(), stream_ptr(dev))
                if need_dh:
                    torch.mm(d, w, out=d_h[s:e])
                if need_dw:
                    _accumulate_wgrad(d_w, d.t(), h[s:e], first=(i == 0))
        loss = (blk_loss * share).sum()
        if need_dw and d_w.dtype != weight.dtype:
            d_w = d_w.to(weight.dtype)
        ctx.save_for_backward(d_h, d_w, _f32(loss_scale.reshape(1)))
        ctx.h_shape = hidden.shape
        ctx.set_materialize_grads(False)
        extras = [logp, ent, f["loss"], f["importance_weight"], f["approx_kl"], b["clip_mask"],
                  b["dual_clip_mask"]]
        if px is not None:
            extras += [f["behave_imp_weight"], f["behave_approx_kl"], b["behave_mask"]]
        if want_minmax:
            extras += [mn, row_max]
        extras = [x.view(lead) for x in extras]
        ctx.mark_non_differentiable(*extras)
        return (loss, *extras)

    @staticmethod
    def backward(ctx, g_loss, *unused):
        if g_loss is None:
            return (None,) * 16
        d_h, d_w, scale = ctx.saved_tensors
        g = _f32(g_loss.reshape(1))
        out = []
        for d in (d_h, d_w):
            if d is None:
                out.append(None)
                continue
            with torch.cuda.device(d.device):
                # returns at once when the upstream scalar equals the announced loss_scale (the contract)
                call("areal_b200_rescale_grad_if_needed", ptr(d), d.shape[0], d.shape[1], d.stride(0), ptr(g),
                     ptr(scale), stream_ptr(d.device))
            out.append(d)
        return (out[0].view(ctx.h_shape) if out[0] is not None else None, out[1]) + (None,) * 14


def fused_lm_head_actor_loss(hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                             temperature: float, proximal_logprobs: Optional[torch.Tensor],
                             old_logprobs: torch.Tensor, advantages: torch.Tensor, loss_mask: torch.Tensor,
                             eps_clip: float, eps_clip_higher: Optional[float] = None,
                             c_clip: Optional[float] = None, behav_imp_weight_cap: Optional[float] = None,
                             with_logits_minmax: bool = True, loss_scale=1.0, chunk_rows: int = 4096,
                             wgrad_dtype: torch.dtype = torch.float32) -> Tuple[torch.Tensor, Dict]:
    """`fused_actor_loss(hidden @ weight.T, ...)` without ever holding `[T, V]` logits or their gradient.

    hidden `[T, H]` bf16 (the model's last hidden state, padding rows already cut off), weight `[V, H]`
    bf16 (`lm_head.weight`; no bias, as in the Qwen2 / Llama heads the reference trains).  Returns the
    same `(loss, stat)` as `fused_actor_loss`.  The gradients w.r.t. hidden and weight are computed in
    the forward, already multiplied by `loss_scale` (float or 1-element tensor; the scalar the engine
    multiplies the loss by before `.backward()`, areal/engine/fsdp_engine.py:578-586; must not be 0);
    `(loss * loss_scale).backward()` hands them out, rescaling them if the upstream scalar differs.

    chunk_rows: tokens per block; peak extra memory is `2 * chunk_rows * V * 2` bytes (2.5 GB at 4096 and
    V = 151,936, against 10 GB for logits + dlogits at T = 16,384).  wgrad_dtype: accumulator of the
    weight gradient across blocks (fp32 by default: one rounding to bf16 at the end, like the single
    GEMM of the unfused path)."""
    if c_clip is not None:
        assert c_clip > 1.0, c_clip
    if not torch.is_tensor(loss_scale):
        loss_scale = torch.full((1,), float(loss_scale), dtype=torch.float32, device=hidden.device)
    res = _LMHeadActor.apply(hidden, weight, labels, temperature, proximal_logprobs, old_logprobs, advantages,
                             loss_mask, eps_clip, eps_clip_higher, c_clip, behav_imp_weight_cap,
                             with_logits_minmax, loss_scale, chunk_rows, wgrad_dtype)
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


@torch.no_grad()
def lm_head_logprobs(hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                     temperature: float = 1.0, chunk_rows: int = 4096, with_entropy: bool = False, impl: str = "auto"):
    """Forward-only companion: what `PPOActor.compute_logp`'s post hook computes from the logits
    (`gather_logprobs(logits, labels, temperature)`, areal/engine/ppo/actor.py:57-63; the recompute of
    proximal / reference log-probs runs under no_grad), from the last hidden state, `chunk_rows` tokens
    at a time: one `[chunk_rows, V]` logits buffer instead of `[T, V]`.  Returns logp `[T]` f32, or
    (logp, entropy) with `with_entropy=True`.  `impl`: "auto" takes the tcgen05 kernel (`lm_head_logprobs_tc`) for
    bf16 operands with a hidden size of at most 2,048 (measured at T = 16,384, V = 151,936: H = 1,536 5.6-6.5 ms against
    5.9-6.1 ms for the library path, the order depending on the power state of the part; H = 3,584 13.9-15.0 ms against
    12.4-13.2 ms: the kernel streams its operands out of L2 at the L2's limit, 128 flop per byte, where the library's
    larger effective tiles gain with K), "library" forces the row-blocked cuBLAS + streaming path, "tcgen05" the
    kernel (no `[chunk_rows, V]` buffer at all)."""
    from .functional import _logits_forward
    dev = _lib.require_cuda(hidden, weight, labels)
    if impl not in ("auto", "library", "tcgen05"):
        raise ValueError(f"impl must be 'auto', 'library' or 'tcgen05', got {impl!r}")
    # bf16 with 16-byte aligned rows: ONE tensor-core kernel (tcgen05 GEMM + online-softmax epilogue, no logits
    # buffer at all); everything else (fp16 / fp32, odd hidden sizes): library GEMM per row block + streaming pass
    tc_ok = (hidden.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and hidden.shape[-1] % 8 == 0)
    if impl == "tcgen05" or (impl == "auto" and tc_ok and hidden.shape[-1] <= 2048):
        return lm_head_logprobs_tc(hidden, weight, labels, temperature, with_entropy=with_entropy)
    if hidden.dtype != weight.dtype or hidden.dtype not in _lib.DTYPE_CODE:
        raise TypeError("lm_head_logprobs: hidden and weight must share a dtype (bf16, fp16 or fp32)")
    lead = hidden.shape[:-1]
    if labels.shape != lead:
        raise RuntimeError(f"labels shape {tuple(labels.shape)} does not match hidden {tuple(hidden.shape)}")
    H, V = hidden.shape[-1], weight.shape[0]
    h = hidden.reshape(-1, H).contiguous()
    lab = labels.reshape(-1).long().contiguous()
    T = h.shape[0]
    R = max(1, min(int(chunk_rows), T))
    logits = torch.empty((R, V), dtype=hidden.dtype, device=dev)
    w_t = weight.t()
    lps, ents = [], []
    for s in range(0, T, R):
        e = min(s + R, T)
        x = logits[:e - s]
        torch.mm(h[s:e], w_t, out=x)
        res = _logits_forward(x, lab[s:e], temperature, with_entropy)
        lps.append(res[0])
        if with_entropy:
            ents.append(res[1])
    logp = (torch.cat(lps) if lps else torch.empty(0, dtype=torch.float32, device=dev)).view(lead)
    if with_entropy:
        return logp, torch.cat(ents).view(lead)
    return logp


@torch.no_grad()
def lm_head_logprobs_tc(hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor, temperature: float = 1.0,
                        with_entropy: bool = False, with_stats: bool = False):
    """`lm_head_logprobs` as ONE tensor-core kernel (areal_b200_lm_head_logprobs): the projection runs on
    tcgen05 with the accumulators in tensor memory and the online softmax as its epilogue, so neither the
    `[T, V]` logits nor a `[chunk_rows, V]` block of them ever exists in memory.  bf16 only (hidden `[T, H]`,
    weight `[V, H]`, H % 8 == 0).  Returns logp `[T]` f32, (logp, entropy) with `with_entropy=True`, or with
    `with_stats=True` a dict(logp, entropy, vocab_min_logits, vocab_max_logits, lse) like `functional.logits_stats`."""
    dev = _lib.require_cuda(hidden, weight, labels)
    if hidden.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16:
        raise TypeError("lm_head_logprobs_tc: hidden and weight must be bfloat16")
    lead = hidden.shape[:-1]
    if labels.shape != lead:
        raise RuntimeError(f"labels shape {tuple(labels.shape)} does not match hidden {tuple(hidden.shape)}")
    H, V = hidden.shape[-1], weight.shape[0]
    if weight.dim() != 2 or weight.shape[1] != H:
        raise RuntimeError(f"weight shape {tuple(weight.shape)} does not match hidden size {H}")
    h = hidden.reshape(-1, H)
    if h.stride(-1) != 1 or h.stride(0) % 8 != 0 or h.data_ptr() % 16 != 0:
        h = h.contiguous()
    w = weight if (weight.stride(1) == 1 and weight.stride(0) % 8 == 0 and weight.data_ptr() % 16 == 0) else weight.contiguous()
    if H % 8 != 0:
        raise RuntimeError("lm_head_logprobs_tc: the hidden size must be a multiple of 8")
    lab = labels.reshape(-1).long().contiguous()
    T = h.shape[0]
    f32 = dict(dtype=torch.float32, device=dev)
    logp, row_max, ln_s = torch.empty(T, **f32), torch.empty(T, **f32), torch.empty(T, **f32)
    ent = torch.empty(T, **f32) if (with_entropy or with_stats) else None
    mn = torch.empty(T, **f32) if with_stats else None
    with torch.cuda.device(dev):
        nbytes = int(_lib.lib().areal_b200_lm_head_workspace_bytes(T, V))
        ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=dev)
        _lib.call("areal_b200_lm_head_logprobs", _lib.ptr(h), _lib.ptr(w), _lib.ptr(lab), T, H, V,
                  h.stride(0) if T > 1 else H, w.stride(0) if V > 1 else H, float(temperature), _lib.ptr(logp), _lib.ptr(ent),
                  _lib.ptr(row_max), _lib.ptr(ln_s), _lib.ptr(mn), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev))
    if with_stats:
        return dict(logp=logp.view(lead), entropy=ent.view(lead), lse=(row_max / temperature + ln_s).view(lead),
                    vocab_min_logits=mn.view(lead), vocab_max_logits=row_max.view(lead))
    if with_entropy:
        return logp.view(lead), ent.view(lead)
    return logp.view(lead)
