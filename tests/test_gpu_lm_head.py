"""GPU parity: LM head + actor loss over row blocks (areal_b200/lm_head.py) vs the oracle
and vs the materialised-logits path (`fused_actor_loss(hidden @ weight.T, ...)`)."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu
D = "cuda:0"
CFG = dict(eps_clip=0.4, eps_clip_higher=None, c_clip=3.0, behav_imp_weight_cap=5.0)


def exact_inputs(T, Hd, V, seed):
    """Hidden states / weights on coarse binary grids: every dot product is exact in fp32 whatever the
    GEMM's summation order, so the bf16 logits are the same bits for every block size and on the CPU."""
    g = torch.Generator().manual_seed(seed)
    h = (torch.randint(-16, 17, (T, Hd), generator=g).float() / 8).bfloat16()
    w = (torch.randint(-16, 17, (V, Hd), generator=g).float() / 32).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g)
    logits = (h.double() @ w.double().t()).float().bfloat16()
    lp_ref, ent_ref, _, mn_ref, mx_ref = O.gather_logprobs_entropy(logits.float().numpy(), labels.numpy(), 0.9)
    prox = (torch.from_numpy(lp_ref) + 0.02 * torch.randn(T, generator=g)).float()
    old = (torch.from_numpy(lp_ref) + 0.05 * torch.randn(T, generator=g)).float()
    adv = torch.randn(T, generator=g)
    mask = torch.rand(T, generator=g) > 0.3
    return h, w, labels, logits, prox, old, adv, mask, (lp_ref, ent_ref, mn_ref, mx_ref)


def unfused(h, w, labels, prox, old, adv, mask, scale):
    from areal_b200 import functional as F
    hh = h.to(D).requires_grad_()
    ww = w.to(D).requires_grad_()
    logits = torch.mm(hh, ww.t())
    loss, st = F.fused_actor_loss(logits, labels.to(D), 0.9, prox.to(D), old.to(D), adv.to(D), mask.to(D),
                                  loss_scale=scale, **CFG)
    (loss * scale).backward()
    return loss.detach(), st, hh.grad, ww.grad


@pytest.mark.parametrize("T,Hd,V,chunk", [(333, 64, 4104, 4096), (333, 64, 4104, 100), (700, 64, 8200, 256),
                                         (64, 64, 151936, 48), (5, 64, 4104, 1)])
def test_lm_head_blocks_equal_materialised_logits_and_oracle(T, Hd, V, chunk):
    from areal_b200.lm_head import fused_lm_head_actor_loss
    old_flag = torch.backends.cuda.matmul.allow_bf16_reduced_precision_reduction
    torch.backends.cuda.matmul.allow_bf16_reduced_precision_reduction = False
    try:
        h, w, labels, logits, prox, old, adv, mask, (lp_ref, ent_ref, mn_ref, mx_ref) = exact_inputs(T, Hd, V, T + V)
        scale = 0.37
        hh = h.to(D).requires_grad_()
        ww = w.to(D).requires_grad_()
        loss, st = fused_lm_head_actor_loss(hh, ww, labels.to(D), 0.9, prox.to(D), old.to(D), adv.to(D),
                                            mask.to(D), loss_scale=scale, chunk_rows=chunk, **CFG)
        (loss * scale).backward()
        loss_u, st_u, dh_u, dw_u = unfused(h, w, labels, prox, old, adv, mask, scale)
    finally:
        torch.backends.cuda.matmul.allow_bf16_reduced_precision_reduction = old_flag
    # per-token outputs do not depend on the block size: same logits bits, same row kernel
    for name in st_u:
        if st_u[name].dtype == torch.bool:
            assert torch.equal(st[name], st_u[name]), name
        else:
            assert torch.allclose(st[name], st_u[name], rtol=2e-5, atol=2e-6), name
    assert torch.allclose(loss, loss_u, rtol=5e-6, atol=1e-7)
    # oracle, end to end from the exact logits
    want, st_ref, coef = O.ppo_actor_loss(lp_ref, prox.numpy().astype(np.float64), old.numpy().astype(np.float64),
                                          adv.numpy().astype(np.float64), loss_mask=mask.numpy(), dtype=np.float64,
                                          **CFG)
    assert abs(float(loss) - want) <= H.TOL_LOGITS * max(1.0, abs(want))
    assert np.allclose(st["logprobs"].cpu().numpy(), lp_ref, rtol=H.TOL_LOGITS, atol=H.TOL_LOGITS)
    assert np.allclose(st["entropy"].cpu().numpy(), ent_ref, rtol=H.TOL_LOGITS, atol=H.TOL_LOGITS)
    assert np.array_equal(st["vocab_min_logits"].cpu().numpy(), mn_ref)
    assert np.array_equal(st["vocab_max_logits"].cpu().numpy(), mx_ref)
    assert np.array_equal(st["clip_mask"].cpu().numpy(), st_ref["clip_mask"])
    # gradients: fp64 chain rule through the oracle's dlogits; the GPU feeds bf16 dlogits to the GEMMs
    dl = O.dlogits(logits.float().numpy(), labels.numpy(), 0.9, scale * coef)
    dh_ref = dl @ w.double().numpy()
    dw_ref = dl.T @ h.double().numpy()
    for got, ref, other in ((hh.grad, dh_ref, dh_u), (ww.grad, dw_ref, dw_u)):
        got = got.float().cpu().numpy().astype(np.float64)
        tol = 2 ** -6 * np.abs(ref) + 2 ** -7 * np.abs(ref).max() + 1e-30
        assert np.all(np.abs(got - ref) <= tol)
        assert np.all(np.abs(got - other.float().cpu().numpy()) <= tol)


def test_lm_head_other_upstream_scalar_and_frozen_inputs():
    """A different upstream scalar than announced rescales the stored gradients; inputs that do not
    require a gradient get none."""
    from areal_b200.lm_head import fused_lm_head_actor_loss
    h, w, labels, logits, prox, old, adv, mask, _ = exact_inputs(200, 64, 4104, 7)
    args = (labels.to(D), 0.9, prox.to(D), old.to(D), adv.to(D), mask.to(D))
    a_h, a_w = h.to(D).requires_grad_(), w.to(D).requires_grad_()
    loss_a, _ = fused_lm_head_actor_loss(a_h, a_w, *args, loss_scale=0.5, chunk_rows=64, **CFG)
    (loss_a * 0.5).backward()
    b_h, b_w = h.to(D).requires_grad_(), w.to(D).requires_grad_()
    loss_b, _ = fused_lm_head_actor_loss(b_h, b_w, *args, loss_scale=0.5, chunk_rows=64, **CFG)
    (loss_b * 1.5).backward()
    assert torch.equal(loss_a, loss_b)
    assert torch.allclose(b_h.grad.float() / 3, a_h.grad.float(), rtol=2 ** -6, atol=1e-12)
    assert torch.allclose(b_w.grad.float() / 3, a_w.grad.float(), rtol=2 ** -6, atol=1e-12)
    c_h, c_w = h.to(D), w.to(D).requires_grad_()
    loss_c, _ = fused_lm_head_actor_loss(c_h, c_w, *args, loss_scale=0.5, chunk_rows=64, wgrad_dtype=torch.bfloat16,
                                         **CFG)
    (loss_c * 0.5).backward()
    assert c_h.grad is None
    assert torch.allclose(c_w.grad.float(), a_w.grad.float(), rtol=2 ** -5, atol=2 ** -7 * float(a_w.grad.abs().max()))


def test_lm_head_fully_masked_block_and_batch():
    from areal_b200.lm_head import fused_lm_head_actor_loss
    h, w, labels, logits, prox, old, adv, mask, _ = exact_inputs(192, 64, 4104, 11)
    mask[64:128] = False                      # one whole block without a loss token
    args = (labels.to(D), 0.9, prox.to(D), old.to(D), adv.to(D))
    hh, ww = h.to(D).requires_grad_(), w.to(D).requires_grad_()
    loss, _ = fused_lm_head_actor_loss(hh, ww, *args, mask.to(D), loss_scale=1.0, chunk_rows=64, **CFG)
    loss.backward()
    loss_u, _, dh_u, dw_u = unfused(h, w, labels, prox, old, adv, mask, 1.0)
    assert torch.allclose(loss, loss_u, rtol=5e-6, atol=1e-7)
    assert torch.all(hh.grad[64:128] == 0)
    assert torch.allclose(hh.grad.float(), dh_u.float(), rtol=2 ** -6, atol=2 ** -7 * float(dh_u.abs().max()))
    none = torch.zeros_like(mask)
    h2, w2 = h.to(D).requires_grad_(), w.to(D).requires_grad_()
    loss0, _ = fused_lm_head_actor_loss(h2, w2, *args, none.to(D), loss_scale=1.0, chunk_rows=64, **CFG)
    loss0.backward()
    assert float(loss0) == 0.0 and not torch.any(h2.grad != 0) and not torch.any(w2.grad != 0)


@pytest.mark.parametrize("chunk,with_entropy", [(4096, False), (77, True), (1, True)])
def test_lm_head_logprobs_forward_only(chunk, with_entropy):
    """compute_logp's post hook (areal/engine/ppo/actor.py:57-63) from the hidden state, block by block."""
    from areal_b200.lm_head import lm_head_logprobs
    T = 300 if chunk > 1 else 6
    h, w, labels, logits, *_, (lp_ref, ent_ref, _, _) = exact_inputs(T, 64, 4104, 5 + chunk)
    hh = h.to(D).requires_grad_()
    res = lm_head_logprobs(hh, w.to(D), labels.to(D), 0.9, chunk_rows=chunk, with_entropy=with_entropy, impl="library")
    lp = res[0] if with_entropy else res
    assert not lp.requires_grad and lp.dtype == torch.float32 and lp.shape == (T,)
    assert np.allclose(lp.cpu().numpy(), lp_ref, rtol=H.TOL_LOGITS, atol=H.TOL_LOGITS)
    if with_entropy:
        assert np.allclose(res[1].cpu().numpy(), ent_ref, rtol=H.TOL_LOGITS, atol=H.TOL_LOGITS)
    # same bits as the materialised path
    from areal_b200 import functional as F
    with torch.no_grad():
        want = F.gather_logprobs(logits.to(D), labels.to(D), 0.9)
    assert torch.equal(lp, want)
    # the default route (one tensor-core kernel, no logits buffer): same numbers up to the summation order
    auto = lm_head_logprobs(hh, w.to(D), labels.to(D), 0.9, with_entropy=with_entropy)
    lp2 = auto[0] if with_entropy else auto
    assert not lp2.requires_grad and torch.allclose(lp2, want, rtol=2e-5, atol=2e-5)
    if with_entropy:
        assert np.allclose(auto[1].cpu().numpy(), ent_ref, rtol=H.TOL_LOGITS, atol=H.TOL_LOGITS)


def test_grpo_loss_fn_from_hidden_exports_what_grpo_loss_fn_exports():
    """The loss function on (hidden state, LM-head weight) records the same tracker keys and values as
    grpo_loss_fn on the materialised logits (areal/engine/ppo/actor.py:313-391)."""
    from areal_b200 import stats_tracker as ST
    from areal_b200.actor import grpo_loss_fn, grpo_loss_fn_from_hidden
    T, Hd, V = 500, 64, 8200
    h, w, labels, logits, prox, old, adv, mask, _ = exact_inputs(T, Hd, V, 31)
    mb = dict(input_ids=torch.roll(labels, 1).to(D), logprobs=old.to(D), prox_logp=prox.to(D),
              advantages=adv.to(D), loss_mask=mask.to(D))
    args = dict(temperature=0.9, eps_clip=0.4, eps_clip_higher=None, c_clip=3.0, behav_imp_weight_cap=5.0)
    tr_a, tr_b = ST.DistributedStatsTracker(), ST.DistributedStatsTracker()
    hh, ww = h.to(D).requires_grad_(), w.to(D).requires_grad_()
    with tr_a.scope("ppo_actor"):
        loss_a = grpo_loss_fn_from_hidden(hh, ww, mb, tracker=tr_a, loss_scale=0.5, chunk_rows=128, **args)
    (loss_a * 0.5).backward()
    x = logits.to(D).requires_grad_()
    with tr_b.scope("ppo_actor"):
        loss_b = grpo_loss_fn(x, mb, tracker=tr_b, **args)
    out_a, out_b = tr_a.export(), tr_b.export()
    assert sorted(out_a) == sorted(out_b) and len(out_a) > 20
    for k in out_a:
        assert abs(out_a[k] - out_b[k]) <= 2e-5 * max(1.0, abs(out_b[k])), (k, out_a[k], out_b[k])
    assert torch.allclose(loss_a.detach(), loss_b.detach(), rtol=5e-6, atol=1e-7)
    assert hh.grad is not None and ww.grad is not None


def test_lm_head_empty_batch():
    from areal_b200.lm_head import fused_lm_head_actor_loss
    h = torch.zeros(0, 64, dtype=torch.bfloat16, device=D, requires_grad=True)
    w = torch.ones(4104, 64, dtype=torch.bfloat16, device=D, requires_grad=True)
    e = torch.zeros(0, device=D)
    loss, st = fused_lm_head_actor_loss(h, w, torch.zeros(0, dtype=torch.long, device=D), 1.0, e, e, e,
                                        torch.zeros(0, dtype=torch.bool, device=D), loss_scale=1.0, **CFG)
    loss.backward()
    assert float(loss) == 0.0 and st["logprobs"].shape == (0,)
    assert h.grad.shape == (0, 64) and not torch.any(w.grad != 0)


def test_lm_head_golden_from_the_reference():
    """Golden vectors of the UNMODIFIED reference (oracle/gen_golden.py::gen_lm_head): logits = F.linear(hidden,
    weight), gather_logprobs_entropy, ppo_actor_loss_fn, (loss * scale).backward() down to hidden and weight."""
    from areal_b200.lm_head import fused_lm_head_actor_loss
    z = H.load("lm_head.npz")
    for k, c in H.cases(z):
        h = H.bf16_from_bits(c["h_bits"]).to(D).requires_grad_()
        w = H.bf16_from_bits(c["w_bits"]).to(D).requires_grad_()
        cfg = {name: (float(c["cfg_" + name]) if "cfg_" + name in c else None)
               for name in ("eps_clip", "eps_clip_higher", "c_clip", "behav_imp_weight_cap")}
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(D)
        scale = float(c["scale"])
        loss, st = fused_lm_head_actor_loss(h, w, t(c["labels"]), float(c["temp"]), t(c["prox"]), t(c["old"]),
                                            t(c["adv"]), t(c["mask"]), loss_scale=scale, chunk_rows=64, **cfg)
        (loss * scale).backward()
        want = float(c["loss"])
        assert abs(float(loss.detach()) - want) <= H.TOL_LOGITS * max(1.0, abs(want)), k
        assert np.allclose(st["logprobs"].cpu().numpy(), c["logp"], rtol=H.TOL_LOGITS, atol=H.TOL_LOGITS), k
        assert np.allclose(st["entropy"].cpu().numpy(), c["entropy"], rtol=H.TOL_LOGITS, atol=H.TOL_LOGITS), k
        assert np.allclose(st["importance_weight"].cpu().numpy(), c["importance_weight"], rtol=1e-4, atol=1e-5), k
        assert np.array_equal(st["clip_mask"].cpu().numpy(), c["clip_mask"]), k
        # both sides feed bf16 dlogits to bf16 GEMMs: one bf16 ulp per element, plus the accumulation noise
        for got, bits in ((h.grad, c["dh_bits"]), (w.grad, c["dw_bits"])):
            ref = H.bf16_from_bits(bits).float().numpy().astype(np.float64)
            got = got.float().cpu().numpy().astype(np.float64)
            assert got.shape == ref.shape
            assert np.all(np.abs(got - ref) <= 2 ** -6 * np.abs(ref) + 2 ** -7 * np.abs(ref).max() + 1e-30), k
