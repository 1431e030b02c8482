"""GPU parity: PPO actor / critic losses (stand-alone and fused with the logits pass) through the
C ABI vs golden vectors of the unmodified reference and its dual-clip known-answer test."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu
D = "cuda:0"


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(D)


def check_stats(st, c, k):
    for name in ("loss", "importance_weight", "approx_kl", "behave_imp_weight", "behave_approx_kl"):
        if "stat_" + name in c:
            assert st[name].dtype == torch.float32
            assert np.allclose(st[name].cpu().numpy(), c["stat_" + name], rtol=1e-5, atol=1e-6), (k, name)
    for name in ("clip_mask", "dual_clip_mask", "behave_mask"):
        if "stat_" + name in c:
            assert st[name].dtype == torch.bool
            assert np.array_equal(st[name].cpu().numpy(), c["stat_" + name]), (k, name)


def test_actor_loss_golden():
    from areal_b200 import functional as F, ppo_functional as P
    z = H.load("actor_loss.npz")
    for k, c in H.cases(z):
        eps, eps_hi, c_clip, cap = (H.opt(x) for x in c["cfg"])
        lp = t(c["lp"]).requires_grad_()
        loss, st = F.ppo_actor_loss_fn(lp, t(c["prox"]), t(c["old"]), t(c["adv"]), eps, t(c["mask"]),
                                       eps_clip_higher=eps_hi, c_clip=c_clip, behav_imp_weight_cap=cap)
        assert loss.dim() == 0 and loss.dtype == torch.float32
        assert abs(float(loss) - float(c["loss"])) <= H.TOL_LOGITS * max(1.0, abs(float(c["loss"]))), k
        assert abs(float(loss) - float(c["loss"])) <= 2e-6 * max(1.0, abs(float(c["loss"]))), k
        (loss * 3.0).backward()
        assert np.allclose(lp.grad.cpu().numpy(), 3.0 * c["grad"], rtol=1e-5, atol=1e-9), k
        check_stats(st, c, k)
        assert set(st) == {"loss", "importance_weight", "approx_kl", "clip_mask", "dual_clip_mask",
                           "behave_imp_weight", "behave_approx_kl", "behave_mask"}
    for k, c in H.cases(z, prefix="l", count_key="n_legacy"):
        eps, c_clip, cap = (H.opt(x) for x in c["cfg"])
        prox = t(c["prox"]) if bool(c["has_prox"]) else None
        lp = t(c["lp"]).requires_grad_()
        loss, st = P.actor_loss_fn(lp, t(c["old"]), t(c["adv"]), eps, loss_mask=t(c["mask"]), c_clip=c_clip,
                                   proximal_logprobs=prox, behav_imp_weight_cap=cap)
        assert abs(float(loss) - float(c["loss"])) <= 2e-6 * max(1.0, abs(float(c["loss"]))), k
        loss.backward()
        assert np.allclose(lp.grad.cpu().numpy(), c["grad"], rtol=1e-5, atol=1e-9), k
        check_stats(st, c, k)
        assert ("behave_mask" in st) == bool(c["has_prox"])


@pytest.mark.parametrize("eps_clip", [0.01, 0.05, 0.1, 0.2, 0.5])
@pytest.mark.parametrize("c_clip", [None, 1.5, 2.0, 5.0])
@pytest.mark.parametrize("size", [1, 10, 100])
def test_dual_clip_known_answer(size, eps_clip, c_clip):
    """Grid + independent formula of realhf/tests/data/test_dual_clip.py:53-63 (legacy signature,
    no mask, no proximal)."""
    from areal_b200 import ppo_functional as P
    rng = np.random.RandomState(size * 7 + int(eps_clip * 100))
    old = -np.abs(rng.randn(size)).astype(np.float32)
    new = -np.abs(rng.randn(size)).astype(np.float32)
    adv = rng.randn(size).astype(np.float32)
    want = O.ppo_dual_clip_known_answer(new, old, adv, eps_clip, c_clip)
    loss, _ = P.actor_loss_fn(t(new), t(old), t(adv), eps_clip, c_clip=c_clip)
    assert np.isclose(float(loss), want, rtol=1e-5, atol=1e-7)


def test_actor_loss_large_random_vs_oracle_and_2d_inputs():
    from areal_b200 import functional as F
    rng = np.random.RandomState(0)
    T = 200003
    lp0 = -rng.rand(T).astype(np.float32) * 5
    lp = (lp0 + 0.3 * rng.randn(T)).astype(np.float32)
    prox = (lp0 + 0.02 * rng.randn(T)).astype(np.float32)
    old = (lp0 + 0.05 * rng.randn(T)).astype(np.float32)
    adv = rng.randn(T).astype(np.float32)
    mask = rng.rand(T) > 0.3
    for cfg in (dict(eps_clip=0.4, behav_imp_weight_cap=5.0), dict(eps_clip=0.2, eps_clip_higher=0.28, c_clip=3.0)):
        want, st_ref, g_ref = O.ppo_actor_loss(lp, prox, old, adv, loss_mask=mask, dtype=np.float64, **cfg)
        x = t(lp).requires_grad_()
        loss, st = F.ppo_actor_loss_fn(x, t(prox), t(old), t(adv), loss_mask=t(mask), **cfg)
        assert abs(float(loss) - want) <= 1e-5 * max(1.0, abs(want))
        loss.backward()
        assert np.allclose(x.grad.cpu().numpy(), g_ref, rtol=1e-4, atol=1e-10)
        l2, _ = F.ppo_actor_loss_fn(t(lp), t(prox), t(old), t(adv), loss_mask=t(mask), **cfg)
        assert torch.equal(l2, loss.detach())          # fixed-order reduction: bit-reproducible
    # [bs, seqlen] inputs keep their shape in the stats
    sh = (7, 11)
    loss, st = F.ppo_actor_loss_fn(t(lp[:77].reshape(sh)), t(prox[:77].reshape(sh)), t(old[:77].reshape(sh)),
                                   t(adv[:77].reshape(sh)), 0.2, t(mask[:77].reshape(sh)))
    assert st["loss"].shape == sh and st["clip_mask"].shape == sh


def test_critic_loss_golden():
    from areal_b200 import functional as F, ppo_functional as P
    z = H.load("critic_loss.npz")
    for k, c in H.cases(z):
        mask = t(c["mask"]) if bool(c["has_mask"]) else None
        kind = "huber" if bool(c["huber"]) else "mse"
        for fn in (F.ppo_critic_loss_fn, P.critic_loss_fn):
            v = t(c["value"]).requires_grad_()
            loss, st = fn(v, t(c["old"]), t(c["target"]), float(c["eps"]), loss_mask=mask, loss_fn_type=kind)
            assert abs(float(loss) - float(c["loss"])) <= 2e-6 * max(1.0, abs(float(c["loss"]))), k
            loss.backward()
            assert np.allclose(v.grad.cpu().numpy(), c["grad"], rtol=1e-5, atol=1e-9), k
            assert np.array_equal(st["clip_mask"].cpu().numpy(), c["stat_clip_mask"]), k
            assert np.allclose(st["loss"].cpu().numpy(), c["stat_loss"], rtol=1e-6, atol=1e-7), k
    with pytest.raises(NotImplementedError):
        F.ppo_critic_loss_fn(t(np.zeros(3, np.float32)), t(np.zeros(3, np.float32)), t(np.zeros(3, np.float32)), 0.5,
                             loss_fn_type="l1")
    with pytest.raises(AssertionError):
        F.ppo_critic_loss_fn(t(np.zeros(3, np.float64)), t(np.zeros(3, np.float32)), t(np.zeros(3, np.float32)), 0.5)
    # all-false mask: nan like the reference (no `or 1`, functional.py:306)
    loss, _ = F.ppo_critic_loss_fn(t(np.ones(3, np.float32)), t(np.zeros(3, np.float32)), t(np.zeros(3, np.float32)),
                                   0.5, loss_mask=t(np.zeros(3, bool)))
    assert np.isnan(float(loss))


@pytest.mark.parametrize("V", [151936, 4099])
def test_fused_actor_equals_split_path_and_oracle(V):
    """fused_actor_loss == gather_logprobs_entropy + ppo_actor_loss_fn + logits.min/max (what
    grpo_loss_fn computes, areal/engine/ppo/actor.py:313-391), incl. loss.backward() -> dlogits."""
    from areal_b200 import functional as F
    g = torch.Generator().manual_seed(V)
    T = 40 if V > 10000 else 333
    logits = (torch.randn(T, V, generator=g) * 2).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g)
    lp_ref, ent_ref, _, mn_ref, mx_ref = O.gather_logprobs_entropy(logits.float().numpy(), labels.numpy(), 0.9)
    prox = (torch.from_numpy(lp_ref) + 0.02 * torch.randn(T, generator=g)).float()
    old = (torch.from_numpy(lp_ref) + 0.05 * torch.randn(T, generator=g)).float()
    adv = torch.randn(T, generator=g)
    mask = torch.rand(T, generator=g) > 0.3
    cfg = dict(eps_clip=0.4, eps_clip_higher=None, c_clip=3.0, behav_imp_weight_cap=5.0)
    x1 = logits.to(D).requires_grad_()
    loss1, st1 = F.fused_actor_loss(x1, labels.to(D), 0.9, prox.to(D), old.to(D), adv.to(D), mask.to(D), **cfg)
    (loss1 * 2.0).backward()
    x2 = logits.to(D).requires_grad_()
    lp2, ent2 = F.gather_logprobs_entropy(x2, labels.to(D), 0.9)
    loss2, st2 = F.ppo_actor_loss_fn(lp2, prox.to(D), old.to(D), adv.to(D), loss_mask=mask.to(D), **cfg)
    (loss2 * 2.0).backward()
    assert torch.allclose(loss1, loss2, rtol=1e-6, atol=1e-7)
    assert torch.equal(st1["logprobs"], lp2.detach()) and torch.equal(st1["entropy"], ent2.detach())
    for name in st2:
        assert torch.equal(st1[name], st2[name]), name
    # same math, scalars folded in a different order: at most one bf16 ulp apart
    assert torch.allclose(x1.grad.float(), x2.grad.float(), rtol=2 ** -7, atol=1e-12)
    assert np.array_equal(st1["vocab_min_logits"].cpu().numpy(), mn_ref)
    assert np.array_equal(st1["vocab_max_logits"].cpu().numpy(), mx_ref)
    # against the oracle end to end
    want, st_ref, coef = O.ppo_actor_loss(lp_ref, prox.numpy().astype(np.float64), old.numpy().astype(np.float64),
                                          adv.numpy().astype(np.float64), loss_mask=mask.numpy(), dtype=np.float64,
                                          **cfg)
    assert abs(float(loss1) - want) <= H.TOL_LOGITS * max(1.0, abs(want))
    ref_grad = O.dlogits(logits.float().numpy(), labels.numpy(), 0.9, 2.0 * coef)
    got = x1.grad.float().cpu().numpy().astype(np.float64)
    scale = np.maximum(np.abs(ref_grad).max(axis=-1, keepdims=True), 1e-3 * np.abs(2.0 * coef)[:, None] / 0.9) + 1e-30
    assert np.all(np.abs(got - ref_grad) <= 2 ** -8 * np.abs(ref_grad) + 2 * H.TOL_LOGITS * scale + 1e-30)


@pytest.mark.parametrize("V,cluster,nbuf", [(151936, 0, 0), (151936, 8, 3), (151936, 4, 3), (151936, 3, 2),
                                           (151936, 4, 2), (32768, 0, 0), (4096, 0, 0), (8200, 2, 2), (8200, 5, 3)])
def test_single_pass_row_resident_equals_two_pass(V, cluster, nbuf):
    """fused_actor_loss(..., loss_scale=s): one kernel, rows resident in a thread-block cluster; must
    give the two-pass path's loss, stats and (loss*s).backward() gradient."""
    from areal_b200 import functional as F
    g = torch.Generator().manual_seed(V + cluster)
    T = 150 if V > 100000 else 333
    logits = (torch.randn(T, V, generator=g) * 2).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g)
    labels[:3] = torch.tensor([0, V - 1, V // 2])
    lp_ref, *_ = O.gather_logprobs_entropy(logits.float().numpy(), labels.numpy(), 0.9)
    prox = (torch.from_numpy(lp_ref) + 0.02 * torch.randn(T, generator=g)).float()
    old = (torch.from_numpy(lp_ref) + 0.05 * torch.randn(T, generator=g)).float()
    adv = torch.randn(T, generator=g)
    mask = torch.rand(T, generator=g) > 0.3
    cfg = dict(eps_clip=0.4, eps_clip_higher=None, c_clip=3.0, behav_imp_weight_cap=5.0)
    scale = 0.37
    x1 = logits.to(D).requires_grad_()
    loss1, st1 = F.fused_actor_loss(x1, labels.to(D), 0.9, prox.to(D), old.to(D), adv.to(D), mask.to(D),
                                    loss_scale=scale, cluster_size=cluster, n_buffers=nbuf, **cfg)
    (loss1 * scale).backward()
    x2 = logits.to(D).requires_grad_()
    loss2, st2 = F.fused_actor_loss(x2, labels.to(D), 0.9, prox.to(D), old.to(D), adv.to(D), mask.to(D), **cfg)
    (loss2 * scale).backward()
    assert torch.allclose(loss1, loss2, rtol=2e-6, atol=1e-7)
    for name in st2:
        a, b = st1[name], st2[name]
        if a.dtype == torch.bool:
            assert torch.equal(a, b), name
        else:
            assert torch.allclose(a, b, rtol=2e-5, atol=2e-6), name
    assert torch.allclose(x1.grad.float(), x2.grad.float(), rtol=2 ** -7, atol=1e-10)
    # a different upstream scalar than announced: still correct (rescaled in place)
    x3 = logits.to(D).requires_grad_()
    loss3, _ = F.fused_actor_loss(x3, labels.to(D), 0.9, prox.to(D), old.to(D), adv.to(D), mask.to(D),
                                  loss_scale=scale, cluster_size=cluster, n_buffers=nbuf, **cfg)
    (loss3 * 1.5).backward()
    assert torch.allclose(x3.grad.float() * (scale / 1.5), x2.grad.float(), rtol=2 ** -6, atol=1e-10)


def test_single_pass_peaked_rows_keep_relative_accuracy_at_the_label():
    """p_label -> 1: the label's gradient gl * (1 - p) is a cancellation; the single-pass kernel writes the
    other elements as 2^(x*c2 + bias) with the scale folded into the exponent, but must keep the label's
    element right to a few percent RELATIVE (same bar as the two-pass test in test_gpu_logits.py)."""
    from areal_b200 import functional as F
    g = torch.Generator().manual_seed(22)
    T, V = 12, 151936
    logits = (torch.randn(T, V, generator=g) * 2.0).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g)
    for t in range(T):
        logits[t, labels[t]] = 21.0 + 0.5 * t      # 1 - p_label from ~8e-4 down to ~3e-6
    lp_ref, *_ = O.gather_logprobs_entropy(logits.float().numpy(), labels.numpy(), 1.0)
    old = torch.from_numpy(lp_ref).float()
    adv = torch.randn(T, generator=g).sign() * (0.5 + torch.rand(T, generator=g))    # both signs of gl
    mask = torch.ones(T, dtype=torch.bool)
    scale = 0.7
    x = logits.to(D).requires_grad_()
    loss, _ = F.fused_actor_loss(x, labels.to(D), 1.0, None, old.to(D), adv.to(D), mask.to(D), eps_clip=0.2,
                                 loss_scale=scale)
    (loss * scale).backward()
    _, _, coef = O.ppo_actor_loss(lp_ref, None, old.numpy().astype(np.float64), adv.numpy().astype(np.float64),
                                  eps_clip=0.2, loss_mask=mask.numpy(), dtype=np.float64)
    ref = O.dlogits(logits.float().numpy(), labels.numpy(), 1.0, scale * coef)
    got = x.grad.float().cpu().numpy().astype(np.float64)
    lab = labels.numpy()
    r_lab = ref[np.arange(T), lab]
    assert np.all(np.abs(got[np.arange(T), lab] - r_lab) <= 0.05 * np.abs(r_lab) + 1e-12)
    # everything else: bf16 rounding of the reference plus 1e-3 of the row scale
    rs = np.abs(ref).max(-1, keepdims=True)
    assert np.all(np.abs(got - ref) <= 2 ** -8 * np.abs(ref) + 0.05 * np.abs(r_lab)[:, None] * (np.arange(V)[None, :] == lab[:, None])
                  + H.TOL_LOGITS * 1e-3 * rs + 1e-30)


def test_single_pass_request_falls_back_where_it_does_not_apply():
    """loss_scale given, but the rows cannot stay on chip (fp16 / fp32 logits, rows that are not 16-byte
    aligned): the same results come from the two-pass entries instead of an 'unsupported' error."""
    from areal_b200 import functional as F
    g = torch.Generator().manual_seed(5)
    T, V = 40, 4104
    base = torch.randn(T, V, generator=g) * 2
    labels = torch.randint(0, V - 8, (T,), generator=g).to(D)
    old = -torch.rand(T, generator=g).to(D) * 6
    adv = torch.randn(T, generator=g).to(D)
    mask = (torch.rand(T, generator=g) > 0.3).to(D)
    for make in (lambda: base.half().to(D), lambda: base.to(D), lambda: base.bfloat16().to(D)[:, 1:V - 7]):
        a = make().detach().requires_grad_()
        b = make().detach().requires_grad_()
        la, sa = F.fused_actor_loss(a, labels, 1.0, None, old, adv, mask, eps_clip=0.2, loss_scale=0.5)
        lb, sb = F.fused_actor_loss(b, labels, 1.0, None, old, adv, mask, eps_clip=0.2)
        (la * 0.5).backward()
        (lb * 0.5).backward()
        assert torch.equal(la, lb) and torch.equal(sa["logprobs"], sb["logprobs"])
        assert torch.equal(a.grad, b.grad)


def test_single_pass_gradient_is_handed_out_once_and_bad_cluster_is_rejected():
    from areal_b200 import functional as F
    from areal_b200._lib import ArealB200Error
    g = torch.Generator().manual_seed(6)
    T, V = 16, 4096
    x = (torch.randn(T, V, generator=g) * 2).bfloat16().to(D).requires_grad_()
    labels = torch.randint(0, V, (T,), generator=g).to(D)
    old = -torch.rand(T, generator=g).to(D) * 6
    adv = torch.randn(T, generator=g).to(D)
    mask = torch.ones(T, dtype=torch.bool, device=D)
    loss, _ = F.fused_actor_loss(x, labels, 1.0, None, old, adv, mask, eps_clip=0.2, loss_scale=0.5)
    (loss * 0.5).backward(retain_graph=True)
    first = x.grad.clone()
    x.grad += 1.0                                    # the caller owns the buffer now
    with pytest.raises(RuntimeError, match="handed"):
        (loss * 0.5).backward()
    assert torch.equal(x.grad, first + 1.0)
    # a forced cluster that would leave a CTA without any part of the row: an argument error, not a hang
    tiny = (torch.randn(4, 32, generator=g)).bfloat16().to(D).requires_grad_()      # 4 vectors per row
    with pytest.raises(ArealB200Error, match="cluster_size"):
        F.fused_actor_loss(tiny, labels[:4] % 32, 1.0, None, old[:4], adv[:4], mask[:4], eps_clip=0.2, loss_scale=0.5,
                           cluster_size=8)
