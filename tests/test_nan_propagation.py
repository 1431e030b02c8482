"""NaN inputs through the clamps and max/min of the loss and reward kernels.

torch.clamp / torch.max / torch.min (what the reference uses: areal/utils/functional.py:193-207, 291-297,
areal/utils/data.py:1363, areal/engine/ppo/actor.py:99-101) return NaN when an operand is NaN; CUDA's
fminf / fmaxf return the other operand.  tests/golden/nan_cases.npz holds outputs of the unmodified
reference functions on inputs with NaNs planted (oracle/gen_golden.py:gen_nan_cases); the oracle is
checked against them on the CPU and the kernels on the GPU: same NaN pattern, same finite values."""
import types

import numpy as np
import pytest
import torch

from oracle import cpu_oracle as O
from tests import adv_cases as A
from tests import helpers as H

F_STATS = ("loss", "importance_weight", "approx_kl", "behave_imp_weight", "behave_approx_kl")
B_STATS = ("clip_mask", "dual_clip_mask", "behave_mask")


def same(got, want, what, rtol=1e-5, atol=1e-6):
    got, want = np.asarray(got), np.asarray(want)
    assert np.array_equal(np.isnan(got), np.isnan(want)), (what, np.argwhere(np.isnan(got) != np.isnan(want))[:5])
    ok = ~np.isnan(want)
    assert np.allclose(got[ok], want[ok], rtol=rtol, atol=atol), what


def actor_cases():
    z = H.load("nan_cases.npz")
    return list(H.cases(z, prefix="a", count_key="n_actor"))


def critic_cases():
    z = H.load("nan_cases.npz")
    return list(H.cases(z, prefix="v", count_key="n_critic"))


def adv_inputs(ci):
    z = H.load("nan_cases.npz")
    ins = {k: z[f"c{ci}_in_{k}"] for k in A.IN_KEYS if f"c{ci}_in_{k}" in z}
    outs = {k: z[f"c{ci}_out_{k}"] for k in A.OUT_KEYS if f"c{ci}_out_{k}" in z}
    return A.CASES[ci], ins, outs


def test_golden_has_nans_where_planted():
    """The fixture really exercises NaN: most planted cases give a NaN loss in the reference; a NaN behaviour
    log-prob under an importance-weight cap does not (`w <= cap` is False, the token drops out), and a NaN
    under a False loss mask does not."""
    n_nan = 0
    for k, c in actor_cases():
        if not bool(c["mask"][3]):
            assert np.isfinite(float(c["loss"])), k
        n_nan += int(np.isnan(float(c["loss"])))
    assert n_nan >= 6
    for k, c in critic_cases():
        assert np.isnan(float(c["loss"])), k


def test_oracle_actor_critic_nan_vs_reference():
    for k, c in actor_cases():
        eps, eps_hi, c_clip, cap = (H.opt(x) for x in c["cfg"])
        with np.errstate(invalid="ignore"):
            loss, st, grad = O.ppo_actor_loss(c["lp"], c["prox"], c["old"], c["adv"], eps, c["mask"],
                                              eps_clip_higher=eps_hi, c_clip=c_clip, behav_imp_weight_cap=cap)
        same(loss, c["loss"], (k, "loss"))
        same(grad, c["grad"], (k, "grad"), atol=1e-9)
        for name in F_STATS:
            if "stat_" + name in c:
                same(st[name], c["stat_" + name], (k, name))
        for name in B_STATS:
            if "stat_" + name in c:
                assert np.array_equal(st[name], c["stat_" + name]), (k, name)
    for k, c in critic_cases():
        with np.errstate(invalid="ignore"):
            loss, st, grad = O.ppo_critic_loss(c["value"], c["old"], c["target"], 0.5, c["mask"],
                                               "huber" if bool(c["huber"]) else "mse")
        same(loss, c["loss"], (k, "loss"))
        same(grad, c["grad"], (k, "grad"), atol=1e-9)
        same(st["loss"], c["stat_loss"], (k, "stat_loss"))
        assert np.array_equal(st["clip_mask"], c["stat_clip_mask"]), k


@pytest.mark.parametrize("ci", [0, 1, 2])
def test_oracle_advantages_nan_vs_reference(ci):
    c, ins, outs = adv_inputs(ci)
    with np.errstate(invalid="ignore"):
        r = O.compute_advantages(ins["rewards"], ins["loss_mask"], ins["logprobs"], ins["attention_mask"],
                                 ref_logp=ins.get("ref_logp"), values=ins.get("values"), roll_logp=True,
                                 reward_bias=c["reward_bias"], reward_scaling=c["reward_scaling"],
                                 reward_clip=c["reward_clip"], kl_ctl=c["kl_ctl"], kl_estimator=c["kl_estimator"],
                                 mask_no_eos_with_zero=c["mask_no_eos_with_zero"], discount=c["discount"],
                                 gae_lambda=c["gae_lambda"])
    assert np.isnan(outs["kl_rewards"]).any() and np.isnan(outs["tot_rewards"]).any()
    for k in ("kl_rewards", "tot_rewards", "advantages", "returns"):
        same(r[k], outs[k], (ci, k), rtol=1e-5, atol=1e-5)


# ---------------------------------------------------------------------------------------------- GPU
def t(a, dev="cuda:0"):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


@pytest.mark.gpu
def test_gpu_actor_loss_nan_vs_reference():
    from areal_b200 import functional as F
    for k, c in actor_cases():
        eps, eps_hi, c_clip, cap = (H.opt(x) for x in c["cfg"])
        lp = t(c["lp"]).requires_grad_()
        loss, st = F.ppo_actor_loss_fn(lp, t(c["prox"]), t(c["old"]), t(c["adv"]), eps, t(c["mask"]),
                                       eps_clip_higher=eps_hi, c_clip=c_clip, behav_imp_weight_cap=cap)
        same(loss.detach().cpu().numpy(), c["loss"], (k, "loss"))
        loss.backward()
        same(lp.grad.cpu().numpy(), c["grad"], (k, "grad"), atol=1e-9)
        for name in F_STATS:
            if "stat_" + name in c:
                same(st[name].cpu().numpy(), c["stat_" + name], (k, name))
        for name in B_STATS:
            if "stat_" + name in c:
                assert np.array_equal(st[name].cpu().numpy(), c["stat_" + name]), (k, name)


@pytest.mark.gpu
def test_gpu_critic_loss_nan_vs_reference():
    from areal_b200 import functional as F
    for k, c in critic_cases():
        v = t(c["value"]).requires_grad_()
        loss, st = F.ppo_critic_loss_fn(v, t(c["old"]), t(c["target"]), 0.5, t(c["mask"]),
                                        "huber" if bool(c["huber"]) else "mse")
        same(loss.detach().cpu().numpy(), c["loss"], (k, "loss"))
        loss.backward()
        same(v.grad.cpu().numpy(), c["grad"], (k, "grad"), atol=1e-9)
        same(st["loss"].cpu().numpy(), c["stat_loss"], (k, "stat_loss"))
        assert np.array_equal(st["clip_mask"].cpu().numpy(), c["stat_clip_mask"]), k


@pytest.mark.gpu
@pytest.mark.parametrize("ci", [0, 1, 2])
def test_gpu_compute_advantages_nan_vs_reference(ci):
    from areal_b200.actor import compute_advantages
    c, ins, outs = adv_inputs(ci)
    data = {k: t(v.copy()) for k, v in ins.items()}
    actor = types.SimpleNamespace(
        config=A.config_of(c), reward_bias=c["reward_bias"], reward_scaling=c["reward_scaling"],
        reward_clip=c["reward_clip"], kl_ctl=c["kl_ctl"],
        kl_estimator=types.SimpleNamespace(kl_estimator=c["kl_estimator"], apply_clamp=True), discount=c["discount"],
        gae_lambda=c["gae_lambda"], mask_no_eos_with_zero=c["mask_no_eos_with_zero"], adv_norm=None, reward_norm=None)
    compute_advantages(actor, data)
    for k in ("kl_rewards", "tot_rewards", "advantages", "returns"):
        same(data[k].cpu().numpy(), outs[k], (ci, k), rtol=1e-5, atol=1e-5)
