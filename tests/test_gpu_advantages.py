"""GPU parity: the fused reward construction + padded GAE (areal_b200_advantages_padded) behind the
mirror of PPOActor.compute_advantages, against golden outputs of the unmodified reference method and
against the pinned oracle at larger shapes."""
import types

import numpy as np
import pytest
import torch

from oracle import cpu_oracle as O
from tests import adv_cases as A
from tests import helpers as H

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def make_actor(c):
    from areal_b200.data import NormConfig, Normalization
    return types.SimpleNamespace(
        config=A.config_of(c), reward_bias=c["reward_bias"], reward_scaling=c["reward_scaling"],
        reward_clip=c["reward_clip"], kl_ctl=c["kl_ctl"],
        kl_estimator=types.SimpleNamespace(kl_estimator=c["kl_estimator"], apply_clamp=True), discount=c["discount"],
        gae_lambda=c["gae_lambda"], mask_no_eos_with_zero=c["mask_no_eos_with_zero"],
        adv_norm=Normalization(NormConfig(mean_level="batch", std_level="batch")) if c["adv_norm"] else None,
        reward_norm=Normalization(NormConfig(mean_level="group", std_level="group", group_size=4)) if c["reward_norm"] else None)


@pytest.mark.parametrize("ci", range(len(A.CASES)))
def test_compute_advantages_golden(ci):
    from areal_b200.actor import compute_advantages
    c, ins, outs = A.load()[ci]
    data = {k: torch.from_numpy(v.copy()).to(DEV) for k, v in ins.items()}
    ref_in = data.get("ref_logp")
    rewards_in = data["rewards"]
    compute_advantages(make_actor(c), data)
    got = {k: data[k].cpu().numpy() for k in outs if k in data}
    assert got["loss_mask"].dtype == np.float32
    for k in ("loss_mask", "logprobs"):
        assert np.array_equal(got[k], outs[k]), k
    if "ref_logp" in outs:
        assert data["ref_logp"] is ref_in and np.array_equal(got["ref_logp"], outs["ref_logp"])   # masked in place
    assert data["rewards"] is rewards_in and np.array_equal(got["rewards"], outs["rewards"])        # penalty in place
    if "prox_logp" in outs:
        assert np.array_equal(got["prox_logp"], outs["prox_logp"])
    exact = c["kl_estimator"] != "k3" and not c["reward_norm"]
    for k in ("kl_rewards", "tot_rewards"):
        if exact:
            assert np.array_equal(got[k], outs[k]), k
        else:
            np.testing.assert_allclose(got[k], outs[k], rtol=1e-5, atol=1e-6)
    A.assert_gae_close(got["returns"], outs["returns"], "returns")
    if c["adv_norm"]:
        np.testing.assert_allclose(got["advantages"], outs["advantages"], rtol=1e-4, atol=1e-4)
    else:
        A.assert_gae_close(got["advantages"], outs["advantages"], "advantages")


@pytest.mark.parametrize("B,S,mask_dtype", [(64, 4099, torch.bool), (3, 1, torch.int64), (5, 2, torch.float32),
                                            (300, 513, torch.int32),
                                            # S % 8 == 0: the 128-bit path (byte masks, float masks, generic)
                                            (33, 4096, torch.bool), (16, 2056, torch.float32), (9, 8, torch.bool),
                                            (20, 1024, torch.int64), (7, 6152, torch.uint8)])
def test_compute_advantages_vs_oracle(B, S, mask_dtype):
    from areal_b200.actor import compute_advantages
    rng = np.random.RandomState(B + S)
    lens = rng.randint(0 if S > 2 else 1, S + 1, B)
    lens[0] = S
    am = np.arange(S)[None, :] < lens[:, None]
    lm = am & (rng.rand(B, S) > 0.3)
    logp = -rng.rand(B, S).astype(np.float32) * 5
    ref = -rng.rand(B, S).astype(np.float32) * 5
    val = rng.randn(B, S).astype(np.float32)
    score = (rng.randn(B) * 4).astype(np.float32)
    c = dict(A.BASE, kl_estimator="k2", kl_ctl=0.3, reward_bias=0.25, reward_scaling=1.5, reward_clip=3.0,
             discount=0.999, gae_lambda=0.97, mask_no_eos_with_zero=True)
    data = dict(input_ids=torch.zeros(B, S, dtype=torch.long, device=DEV),
                attention_mask=torch.from_numpy(am).to(DEV).to(mask_dtype),
                loss_mask=torch.from_numpy(lm).to(DEV).to(mask_dtype), logprobs=torch.from_numpy(logp).to(DEV),
                ref_logp=torch.from_numpy(ref).to(DEV), values=torch.from_numpy(val).to(DEV),
                rewards=torch.from_numpy(score).to(DEV))
    compute_advantages(make_actor(c), data)
    r = O.compute_advantages(score, lm, logp, am, ref_logp=ref, values=val, roll_logp=True, reward_bias=0.25,
                             reward_scaling=1.5, reward_clip=3.0, kl_ctl=0.3, kl_estimator="k2",
                             mask_no_eos_with_zero=True, discount=0.999, gae_lambda=0.97)
    for k in ("loss_mask", "logprobs", "ref_logp", "kl_rewards", "tot_rewards"):
        assert np.array_equal(data[k].cpu().numpy(), r[k]), k
    A.assert_gae_close(data["advantages"].cpu().numpy(), r["advantages"], "advantages")
    A.assert_gae_close(data["returns"].cpu().numpy(), r["returns"], "returns")
    assert np.array_equal(data["rewards"].cpu().numpy(), score)          # no penalty configured: untouched


@pytest.mark.parametrize("B,S,mask_dtype,overlong,with_values,with_ref,kl",
                         # few long rows: a cluster of up to 8 CTAs per row, lengths summed across the cluster
                         [(16, 32768, torch.bool, False, True, True, "k2"), (5, 20001, torch.int64, True, True, True, "k2"),
                          (3, 32769, torch.float32, True, False, True, "k2"),     # column S-1 alone beyond the last tile
                          (9, 4098, torch.bool, False, True, False, "k2"), (4, 8200, torch.uint8, True, True, True, "k1"),
                          # byte masks, k1, no penalty: the kernel's instance for the common configuration
                          (16, 32768, torch.bool, False, False, True, "k1"), (7, 12296, torch.uint8, False, True, True, "k1"),
                          (3, 20001, torch.bool, False, True, False, "k1"),
                          (2, 40000, torch.bool, True, True, True, "k2")])       # beyond a cluster: one CTA per row again
def test_compute_advantages_long_rows_vs_oracle(B, S, mask_dtype, overlong, with_values, with_ref, kl):
    from areal_b200.actor import compute_advantages
    rng = np.random.RandomState(B + S)
    lens = rng.randint(1, S + 1, B)
    lens[0] = S                                                          # no EOS: the bootstrap value enters
    if B > 2:
        lens[1], lens[2] = 1, 4097                                       # reward / zeroed column in the LEFTMOST tiles
    am = np.arange(S)[None, :] < lens[:, None]
    lm = am & (rng.rand(B, S) > 0.3)
    lm[-1, S // 2:] = False                                              # whole tiles without a valid token
    logp = -rng.rand(B, S).astype(np.float32) * 5
    ref = -rng.rand(B, S).astype(np.float32) * 5
    val = rng.randn(B, S).astype(np.float32)
    score = (rng.randn(B) * 4).astype(np.float32)
    c = dict(A.BASE, kl_estimator=kl, kl_ctl=0.3, reward_bias=0.25, reward_scaling=1.5, reward_clip=3.0,
             discount=0.999, gae_lambda=0.97, mask_no_eos_with_zero=True, overlong_reward_penalty=overlong,
             overlong_tokens=1000, overlong_penalty_factor=0.75, max_new_tokens=S // 2)
    data = dict(input_ids=torch.zeros(B, S, dtype=torch.long, device=DEV),
                attention_mask=torch.from_numpy(am).to(DEV).to(mask_dtype),
                loss_mask=torch.from_numpy(lm).to(DEV).to(mask_dtype), logprobs=torch.from_numpy(logp).to(DEV),
                rewards=torch.from_numpy(score).to(DEV))
    if with_ref:
        data["ref_logp"] = torch.from_numpy(ref).to(DEV)
    if with_values:
        data["values"] = torch.from_numpy(val).to(DEV)
    again = {k: v.clone() for k, v in data.items()}
    compute_advantages(make_actor(c), data)
    compute_advantages(make_actor(c), again)                 # run-to-run: the exchanges combine in a fixed order
    for k in ("advantages", "returns", "kl_rewards", "tot_rewards"):
        assert torch.equal(data[k], again[k]), k
    sc = O.overlong_penalty(score, lm, 1000, 0.75, S // 2) if overlong else score
    assert np.array_equal(data["rewards"].cpu().numpy(), sc.astype(np.float32))        # penalty in place, once
    if overlong:
        assert not np.array_equal(sc, score)
    r = O.compute_advantages(sc, lm, logp, am, ref_logp=ref if with_ref else None, values=val if with_values else None,
                             roll_logp=True, reward_bias=0.25, reward_scaling=1.5, reward_clip=3.0, kl_ctl=0.3,
                             kl_estimator=kl, mask_no_eos_with_zero=True, discount=0.999, gae_lambda=0.97)
    keys = ("loss_mask", "logprobs", "kl_rewards", "tot_rewards") + (("ref_logp",) if with_ref else ())
    for k in keys:
        assert np.array_equal(data[k].cpu().numpy(), r[k]), k
    A.assert_gae_close(data["advantages"].cpu().numpy(), r["advantages"], "advantages")
    A.assert_gae_close(data["returns"].cpu().numpy(), r["returns"], "returns")


def test_compute_advantages_errors():
    from areal_b200.actor import compute_advantages
    c = dict(A.BASE, kl_estimator="k4")
    data = dict(input_ids=torch.zeros(2, 4, dtype=torch.long, device=DEV),
                attention_mask=torch.ones(2, 4, dtype=torch.bool, device=DEV),
                loss_mask=torch.ones(2, 4, dtype=torch.bool, device=DEV), logprobs=torch.zeros(2, 4, device=DEV),
                rewards=torch.zeros(2, device=DEV))
    with pytest.raises(ValueError, match="Invalid KL estimator"):
        compute_advantages(make_actor(c), data)
    cpu = {k: v.cpu() for k, v in data.items()}
    with pytest.raises(RuntimeError):
        compute_advantages(make_actor(A.BASE), cpu)
