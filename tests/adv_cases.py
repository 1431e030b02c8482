"""Shared by the CPU and GPU tests of compute_advantages: how a golden case of
tests/golden/advantages.npz (oracle/gen_golden.py:gen_advantages) maps to arguments."""
import types

import numpy as np

from tests import helpers as H

# the grid of oracle/gen_golden.py:adv_case_grid, restated (the generator needs /root/reference)
BASE = dict(reward_bias=0.0, reward_scaling=1.0, reward_clip=20.0, kl_ctl=0.1, kl_estimator="k1",
            discount=1.0, gae_lambda=1.0, mask_no_eos_with_zero=False, use_decoupled_loss=True,
            recompute_logprob=True, overlong_reward_penalty=False, overlong_tokens=8, overlong_penalty_factor=1.0,
            max_new_tokens=24, with_values=False, with_ref=True, adv_norm=False, reward_norm=False)
GRID = [dict(),
        dict(kl_estimator="k2", kl_ctl=0.5, discount=0.99, gae_lambda=0.95, with_values=True),
        dict(kl_estimator="k3", kl_ctl=0.05, reward_bias=-0.5, reward_scaling=3.0, reward_clip=2.5,
             mask_no_eos_with_zero=True, with_values=True, discount=0.9, gae_lambda=0.8),
        dict(with_ref=False, kl_ctl=0.0, adv_norm=True),
        dict(use_decoupled_loss=False, recompute_logprob=True, with_values=True, discount=0.97, gae_lambda=0.9),
        dict(use_decoupled_loss=False, recompute_logprob=False, kl_estimator="k2"),
        dict(overlong_reward_penalty=True, overlong_tokens=6, overlong_penalty_factor=0.75, max_new_tokens=20,
             reward_scaling=2.0, with_values=True, discount=0.95, gae_lambda=0.9),
        dict(reward_norm=True, adv_norm=True, kl_estimator="k3", with_values=True, gae_lambda=0.7)]
CASES = [dict(BASE, **g) for g in GRID]
IN_KEYS = ("input_ids", "attention_mask", "loss_mask", "logprobs", "prox_logp", "ref_logp", "values", "rewards")
OUT_KEYS = ("advantages", "returns", "kl_rewards", "tot_rewards", "loss_mask", "logprobs", "rewards", "ref_logp",
            "prox_logp")


def load():
    z = H.load("advantages.npz")
    assert int(z["n_cases"]) == len(CASES)
    out = []
    for ci, c in enumerate(CASES):
        ins = {k: z[f"c{ci}_in_{k}"] for k in IN_KEYS if f"c{ci}_in_{k}" in z}
        outs = {k: z[f"c{ci}_out_{k}"] for k in OUT_KEYS if f"c{ci}_out_{k}" in z}
        out.append((c, ins, outs))
    return out


def config_of(c):
    return types.SimpleNamespace(**{k: c[k] for k in ("use_decoupled_loss", "recompute_logprob", "overlong_reward_penalty",
                                                      "overlong_tokens", "overlong_penalty_factor", "max_new_tokens")})


def assert_gae_close(got, ref, what):
    """Advantages / returns: the reference loop drifts in fp32, ours is the correctly rounded fp64
    scan, so compare with the GAE tolerance relative to the row scale."""
    scale = np.maximum(1.0, np.abs(ref).max(axis=-1, keepdims=True))
    err = float((np.abs(got.astype(np.float64) - ref) / scale).max())
    assert err <= H.TOL_GAE, (what, err)
