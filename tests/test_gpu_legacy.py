"""GPU parity: the legacy (realhf) spelling of P2 -- gather_packed_shifted_log_probs, the index builders
and calc_entropy -- through the C ABI vs golden vectors of the unmodified reference functions."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu
D = "cuda:0"


def test_legacy_packed_golden():
    from areal_b200 import realhf_functional as RF
    z = H.load("legacy_packed.npz")
    for k, c in H.cases(z):
        bf16 = bool(c["is_bf16"])
        logits = (H.bf16_from_bits(c["logits"]) if bf16 else torch.from_numpy(c["logits"])).to(D)
        cu = torch.from_numpy(c["cu"]).to(D)
        ids = torch.from_numpy(c["ids"]).to(D)
        assert np.array_equal(RF.build_leave_one_indices(logits, cu).cpu().numpy(), c["leave_one"])
        assert np.array_equal(RF.build_shift_one_indices(logits, cu).cpu().numpy(), c["shift_one"])
        x = logits.clone().requires_grad_()
        lp = RF.gather_packed_shifted_log_probs(x, cu, ids)
        assert lp.dtype == logits.dtype and lp.shape[0] == logits.shape[0] - (cu.numel() - 1)
        ent = RF.calc_entropy(logits, cu)
        assert ent.dtype == logits.dtype and not ent.requires_grad
        got_lp, got_ent = lp.detach().float().cpu().numpy(), ent.float().cpu().numpy()
        if bf16:
            assert np.all(np.abs(got_lp - c["logp"]) <= 2 ** -7 * np.abs(c["logp"]) + 1e-6), k
            assert np.all(np.abs(got_ent - c["entropy"]) <= 0.05 * np.abs(c["entropy"]) + 1e-3), k
        else:
            assert H.rel_err(got_lp, c["logp"]) <= H.TOL_LOGITS, k
            assert H.rel_err(got_ent, c["entropy"]) <= H.TOL_LOGITS, k
            assert np.allclose(got_ent, c["entropy"], rtol=1e-4, atol=1e-5), k
        # gradient flows to the logits through the index select
        w = torch.randn(lp.shape, device=D)
        (lp.float() * w).sum().backward()
        lab = np.concatenate([c["ids"][1:], [0]])
        gfull = np.zeros(logits.shape[0])
        gfull[c["leave_one"]] = w.cpu().numpy()
        ref = O.dlogits(logits.float().cpu().numpy(), lab, 1.0, gfull)
        got = x.grad.float().cpu().numpy()
        scale = np.maximum(np.abs(ref).max(-1, keepdims=True), 1e-3 * np.abs(gfull)[:, None]) + 1e-30
        slack = 2 ** -8 if bf16 else 1e-6
        assert np.all(np.abs(got - ref) <= slack * np.abs(ref) + (3e-3 if bf16 else 1e-3) * scale + 1e-30), k


def test_packed_rewards_golden_and_large():
    """get_packed_rewards / get_packed_reward_dense: bit-exact vs golden outputs of the reference
    functions, and vs the pinned oracle on a 200k-token batch with 1-token sequences."""
    from areal_b200 import ppo_functional as P
    z = H.load("packed_rewards.npz")
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)   # noqa: E731
    for k in range(int(z["n_cases"])):
        kl_ctl, clip, mz, delta = [float(v) for v in z[f"c{k}_cfg"]]
        a, b = P.get_packed_rewards(kl_ctl, clip, t(z[f"c{k}_lp"]), t(z[f"c{k}_rf"]), t(z[f"c{k}_score"]),
                                    t(z[f"c{k}_cu"]), t(z[f"c{k}_ne"]), bool(mz))
        assert np.array_equal(a.cpu().numpy(), z[f"c{k}_kl"]) and np.array_equal(b.cpu().numpy(), z[f"c{k}_tot"]), k
        c, d = P.get_packed_reward_dense(kl_ctl, clip, t(z[f"c{k}_lp"]), t(z[f"c{k}_rf"]), t(z[f"c{k}_dense"]),
                                         t(z[f"c{k}_cu"]), t(z[f"c{k}_ne"]).float(), bool(delta))
        assert np.array_equal(c.cpu().numpy(), z[f"c{k}_dkl"]) and np.array_equal(d.cpu().numpy(), z[f"c{k}_dtot"]), k
    rng = np.random.RandomState(4)
    lens = np.concatenate([rng.randint(1, 3000, 120), np.ones(500, np.int64), rng.randint(1, 9, 3000)])
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    B, Ts = len(lens), int(cu[-1])
    lp, rf = -rng.rand(Ts).astype(np.float32) * 5, -rng.rand(Ts).astype(np.float32) * 5
    score, dense = (rng.randn(B) * 6).astype(np.float32), (rng.randn(Ts + B) * 6).astype(np.float32)
    ne = rng.rand(B) > 0.5
    a, b = P.get_packed_rewards(0.05, 3.0, t(lp), t(rf), t(score), t(cu), t(ne), True)
    ra, rb = O.packed_rewards(0.05, 3.0, lp, rf, score, cu, ne, True)
    assert np.array_equal(a.cpu().numpy(), ra) and np.array_equal(b.cpu().numpy(), rb)
    for delta in (False, True):
        c, d = P.get_packed_reward_dense(0.05, 3.0, t(lp), t(rf), t(dense), t(cu), t(ne).float(), delta)
        rc, rd = O.packed_rewards(0.05, 3.0, lp, rf, dense, cu, dense=True, reward_delta=delta)
        assert np.array_equal(c.cpu().numpy(), rc) and np.array_equal(d.cpu().numpy(), rd)


def test_legacy_masked_normalization_golden():
    """realhf masked_normalization (dim=None), with and without mask, biased / unbiased: golden outputs of
    the unmodified reference function."""
    from areal_b200 import realhf_functional as RF
    z = H.load("legacy_masked_norm.npz")
    dev = torch.device("cuda:0")
    for k in range(int(z["n_cases"])):
        has_mask, unbiased = [bool(v) for v in z[f"c{k}_cfg"]]
        x = torch.from_numpy(z[f"c{k}_x"]).to(dev)
        m = torch.from_numpy(z[f"c{k}_mask"]).to(dev) if has_mask else None
        out = RF.masked_normalization(x, m, unbiased=unbiased, all_reduce=False)
        assert out.dtype == torch.float32 and out.shape == x.shape
        np.testing.assert_allclose(out.cpu().numpy(), z[f"c{k}_out"], rtol=1e-5, atol=1e-5)


def test_gather_shifted_log_probs_padded():
    """Padded legacy spelling: [bs, seqlen, V] -> [bs, seqlen-1]; torch's own log_softmax as the check
    (the function is three torch calls upstream, realhf functional.py:84-88)."""
    from areal_b200 import realhf_functional as RF
    g = torch.Generator().manual_seed(2)
    logits = (torch.randn(3, 17, 1000, generator=g) * 2).bfloat16()
    labels = torch.randint(0, 1000, (3, 17), generator=g)
    got = RF.gather_shifted_log_probs(logits.to("cuda:0").requires_grad_(), labels.to("cuda:0"))
    ref = torch.log_softmax(logits[:, :-1].float(), dim=-1).gather(-1, labels[:, 1:].unsqueeze(-1)).squeeze(-1)
    assert got.shape == (3, 16) and got.dtype == torch.bfloat16
    assert torch.allclose(got.float().cpu(), ref, rtol=2 ** -7, atol=2 ** -7)
    got.float().sum().backward()
