"""GPU parity: Normalization / masked_normalization through the C ABI vs golden vectors of the
unmodified reference class over the grid its own tests walk (areal/tests/test_adv_norm_config.py)."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu
D = "cuda:0"
LEVEL = {0: "batch", 1: "group", 2: None}


def test_normalization_golden_grid():
    from areal_b200.data import NormConfig, Normalization
    z = H.load("normalization.npz")
    n = 0
    for k, c in H.cases(z):
        ml, sl, loo, unb, gs = (int(v) for v in c["cfg"])
        cfg = NormConfig(mean_level=LEVEL[ml], mean_leave1out=bool(loo), std_level=LEVEL[sl], std_unbiased=bool(unb),
                         eps=1e-5, group_size=gs)
        x = torch.from_numpy(c["x"]).to(D)
        m = torch.from_numpy(c["mask"]).to(D) if bool(c["has_mask"]) else None
        y = Normalization(cfg)(x, m)
        assert y.dtype == torch.float32 and y.shape == x.shape
        got = y.cpu().numpy()
        assert np.allclose(got, c["y"], rtol=1e-5, atol=1e-5), (k, c["cfg"], float(np.abs(got - c["y"]).max()))
        n += 1
    assert n == 304


def test_masked_normalization_golden_and_large():
    from areal_b200.data import masked_normalization
    z = H.load("normalization.npz")
    for k, c in H.cases(z, prefix="m", count_key="n_masked_norm"):
        x = torch.from_numpy(c["x"]).to(D)
        m = torch.from_numpy(c["mask"]).to(D) if bool(c["has_mask"]) else None
        y = masked_normalization(x, m, unbiased=bool(c["unbiased"]))
        assert np.allclose(y.cpu().numpy(), c["y"], rtol=1e-5, atol=1e-5), k
    # a realistic size: [512, 4096] advantages, several slabs per group
    rng = np.random.RandomState(0)
    x = (rng.randn(512, 4096) * 2 + 0.3).astype(np.float32)
    m = (rng.rand(512, 4096) > 0.4).astype(np.float32)
    want = O.masked_normalization(x, m)
    got = masked_normalization(torch.from_numpy(x).to(D), torch.from_numpy(m).to(D)).cpu().numpy()
    assert np.allclose(got, want, rtol=1e-5, atol=1e-5)
    from areal_b200.data import NormConfig, Normalization
    for cfg in (NormConfig("batch", False, "batch", True, 1e-5, 1), NormConfig("group", True, "group", False, 1e-5, 8),
                NormConfig("group", False, "batch", True, 1e-5, 16)):
        want = O.normalization(x, m, cfg.mean_level, cfg.std_level, cfg.mean_leave1out, cfg.std_unbiased, cfg.eps,
                               cfg.group_size)
        got = Normalization(cfg)(torch.from_numpy(x).to(D), torch.from_numpy(m).to(D)).cpu().numpy()
        assert np.allclose(got, want, rtol=2e-5, atol=2e-5), cfg


def test_normalization_errors():
    from areal_b200.data import NormConfig, Normalization
    with pytest.raises(ValueError):
        Normalization(NormConfig(mean_level="token"))
    with pytest.raises(ValueError):
        Normalization(NormConfig(std_level="seq"))
    with pytest.raises(ValueError):
        Normalization(NormConfig(mean_level="group", group_size=None))


def test_pack_tensor_dict_golden_and_round_trip():
    """pack_tensor_dict through the interval gather kernel == the reference's Python loop
    (golden from the unmodified function), bit for bit, for int64 / bool / fp32 / bf16 [B,S,3]."""
    from areal_b200.data import pack_tensor_dict, unpack_to_padded
    z = H.load("pack.npz")
    B, S = z["attention_mask"].shape
    data = dict(attention_mask=torch.from_numpy(z["attention_mask"]).to(D),
                input_ids=torch.from_numpy(z["in_input_ids"]).to(D),
                loss_mask=torch.from_numpy(z["in_loss_mask"]).to(D),
                logprobs=torch.from_numpy(z["in_logprobs"]).to(D),
                hidden=H.bf16_from_bits(z["in_hidden"].reshape(-1)).reshape(B, S, 3).to(D),
                rewards=torch.from_numpy(z["in_rewards"]).to(D), tag="keep")
    out = pack_tensor_dict(data)
    assert "attention_mask" not in out and out["tag"] == "keep" and out["max_seqlen"] == int(z["max_seqlen"])
    assert np.array_equal(out["cu_seqlens"].cpu().numpy(), z["cu_seqlens"]) and out["cu_seqlens"].dtype == torch.int32
    for k in ("input_ids", "loss_mask", "logprobs", "rewards"):
        assert out[k].dtype == data[k].dtype
        assert np.array_equal(out[k].cpu().numpy(), z["out_" + k]), k
    assert np.array_equal(H.torch_bits(out["hidden"].reshape(-1)).reshape(-1, 3), z["out_hidden"])
    back = unpack_to_padded(out["logprobs"], out["cu_seqlens"], S)
    am = data["attention_mask"].bool()
    assert torch.equal(back[am], data["logprobs"][am]) and bool((back[~am] == 0).all())
    hb = unpack_to_padded(out["hidden"], out["cu_seqlens"], S)
    assert torch.equal(hb[am], data["hidden"][am])


def test_legacy_running_mean_std_vs_reference():
    """ExponentialRunningMeanStd / MovingAverageRunningMeanStd, float64 and float32 state: after each of
    three updates (no mask, 1-D mask, 2-D mask) mean/std and normalize/denormalize of a probe vector match
    golden outputs of the unmodified reference modules; state_dict keys are the reference's mangled names."""
    from areal_b200 import rms
    z = H.load("rms.npz")
    dev = torch.device("cuda:0")
    probe = torch.from_numpy(z["probe"]).to(dev)
    for k in range(int(z["n_cases"])):
        exp, hp = [bool(v) for v in z[f"c{k}_cfg"]]
        mod = rms.ExponentialRunningMeanStd(beta=0.9, epsilon=1e-5, high_precision=hp) if exp else \
            rms.MovingAverageRunningMeanStd(high_precision=hp)
        tol = 1e-12 if hp else 2e-6
        for step in range(3):
            x = torch.from_numpy(z[f"x{step}"]).to(dev)
            m = torch.from_numpy(z[f"m{step}"]).to(dev) if f"m{step}" in z else None
            mod.update(x, m)
            mean, std = mod.mean_std()
            assert mean.dtype == (torch.float64 if hp else torch.float32) and mean.shape == (1,)
            np.testing.assert_allclose(mean.double().cpu().numpy(), z[f"c{k}_mean{step}"], rtol=tol, atol=tol)
            np.testing.assert_allclose(std.double().cpu().numpy(), z[f"c{k}_std{step}"], rtol=tol, atol=tol)
            n, d = mod.normalize(probe), mod.denormalize(probe)
            assert n.dtype == torch.float32 and d.dtype == torch.float32
            np.testing.assert_allclose(n.cpu().numpy(), z[f"c{k}_norm{step}"], rtol=2e-6, atol=2e-6)
            np.testing.assert_allclose(d.cpu().numpy(), z[f"c{k}_denorm{step}"], rtol=2e-6, atol=2e-6)
            if hp:                                   # same fp64 arithmetic: identical after the f32 rounding
                assert (n.cpu().numpy() != z[f"c{k}_norm{step}"]).mean() < 1e-3
        keys = set(mod.state_dict().keys())
        cls = "ExponentialRunningMeanStd" if exp else "MovingAverageRunningMeanStd"
        want = {f"_{cls}__mean", f"_{cls}__mean_sq"} | ({f"_{cls}__debiasing_term"} if exp else set())
        assert keys == want, keys
        # a checkpoint round trip keeps the kernels' buffer and the parameters tied together
        clone = (rms.ExponentialRunningMeanStd(beta=0.9, epsilon=1e-5, high_precision=hp) if exp
                 else rms.MovingAverageRunningMeanStd(high_precision=hp))
        clone.load_state_dict(mod.state_dict())
        if exp:
            assert torch.equal(clone.normalize(probe), mod.normalize(probe))
    with pytest.raises(NotImplementedError):
        mod.forward(probe)


def test_dynamic_sampling_golden():
    """dynamic_sampling: bit-exact vs the reference function (mixed / nothing filtered / everything equal),
    pass-through of non-batch entries, the early exits."""
    import warnings
    from areal_b200.data import dynamic_sampling
    z = H.load("dynamic_sampling.npz")
    dev = torch.device("cuda:0")
    for ci in range(3):
        data = {k: torch.from_numpy(z[f"c{ci}_in_{k}"]).to(dev) for k in ("rewards", "input_ids", "loss_mask", "logprobs", "other")}
        data["hidden"] = H.bf16_from_bits(z[f"c{ci}_in_hidden"]).to(dev)
        data["meta"] = "keep"
        out, stat = dynamic_sampling(data, 4)
        assert [stat["n_group_kept"], stat["n_group_filtered"]] == z[f"c{ci}_stat"].tolist()
        assert out["meta"] == "keep"
        for k in ("rewards", "input_ids", "loss_mask", "logprobs", "other"):
            got = out[k].cpu().numpy()
            assert got.dtype == z[f"c{ci}_out_{k}"].dtype and np.array_equal(got, z[f"c{ci}_out_{k}"]), (ci, k)
        assert np.array_equal(H.torch_bits(out["hidden"].reshape(-1)).reshape(-1, 11, 3), z[f"c{ci}_out_hidden"])
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        same, stat = dynamic_sampling(data, 5)                       # 24 % 5 != 0: untouched
        assert same is data and stat == dict(n_group_kept=4, n_group_filtered=0) and w
        same, stat = dynamic_sampling(data, 0)
        assert same is data and stat == dict(n_group_kept=0, n_group_filtered=0)
    with pytest.raises(TypeError):
        dynamic_sampling(dict(rewards=[1, 2]), 2)
