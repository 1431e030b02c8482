"""GPU parity: areal_b200_masked_stats, the stats-tracker mirror and grpo_loss_fn against the export
of the unmodified reference (tests/golden/grpo_stats.npz) and numpy fp64."""
import ctypes

import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

GRPO_ARGS = [dict(temperature=1.0, eps_clip=0.2, eps_clip_higher=None, c_clip=None, behav_imp_weight_cap=None),
             dict(temperature=0.7, eps_clip=0.1, eps_clip_higher=0.3, c_clip=3.0, behav_imp_weight_cap=2.0)]


def raw_stats(vals, masks, n, accumulate=False, out=None):
    from areal_b200 import _lib
    from areal_b200._lib import call, ptr, stream_ptr
    K = len(vals)
    dev = torch.device(DEV)
    if out is None:
        out = torch.tensor([0.0, 0.0, np.inf, -np.inf], dtype=torch.float64, device=dev).repeat(K, 1)
    vp, mp, op = ((ctypes.c_void_p * K)() for _ in range(3))
    bits = 0
    for k in range(K):
        if vals[k] is not None:
            vp[k] = vals[k].data_ptr()
            if vals[k].dtype != torch.float32:
                bits |= 1 << k
        if masks[k] is not None:
            mp[k] = masks[k].data_ptr()
        op[k] = out.data_ptr() + 32 * k
    nbytes = _lib.lib().areal_b200_masked_stats_workspace_bytes(K, n)
    ws = _lib.workspace("masked_stats_test", dev, nbytes)
    with torch.cuda.device(dev):
        call("areal_b200_masked_stats", vp, mp, op, bits, K, n, int(accumulate), ptr(ws), ws.numel(), stream_ptr(dev))
    return out


@pytest.mark.parametrize("n", [1, 255, 8192, 8193, 1_000_003])
def test_masked_stats_vs_numpy(n):
    rng = np.random.RandomState(n % 1000)
    K = 7
    xs = [rng.randn(n).astype(np.float32) * (10 ** (k - 3)) for k in range(K)]
    ms = [rng.rand(n) > p for p in (0.0, 0.5, 0.99, 1.1, 0.2, 0.7, 0.5)]
    xs[4][rng.randint(0, n, 3)] = np.nan                           # NaNs: masked out or propagated
    tv = [torch.from_numpy(x).to(DEV) for x in xs]
    tm = [torch.from_numpy(m).to(DEV) for m in ms]
    tv[5] = torch.from_numpy(ms[0] ^ ms[1]).to(DEV)                # a bool value array (mask.float())
    xs[5] = (ms[0] ^ ms[1]).astype(np.float32)
    tv[6], tm[6] = None, tm[6]                                     # count only
    xs[6] = np.ones(n, np.float32)
    tm[0] = None                                                   # no mask: everything counts
    ms[0] = np.ones(n, bool)
    out = raw_stats(tv, tm, n).cpu().numpy()
    out2 = raw_stats(tv, tm, n).cpu().numpy()
    assert np.array_equal(out, out2, equal_nan=True)               # fixed order, workspace left clean
    for k in range(K):
        sel = xs[k][ms[k]].astype(np.float64)
        assert out[k, 1] == sel.size
        if np.isnan(sel).any():
            assert np.isnan(out[k, 0]) and np.isnan(out[k, 2]) and np.isnan(out[k, 3])
            continue
        assert abs(out[k, 0] - sel.sum()) <= 1e-12 * max(1.0, np.abs(sel).sum())
        if sel.size:
            assert out[k, 2] == sel.min() and out[k, 3] == sel.max()
        else:
            assert out[k, 2] == np.inf and out[k, 3] == -np.inf
    # running totals: a second batch folded into the first
    acc = raw_stats(tv, tm, n)
    acc = raw_stats(tv, tm, n, accumulate=True, out=acc).cpu().numpy()
    ok = ~np.isnan(out[:, 0])
    assert np.allclose(acc[ok, 0], 2 * out[ok, 0], rtol=1e-15) and np.array_equal(acc[ok, 1], 2 * out[ok, 1])
    assert np.array_equal(acc[ok, 2:], out[ok, 2:])


def test_masked_stats_rejects_bad_arguments():
    from areal_b200 import _lib
    from areal_b200._lib import ptr, stream_ptr
    dev = torch.device(DEV)
    one = torch.zeros(64, device=dev)
    arr = (ctypes.c_void_p * 33)()
    with torch.cuda.device(dev):
        rc = _lib.lib().areal_b200_masked_stats(arr, arr, arr, 0, 33, 8, 0, ptr(one), 256, stream_ptr(dev))
        assert rc == 1 and "at most 32 keys" in _lib.lib().areal_b200_last_error().decode()
        arr1 = (ctypes.c_void_p * 1)()
        out1 = (ctypes.c_void_p * 1)(one.data_ptr())
        rc = _lib.lib().areal_b200_masked_stats(arr1, arr1, out1, 0, 1, 1 << 20, 0, ptr(one), 256, stream_ptr(dev))
        assert rc == 3 and "too small" in _lib.lib().areal_b200_last_error().decode()
        rc = _lib.lib().areal_b200_masked_stats(arr1, arr1, out1, 0, 1, 0, 0, ptr(one), 256, stream_ptr(dev))
        assert rc == 1 and "non-empty" in _lib.lib().areal_b200_last_error().decode()


def load_mb(z, mb):
    c = {}
    for k in ("logits", "input_ids", "logprobs", "prox_logp", "advantages", "loss_mask"):
        v = z[f"mb{mb}_{k}"]
        if k == "logits":
            c[k] = torch.from_numpy(v.view(np.int16)).view(torch.bfloat16).to(DEV)
        else:
            c[k] = torch.from_numpy(v).to(DEV)
    return c


@pytest.mark.parametrize("ai", [0, 1])
def test_grpo_loss_fn_and_export_vs_reference(ai):
    """Three micro-batches through grpo_loss_fn inside a scope, then export(): every key the reference
    exports, same values (fp32 reference sums vs our fp64 totals: 2e-6 relative)."""
    from areal_b200 import stats_tracker as ST
    from areal_b200.actor import grpo_loss_fn
    z = H.load("grpo_stats.npz")
    tr = ST.DistributedStatsTracker()
    losses = []
    with tr.scope("ppo_actor"):
        for mb in range(3):
            c = load_mb(z, mb)
            logits = c.pop("logits").requires_grad_(mb == 0)
            loss = grpo_loss_fn(logits, c, tracker=tr, **GRPO_ARGS[ai])
            if mb == 0:
                loss.backward()
                assert logits.grad is not None and torch.isfinite(logits.grad.float()).all()
            losses.append(float(loss.detach()))
        tr.scalar(lr=1e-3)
        tr.scalar(lr=3e-3)
    out = tr.export()
    np.testing.assert_allclose(losses, z[f"a{ai}_losses"], rtol=1e-4, atol=1e-6)
    keys = [str(k) for k in z[f"a{ai}_keys"]]
    assert sorted(out.keys()) == keys
    for k, ref in zip(keys, z[f"a{ai}_vals"]):
        tol = 2e-3 if ("logp" in k or "entropy" in k or "kl" in k or "weight" in k or "loss" in k) else 2e-6
        assert abs(out[k] - ref) <= tol * max(1.0, abs(ref)), (k, out[k], ref)
    assert tr.export() == {}


def test_grpo_loss_fn_distributed_branch_equals_fused(monkeypatch):
    """With a vocabulary-shard (or Ulysses) group active grpo_loss_fn takes the split path -- log-probs through
    gather_logprobs_entropy's distributed branch, then ppo_actor_loss_fn -- and must export what the fused
    single-process path exports.  A world of one rank exercises that path on one GPU."""
    from areal_b200 import functional as F
    from areal_b200 import stats_tracker as ST
    from areal_b200.actor import grpo_loss_fn
    z = H.load("grpo_stats.npz")
    outs, grads = [], []
    for distributed in (False, True):
        if distributed:
            monkeypatch.setattr(F, "_tp_group", lambda: "one-rank world")
        tr = ST.DistributedStatsTracker()
        with tr.scope("ppo_actor"):
            for mb in range(3):
                c = load_mb(z, mb)
                logits = c.pop("logits").requires_grad_(mb == 0)
                loss = grpo_loss_fn(logits, c, tracker=tr, **GRPO_ARGS[1])
                if mb == 0:
                    loss.backward()
                    grads.append(logits.grad.float())
        outs.append(tr.export())
    assert sorted(outs[0]) == sorted(outs[1])
    for k in outs[0]:
        assert abs(outs[0][k] - outs[1][k]) <= 2e-5 * max(1.0, abs(outs[0][k])), (k, outs[0][k], outs[1][k])
    assert torch.allclose(grads[0], grads[1], rtol=2 ** -7, atol=1e-9)


def test_tracker_api_matches_record():
    """denominator()/stat() one by one == record() in one launch; masks change between micro-batches."""
    from areal_b200 import stats_tracker as ST
    g = torch.Generator(device=DEV).manual_seed(3)
    a, b = ST.DistributedStatsTracker(), ST.DistributedStatsTracker()
    for mb, n in enumerate((1000, 17)):
        m = torch.rand(n, device=DEV, generator=g) > 0.4
        x = torch.randn(n, device=DEV, generator=g)
        y = torch.randn(n, 1, device=DEV, generator=g).expand(n, 1)
        a.denominator(valid=m)
        a.stat(denominator="valid", x=x)
        a.stat(denominator="valid", reduce_type=ST.ReduceType.SUM, y=y.reshape(-1))
        b.record(dict(valid=m), [("valid", dict(x=x))])
        b.stat(denominator="valid", reduce_type=ST.ReduceType.SUM, y=y.reshape(-1))
        with pytest.raises(AssertionError):
            a.stat(denominator="valid", z=torch.zeros(n + 1, device=DEV))
    ea, eb = a.export(), b.export()
    assert ea == eb and set(ea) == {"valid", "x/avg", "x/min", "x/max", "y"}


def test_critic_ppo_loss_fn_and_export_vs_reference():
    """The critic's ppo_loss_fn over three micro-batches (one fully masked) + export: the reference's
    losses, every exported key, and a gradient that flows to the [T,1] value head output."""
    from areal_b200 import stats_tracker as ST
    from areal_b200.critic import ppo_loss_fn
    z = H.load("critic_stats.npz")
    tr = ST.DistributedStatsTracker()
    losses = []
    with tr.scope("ppo_critic"):
        for mb in range(3):
            c = {k: torch.from_numpy(z[f"mb{mb}_{k}"]).to(DEV) for k in ("value", "values", "returns", "loss_mask")}
            value = c.pop("value").requires_grad_()
            loss = ppo_loss_fn(value, c, 0.2, tracker=tr)
            loss.backward()
            assert value.grad.shape == value.shape
            if mb < 2:
                assert torch.isfinite(value.grad).all()
            losses.append(float(loss.detach()))
    # the fully masked micro-batch: the reference divides by count_nonzero() = 0 (functional.py:309), so does this
    assert np.isnan(z["losses"][2]) and np.isnan(losses[2])
    np.testing.assert_allclose(losses[:2], z["losses"][:2], rtol=2e-6, atol=1e-7)
    out = tr.export()
    keys = [str(k) for k in z["keys"]]
    assert sorted(out.keys()) == keys
    for k, ref in zip(keys, z["vals"]):
        assert abs(out[k] - ref) <= 3e-6 * max(1.0, abs(ref)), (k, out[k], ref)


def test_sft_loss_fn_and_export_vs_reference():
    """compute_packed_sft_loss over two packed micro-batches (one sequence without any target token):
    losses, every exported key (ppl per sequence, loss per token, vocab min / max), and a finite gradient."""
    from areal_b200 import stats_tracker as ST
    from areal_b200.sft import compute_packed_sft_loss, segment_masked_mean
    z = H.load("sft_stats.npz")
    tr = ST.DistributedStatsTracker()
    losses = []
    with tr.scope("sft"):
        for mb in range(2):
            c = {k: torch.from_numpy(z[f"mb{mb}_{k}"]).to(DEV) for k in ("input_ids", "cu_seqlens", "loss_mask")}
            logits = torch.from_numpy(z[f"mb{mb}_logits"].view(np.int16)).view(torch.bfloat16).to(DEV).requires_grad_()
            loss = compute_packed_sft_loss(logits, c, tracker=tr)
            loss.backward()
            assert torch.isfinite(logits.grad.float()).all() and float(logits.grad.float().abs().max()) > 0
            losses.append(float(loss.detach()))
    np.testing.assert_allclose(losses, z["losses"], rtol=1e-4)
    out = tr.export()
    keys = [str(k) for k in z["keys"]]
    assert sorted(out.keys()) == keys
    for k, ref in zip(keys, z["vals"]):
        tol = 2e-3 if ("ppl" in k or "loss" in k) else 2e-6
        assert abs(out[k] - ref) <= tol * max(1.0, abs(ref)), (k, out[k], ref)
    # the segment kernel on its own: empty sequences, no mask hits
    v = torch.arange(10, dtype=torch.float32, device=DEV)
    m = torch.tensor([1, 1, 0, 0, 0, 1, 0, 1, 1, 1], dtype=torch.bool, device=DEV)
    cu = torch.tensor([0, 2, 2, 5, 10], dtype=torch.int32, device=DEV)
    mean, valid = segment_masked_mean(v, m, cu)
    assert mean.tolist() == [0.5, 0.0, 0.0, (5 + 7 + 8 + 9) / 4] and valid.tolist() == [True, False, False, True]
