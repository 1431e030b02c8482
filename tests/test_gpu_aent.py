"""GPU parity: AEnt clamped entropy (areal_b200_clamped_entropy_fwd / _bwd) against golden outputs of the
unmodified reference functions and against the pinned oracle."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as O
from tests import clamp_cases as CC
from tests import helpers as H

pytestmark = pytest.mark.gpu
D = torch.device("cuda:0")


@pytest.mark.parametrize("ci", range(6))
def test_clamped_entropy_golden(ci):
    from areal_b200 import aent_functional as A
    c = CC.load()[ci]
    logits = H.bf16_from_bits(c["bits"]).to(D).requires_grad_()
    labels = torch.from_numpy(c["labels"]).to(D)
    lp, ent = A.gather_logprobs_clamped_entropy(logits, labels, c["clamp"], c["temp"])
    assert lp.dtype == torch.float32 and ent.shape == labels.shape
    assert H.rel_err(lp.detach().cpu().numpy(), c["logp"]) <= H.TOL_LOGITS
    assert H.rel_err(ent.detach().cpu().numpy(), c["ent"]) <= H.TOL_LOGITS
    w_lp, w_h = torch.from_numpy(c["w_lp"]).to(D), torch.from_numpy(c["w_h"]).to(D)
    ((lp * w_lp).sum() + (ent * w_h).sum()).backward()
    # bf16 gradient: one ulp (2^-8 relative) on top of the 1e-3 bar, relative to the row's largest entry
    CC.check_grad(logits.grad.float().cpu().numpy(), c["grad"].astype(np.float64),
                  H.bf16_from_bits(c["bits"]).float().numpy(), c["clamp"], 6e-3, "cuda")


@pytest.mark.parametrize("dtype,T,V,clamp,extra", [(torch.bfloat16, 40, 151936, 0.3, 0), (torch.float16, 33, 9000, 0.6, 0),
                                                   (torch.bfloat16, 17, 1003, 0.2, 0), (torch.float16, 9, 4104, 0.05, 8),
                                                   (torch.bfloat16, 5, 8, 0.5, 0), (torch.bfloat16, 3, 50000, 0.999, 3)])
def test_clamped_entropy_vs_oracle(dtype, T, V, clamp, extra):
    """Aligned and unaligned rows (odd V, column slices of a wider matrix), fp16 keys, extreme clamps."""
    from areal_b200 import aent_functional as A
    g = torch.Generator().manual_seed(V + T)
    wide = (torch.randn(T, V + extra, generator=g) * 3).to(dtype)
    labels = torch.randint(0, V, (T,), generator=g)
    x = wide.to(D)[:, :V].requires_grad_()
    lp, ent = A.gather_logprobs_clamped_entropy(x, labels.to(D), clamp, 0.8)
    xf = wide[:, :V].float().numpy()
    rlp, rent = O.gather_logprobs_clamped_entropy(xf, labels.numpy(), clamp, 0.8)
    assert H.rel_err(lp.detach().cpu().numpy(), rlp) <= H.TOL_LOGITS
    assert H.rel_err(ent.detach().cpu().numpy(), rent) <= H.TOL_LOGITS
    # the stand-alone entropy entry gives the same numbers (logits already divided by the temperature)
    ent2 = A.clamped_softmax_entropy(x.detach(), clamp)
    rent2 = O.gather_logprobs_clamped_entropy(xf, labels.numpy(), clamp, 1.0)[1]
    assert H.rel_err(ent2.cpu().numpy(), rent2) <= H.TOL_LOGITS
    w_lp, w_h = torch.randn(T, generator=g), torch.randn(T, generator=g)
    w_h[0] = 0.0
    ((lp * w_lp.to(D)).sum() + (ent * w_h.to(D)).sum()).backward()
    ref = O.clamped_entropy_dlogits(xf, labels.numpy(), clamp, 0.8, w_lp.numpy().astype(np.float64),
                                    w_h.numpy().astype(np.float64))
    ulp = 6e-3 if dtype == torch.bfloat16 else 2e-3
    CC.check_grad(x.grad.float().cpu().numpy(), ref, xf, clamp, ulp, "cuda-vs-oracle")
    # ours spreads the tie weight evenly, which is exactly what the oracle does: compare the ties too
    got = x.grad.float().cpu().numpy().astype(np.float64)
    scale = np.abs(ref).max(-1, keepdims=True) + 1e-30
    assert np.all(np.abs(got - ref) <= ulp * scale + 2 ** -8 * np.abs(ref))


def test_clamped_entropy_errors_and_zero_clamp():
    from areal_b200 import aent_functional as A
    from areal_b200 import functional as F
    x = (torch.randn(6, 512) * 2).bfloat16().to(D)
    lab = torch.randint(0, 512, (6,)).to(D)
    lp0, ent0 = A.gather_logprobs_clamped_entropy(x, lab, 0.0, 1.0)       # nothing removed: the plain entropy
    lp1, ent1 = F.gather_logprobs_entropy(x, lab, 1.0)
    assert torch.allclose(lp0, lp1, rtol=1e-5, atol=1e-5) and torch.allclose(ent0, ent1, rtol=1e-5, atol=1e-5)
    with pytest.raises(ValueError, match="need 0 <= k < V"):
        A.gather_logprobs_clamped_entropy(x, lab, 1.0, 1.0)
    with pytest.raises(TypeError, match="bf16 or fp16"):
        A.gather_logprobs_clamped_entropy(x.float(), lab, 0.1, 1.0)
    with pytest.raises(Exception):
        A.gather_logprobs_clamped_entropy(x.cpu(), lab.cpu(), 0.1, 1.0)


@pytest.mark.parametrize("ai", [0, 1])
def test_aent_grpo_loss_fn_and_export_vs_reference(ai):
    """aent_grpo_loss_fn over three micro-batches (one fully masked) + tracker export: the reference's
    losses and every exported key (inputs shared with tests/golden/grpo_stats.npz)."""
    from areal_b200 import aent_functional as A
    from areal_b200 import stats_tracker as ST
    from tests.test_gpu_stats import GRPO_ARGS, load_mb
    zin, z = H.load("grpo_stats.npz"), H.load("aent_stats.npz")
    coeff, clamp = [float(v) for v in z[f"a{ai}_cfg"]]
    tr = ST.DistributedStatsTracker()
    losses = []
    with tr.scope("aent"):
        for mb in range(3):
            c = load_mb(zin, mb)
            logits = c.pop("logits").requires_grad_(mb == 0)
            loss = A.aent_grpo_loss_fn(logits, c, entropy_coeff=coeff, entropy_clamp=clamp, tracker=tr, **GRPO_ARGS[1])
            if mb == 0:
                loss.backward()
                assert torch.isfinite(logits.grad.float()).all() and float(logits.grad.float().abs().max()) > 0
            losses.append(float(loss.detach()))
    np.testing.assert_allclose(losses, z[f"a{ai}_losses"], rtol=1e-4, atol=1e-6)
    out = tr.export()
    keys = [str(k) for k in z[f"a{ai}_keys"]]
    assert sorted(out.keys()) == keys
    for k, ref in zip(keys, z[f"a{ai}_vals"]):
        tol = 2e-3 if any(s in k for s in ("logp", "entropy", "kl", "weight", "loss")) else 2e-6
        assert abs(out[k] - ref) <= tol * max(1.0, abs(ref)), (k, out[k], ref)


def test_one_pass_forward_brackets_ties_and_fallbacks():
    """The one-pass forward brackets the k-th smallest logit from a sample of the row and falls back to the three-pass
    kernel where that bracket fails.  Rows built to hit every branch: a sample that lies (the sampled vectors hold
    +40: the bracket misses the threshold), one value repeated over the whole row and over half of it (the threshold
    sits inside a huge tie class), random bit patterns (the bracket spans far more than 1,024 keys), and plain rows."""
    from areal_b200 import aent_functional as A
    g = torch.Generator().manual_seed(11)
    T, V = 7, 151936
    x = (torch.randn(T, V, generator=g) * 3).bfloat16()
    nv = V // 8
    for t in range(256):                                   # the vectors the kernel samples
        v = (t * nv) // 256
        x[0, v * 8:(v + 1) * 8] = 40.0
    x[1] = 1.5
    x[2, : V // 2] = -5.0
    bits = torch.randint(0, 0x7f7f, (V,), generator=g, dtype=torch.int32)           # finite bf16 magnitudes ...
    sign = torch.randint(0, 2, (V,), generator=g, dtype=torch.int32) << 15
    x[3] = ((bits | sign).to(torch.int16)).view(torch.bfloat16).float().clamp(-60, 60).bfloat16()   # ... of every exponent
    x[4, ::3] = 0.0
    x[4, 1::3] = -0.0
    labels = torch.randint(0, V, (T,), generator=g)
    for clamp in (0.3, 0.05, 0.9):
        lp, ent = A.gather_logprobs_clamped_entropy(x.to(D), labels.to(D), clamp, 1.0)
        rlp, rent = O.gather_logprobs_clamped_entropy(x.float().numpy(), labels.numpy(), clamp, 1.0)
        assert H.rel_err(lp.cpu().numpy(), rlp) <= H.TOL_LOGITS, clamp
        assert H.rel_err(ent.cpu().numpy(), rent) <= H.TOL_LOGITS, (clamp, ent.cpu().numpy(), rent)
