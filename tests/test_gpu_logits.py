"""GPU parity: fused log-prob / entropy forward and dlogits backward through the C ABI, vs golden
vectors from the unmodified reference and vs the fp64 oracle.  Tolerance (north_star): 1e-3
relative, bf16 in / fp32 accumulate."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu
D = "cuda:0"


def grad_close(got_bf16: torch.Tensor, ref: np.ndarray, g_scale=None, tol=H.TOL_LOGITS):
    """dlogits is rounded to bf16 (2^-9 relative per element) on both sides: compare against the
    fp64 value with one bf16 ulp of slack plus the 1e-3 bar relative to the row's largest entry.
    `g_scale` = |upstream| / temperature per row: entries are g*(1[v=y] - p_v), so when p_label -> 1
    every entry is a cancellation result ~1e-9*g while fp32 arithmetic (the reference's too) carries
    ~6e-8*g of noise; the row scale is therefore floored at 1e-3 * g_scale."""
    got = got_bf16.float().cpu().numpy().astype(np.float64)
    scale = np.abs(ref).max(axis=-1, keepdims=True)
    if g_scale is not None:
        scale = np.maximum(scale, 1e-3 * np.abs(np.asarray(g_scale, dtype=np.float64))[:, None])
    scale = scale + 1e-30
    err = np.abs(got - ref)
    return bool(np.all(err <= 2 ** -8 * np.abs(ref) + tol * scale + 1e-30)), float((err / scale).max())


def test_golden_forward_and_backward():
    from areal_b200 import functional as F
    z = H.load("logprobs.npz")
    for k, c in H.cases(z):
        logits = H.bf16_from_bits(c["logits_bits"]).to(D).requires_grad_()
        labels = torch.from_numpy(c["labels"]).to(D)
        temp = float(c["temp"])
        lp, ent = F.gather_logprobs_entropy(logits, labels, temp)
        assert lp.dtype == torch.float32 and ent.dtype == torch.float32 and lp.shape == labels.shape
        assert H.rel_err(lp.detach().cpu().numpy(), c["logp"]) <= H.TOL_LOGITS, k
        assert H.rel_err(ent.detach().cpu().numpy(), c["entropy"]) <= H.TOL_LOGITS, k
        lp_only = F.gather_logprobs(logits, labels, temp)
        assert torch.equal(lp_only.detach(), lp.detach())
        g_lp = torch.from_numpy(c["g_lp"]).to(D)
        g_H = torch.from_numpy(c["g_H"]).to(D)
        (grad,) = torch.autograd.grad((lp * g_lp).sum(), logits, retain_graph=True)
        assert grad.dtype == torch.bfloat16 and grad.shape == logits.shape
        x = H.bf16_from_bits(c["logits_bits"]).float().numpy()
        ok, worst = grad_close(grad, O.dlogits(x, c["labels"], temp, c["g_lp"]), np.abs(c["g_lp"]) / temp)
        assert ok, (k, worst)
        # and directly against the reference's own (bf16-rounded) gradient
        ref_bf16 = H.bf16_from_bits(c["grad_lp_bits"]).float().numpy().astype(np.float64)
        ok, worst = grad_close(grad, ref_bf16, np.abs(c["g_lp"]) / temp, tol=2 * H.TOL_LOGITS)
        assert ok, (k, worst)
        (grad2,) = torch.autograd.grad((lp * g_lp).sum() + (ent * g_H).sum(), logits)
        ok, worst = grad_close(grad2, O.dlogits(x, c["labels"], temp, c["g_lp"], c["g_H"]),
                               (np.abs(c["g_lp"]) + np.abs(c["g_H"]) * 12) / temp)
        assert ok, (k, worst)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("V", [151936, 18992, 4099, 7])
def test_forward_vs_fp64_oracle(dtype, V):
    from areal_b200 import functional as F
    g = torch.Generator().manual_seed(V)
    T = 24 if V > 10000 else 67
    logits = (torch.randn(T, V, generator=g) * 2.0).to(dtype)
    labels = torch.randint(0, V, (T,), generator=g)
    logits[0, labels[0]] += 30.0
    logits[1] *= 4.0
    for temp in (1.0, 0.7):
        st = F.logits_stats(logits.to(D), labels.to(D), temp)
        lp, ent, lse, mn, mx = O.gather_logprobs_entropy(logits.float().numpy(), labels.numpy(), temp)
        assert H.rel_err(st["logp"].cpu().numpy(), lp) <= H.TOL_LOGITS
        assert H.rel_err(st["entropy"].cpu().numpy(), ent) <= H.TOL_LOGITS
        assert H.rel_err(st["lse"].cpu().numpy(), lse) <= H.TOL_LOGITS
        assert np.array_equal(st["vocab_min_logits"].cpu().numpy(), mn)
        assert np.array_equal(st["vocab_max_logits"].cpu().numpy(), mx)


def test_strided_rows_and_views():
    """What the engine passes: `logits.squeeze(0)[:-pad]`, a row-prefix view of a padded tensor
    (areal/engine/fsdp_engine.py:573-578); also a column slice (row stride > V) and 3-D input."""
    from areal_b200 import functional as F
    g = torch.Generator().manual_seed(3)
    full = (torch.randn(1, 40, 4104, generator=g) * 2).bfloat16().to(D)
    labels = torch.randint(0, 4096, (33,), generator=g).to(D)
    view = full.squeeze(0)[:-7]
    cols = view[:, :4096]                       # stride(0) = 4104 > V = 4096, 16-byte aligned rows
    odd = view[:, 3:4099]                       # misaligned rows -> generic kernel
    for x in (view[:, :4104], cols, odd):
        V = x.shape[1]
        lab = labels % V
        x = x.detach().requires_grad_()
        lp, ent = F.gather_logprobs_entropy(x, lab, 0.9)
        ref_lp, ref_ent, *_ = O.gather_logprobs_entropy(x.detach().float().cpu().numpy(), lab.cpu().numpy(), 0.9)
        assert H.rel_err(lp.detach().cpu().numpy(), ref_lp) <= H.TOL_LOGITS
        assert H.rel_err(ent.detach().cpu().numpy(), ref_ent) <= H.TOL_LOGITS
        w = torch.randn(lp.shape, device=D)
        (grad,) = torch.autograd.grad((lp * w).sum(), x)
        ok, worst = grad_close(grad, O.dlogits(x.detach().float().cpu().numpy(), lab.cpu().numpy(), 0.9,
                                               w.cpu().numpy()), np.abs(w.cpu().numpy()) / 0.9)
        assert ok, worst
    x3 = full[:, :32, :4096].reshape(4, 8, 4096)
    lab3 = labels[:32].reshape(4, 8)
    lp3 = F.gather_logprobs(x3, lab3)
    assert lp3.shape == (4, 8)
    ref, *_ = O.gather_logprobs_entropy(x3.reshape(32, 4096).float().cpu().numpy(), lab3.reshape(-1).cpu().numpy(), 1.0)
    assert H.rel_err(lp3.reshape(-1).cpu().numpy(), ref) <= H.TOL_LOGITS


def test_zero_gradient_rows_are_zero_and_label_fixup():
    from areal_b200 import functional as F
    g = torch.Generator().manual_seed(11)
    T, V = 300, 16384 + 8
    logits = (torch.randn(T, V, generator=g) * 3).bfloat16().to(D).requires_grad_()
    labels = torch.randint(0, V, (T,), generator=g).to(D)
    labels[:4] = torch.tensor([0, V - 1, 8191, 8192])
    w = torch.randn(T, generator=g).to(D)
    w[5::3] = 0.0
    lp = F.gather_logprobs(logits, labels, 1.0)
    (grad,) = torch.autograd.grad((lp * w).sum(), logits)
    assert torch.count_nonzero(grad[5::3]) == 0
    ok, worst = grad_close(grad, O.dlogits(logits.detach().float().cpu().numpy(), labels.cpu().numpy(), 1.0,
                                           w.cpu().numpy()), np.abs(w.cpu().numpy()))
    assert ok, worst
    # d logp / d logits sums to zero over the vocabulary
    rs = grad.float().sum(-1).abs().cpu().numpy()
    assert (rs <= 2e-2 * np.abs(w.cpu().numpy()) + 1e-6).all()


def test_vocab_sharded_partials_combine():
    """cfg 4 in miniature on one GPU: 8 shards of V/8 = 18,992, partial kernel per shard + combine
    == the unsharded kernel and the fp64 oracle; the backward uses the global lse per shard."""
    import ctypes
    from areal_b200 import _lib, functional as F
    from areal_b200._lib import call, ptr, stream_ptr
    g = torch.Generator().manual_seed(5)
    T, V, R = 19, 151936, 8
    Vs = V // R
    logits = (torch.randn(T, V, generator=g) * 2).bfloat16().to(D)
    labels = torch.randint(0, V, (T,), generator=g).to(D)
    temp = 0.8
    parts = torch.empty(R, 4, T, dtype=torch.float32, device=D)
    dev = torch.device(D)
    for r in range(R):
        shard = logits[:, r * Vs:(r + 1) * Vs]
        call("areal_b200_logits_partial_fwd", ptr(shard), 0, T, Vs, shard.stride(0), ptr(labels), r * Vs, temp,
             ptr(parts[r]), None, stream_ptr(dev))
    lp = torch.empty(T, device=D)
    ent = torch.empty(T, device=D)
    row_max = torch.empty(T, device=D)
    ln_s = torch.empty(T, device=D)
    call("areal_b200_vocab_combine", ptr(parts), R, T, temp, ptr(lp), ptr(ent), ptr(row_max), ptr(ln_s),
         stream_ptr(dev))
    lse = row_max / temp + ln_s
    ref_lp, ref_ent, ref_lse, *_ = O.gather_logprobs_entropy(logits.float().cpu().numpy(), labels.cpu().numpy(), temp)
    assert H.rel_err(lp.cpu().numpy(), ref_lp) <= H.TOL_LOGITS
    assert H.rel_err(ent.cpu().numpy(), ref_ent) <= H.TOL_LOGITS
    assert H.rel_err(lse.cpu().numpy(), ref_lse) <= H.TOL_LOGITS
    st = F.logits_stats(logits, labels, temp)
    assert torch.allclose(st["logp"], lp, rtol=1e-5, atol=1e-5) and torch.allclose(st["entropy"], ent, rtol=1e-5, atol=1e-5)
    # sharded backward
    w = torch.randn(T, generator=g).to(D)
    full = O.dlogits(logits.float().cpu().numpy(), labels.cpu().numpy(), temp, w.cpu().numpy())
    for r in (0, 3, 7):
        shard = logits[:, r * Vs:(r + 1) * Vs]
        d = torch.empty(T, Vs, dtype=torch.bfloat16, device=D)
        call("areal_b200_logits_bwd", ptr(shard), 0, T, Vs, shard.stride(0), ptr(labels), r * Vs, temp,
             ptr(row_max), ptr(ln_s), None, ptr(w), None, None, None, ptr(d), Vs, stream_ptr(dev))
        ok, worst = grad_close(d, full[:, r * Vs:(r + 1) * Vs], np.abs(w.cpu().numpy()) / temp)
        assert ok, (r, worst)


def test_full_size_config2_spot_check():
    """BASELINE.json configs[1] at full size (T=16,384, V=151,936 bf16, 4.98 GB): sampled rows vs the
    fp64 oracle, plus size-independent properties (logp <= 0, 0 <= H <= ln V, dlogits rows sum to 0,
    masked rows exactly zero)."""
    from areal_b200 import functional as F
    T, V = 16384, 151936
    g = torch.Generator(device=D).manual_seed(0)
    logits = torch.empty(T, V, dtype=torch.bfloat16, device=D)
    for i in range(0, T, 2048):
        logits[i:i + 2048] = (torch.randn(2048, V, generator=g, device=D) * 2.0).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g, device=D)
    logits.requires_grad_()
    lp, ent = F.gather_logprobs_entropy(logits, labels, 1.0)
    assert bool((lp <= 0).all()) and bool((ent >= 0).all()) and bool((ent <= np.log(V) + 1e-3).all())
    rows = torch.randint(0, T, (24,), generator=torch.Generator().manual_seed(1))
    rows[0], rows[1] = 0, T - 1
    sub = logits.detach()[rows.to(D)].float().cpu().numpy()
    ref_lp, ref_ent, *_ = O.gather_logprobs_entropy(sub, labels[rows.to(D)].cpu().numpy(), 1.0)
    assert H.rel_err(lp.detach()[rows.to(D)].cpu().numpy(), ref_lp) <= H.TOL_LOGITS
    assert H.rel_err(ent.detach()[rows.to(D)].cpu().numpy(), ref_ent) <= H.TOL_LOGITS
    w = torch.randn(T, generator=g, device=D)
    w[torch.rand(T, generator=g, device=D) < 0.3] = 0.0
    (grad,) = torch.autograd.grad((lp * w).sum(), logits)
    ref_g = O.dlogits(sub, labels[rows.to(D)].cpu().numpy(), 1.0, w[rows.to(D)].cpu().numpy())
    ok, worst = grad_close(grad[rows.to(D)], ref_g, np.abs(w[rows.to(D)].cpu().numpy()))
    assert ok, worst
    zero_rows = (w == 0)
    assert int(torch.count_nonzero(grad[zero_rows][:64])) == 0
    rs = grad[:512].float().sum(-1).abs()
    assert bool((rs <= 3e-2 * w[:512].abs() + 1e-6).all())


def test_peaked_rows_keep_relative_accuracy():
    """p_label -> 1 (most tokens late in RL training): logp ~ -1e-6, H ~ 1e-5 and the label gradient
    g*(1-p) are cancellation results.  The kernels subtract the row max first (like log_softmax), so
    these must be right to a few percent RELATIVE, not just 1e-3 absolute."""
    from areal_b200 import functional as F
    g = torch.Generator().manual_seed(21)
    T, V = 12, 151936
    logits = (torch.randn(T, V, generator=g) * 2.0).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g)
    for t in range(T):
        logits[t, labels[t]] = 21.0 + 0.5 * t      # 1 - p_label from ~8e-4 down to ~3e-6 (fp32 eps: 1.2e-7)
    x = logits.to(D).requires_grad_()
    lp, ent = F.gather_logprobs_entropy(x, labels.to(D), 1.0)
    ref_lp, ref_ent, *_ = O.gather_logprobs_entropy(logits.float().numpy(), labels.numpy(), 1.0)
    assert np.all(np.abs(lp.detach().cpu().numpy() - ref_lp) <= 0.05 * np.abs(ref_lp) + 1.5e-7)
    assert np.all(np.abs(ent.detach().cpu().numpy() - ref_ent) <= 0.05 * np.abs(ref_ent) + 2e-7)
    (grad,) = torch.autograd.grad(lp.sum(), x)
    ref_g = O.dlogits(logits.float().numpy(), labels.numpy(), 1.0, np.ones(T))
    got = grad.float().cpu().numpy()
    lab = labels.numpy()
    at_label = got[np.arange(T), lab]
    assert np.all(np.abs(at_label - ref_g[np.arange(T), lab]) <= 0.05 * np.abs(ref_g[np.arange(T), lab]) + 1.5e-7)


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32])
@pytest.mark.parametrize("V", [18992, 4099])
def test_backward_other_dtypes(dtype, V):
    """fp16 / fp32 logits (Megatron can hand fp32, areal/utils/data.py:762-763): TMA-staged kernels when
    rows are 16-byte aligned, generic kernel otherwise; gradient in the logits' dtype."""
    from areal_b200 import functional as F
    g = torch.Generator().manual_seed(V + 1)
    T = 41
    logits = (torch.randn(T, V, generator=g) * 2.0).to(dtype)
    labels = torch.randint(0, V, (T,), generator=g)
    labels[0], labels[1] = 0, V - 1
    x = logits.to(D).requires_grad_()
    lp, ent = F.gather_logprobs_entropy(x, labels.to(D), 0.8)
    w = torch.randn(T, generator=g)
    wh = torch.randn(T, generator=g) * 0.1
    ((lp * w.to(D)).sum() + (ent * wh.to(D)).sum()).backward()
    assert x.grad.dtype == dtype
    ref = O.dlogits(logits.float().numpy(), labels.numpy(), 0.8, w.numpy(), wh.numpy())
    got = x.grad.float().cpu().numpy().astype(np.float64)
    scale = np.maximum(np.abs(ref).max(-1, keepdims=True), 1e-3 * (np.abs(w.numpy()) + 12 * np.abs(wh.numpy()))[:, None] / 0.8)
    ulp = 2 ** -11 if dtype == torch.float16 else 2 ** -23
    assert np.all(np.abs(got - ref) <= ulp * np.abs(ref) + H.TOL_LOGITS * scale + 1e-30)


def test_full_size_runs_are_bit_identical():
    """Full-size config 2, three runs of every logits path (split forward + backward, fused two-pass,
    fused single-pass) with the allocator's free memory overwritten in between: outputs and gradients
    must not change by a bit.  Catches shared-memory ring hand-off races (a stage released before its
    last 128-byte wavefront was read shows up as a few wrong 64-column runs per launch)."""
    from areal_b200 import functional as F
    T, V = 16384, 151936
    g = torch.Generator(device=D).manual_seed(5)
    logits = torch.empty(T, V, dtype=torch.bfloat16, device=D)
    for i in range(0, T, 2048):
        logits[i:i + 2048] = (torch.randn(2048, V, generator=g, device=D) * 2.0).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g, device=D)
    w = torch.randn(T, generator=g, device=D)
    w[torch.rand(T, generator=g, device=D) < 0.3] = 0.0
    old = -torch.rand(T, generator=g, device=D) * 12 - 6
    prox = old + 0.1 * torch.randn(T, generator=g, device=D)
    adv = torch.randn(T, generator=g, device=D)
    mask = torch.rand(T, generator=g, device=D) > 0.2
    scale = torch.full((1,), 0.5, device=D)

    def bits(t):
        return t.view(torch.int16) if t.dtype == torch.bfloat16 else t.view(torch.int32)

    def run(path):
        x = logits.detach().requires_grad_()
        if path == "split":
            lp, ent = F.gather_logprobs_entropy(x, labels, 1.0)
            (grad,) = torch.autograd.grad((lp * w).sum(), x)
            return [lp.detach(), ent.detach(), grad]
        loss, st = F.fused_actor_loss(x, labels, 1.0, prox, old, adv, mask, eps_clip=0.4, behav_imp_weight_cap=5.0,
                                      loss_scale=scale if path == "single" else None)
        (loss * scale[0]).backward()
        return [loss.detach().reshape(1), st["logprobs"], st["entropy"], x.grad]

    for path in ("split", "twopass", "single"):
        ref = None
        for it in range(3):
            junk = torch.empty(T, V, dtype=torch.bfloat16, device=D)
            junk.view(torch.int16).fill_(0x7FC1 if it % 2 else 0x4142)
            del junk
            cur = run(path)
            if ref is None:
                ref = [c.clone() for c in cur]
                if path == "split":
                    assert int(torch.count_nonzero(cur[2][w == 0])) == 0
            else:
                for a, b in zip(cur, ref):
                    assert torch.equal(bits(a), bits(b)), (path, it)
            del cur
        del ref
