"""GPU: seeded randomised sweeps against the oracle -- many small shapes with awkward sizes, offsets and
alignments that the hand-picked cases do not enumerate (views that start at odd byte offsets, empty and
one-token sequences, vocabularies that are not multiples of 8, strided rows)."""
import numpy as np
import pytest
import torch

from oracle import c_oracle as C
from oracle import cpu_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu
D = torch.device("cuda:0")


def _lens(rng, total_cap):
    kind = rng.randint(0, 4)
    n = int(rng.randint(1, 200))
    if kind == 0:
        lens = rng.randint(0, 4, n)                         # many empty / tiny sequences
    elif kind == 1:
        lens = rng.randint(1, 3000, n)
    elif kind == 2:
        lens = np.concatenate([rng.randint(0, 50, n), [int(rng.randint(2000, 60000))], rng.randint(0, 50, n)])
    else:
        lens = np.array([int(rng.randint(1, 100000))])
    while lens.sum() > total_cap:
        lens = lens[: max(1, len(lens) // 2)]
        if len(lens) == 1:
            lens = np.minimum(lens, total_cap)
    return lens.astype(np.int64)


def test_gae1d_random_batches():
    from areal_b200 import ppo_functional as P
    rng = np.random.RandomState(20240)
    for case in range(40):
        lens = _lens(rng, 250_000)
        cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        B, T = len(lens), int(cu[-1])
        gamma, lam = float(rng.choice([1.0, 0.99, 0.9, 0.5])), float(rng.choice([1.0, 0.95, 0.3, 0.0]))
        pad = int(rng.randint(0, 4))                        # views that start 4*pad bytes into a buffer
        rb = torch.randn(T + pad, device=D)
        vb = torch.randn(T + B + pad, device=D)
        r, v = rb[pad:], vb[pad:]
        boot = torch.rand(B, device=D) > 0.5
        if T == 0:
            continue
        adv, ret = P.cugae1d_nolp_misalign_func(r, v, torch.from_numpy(cu).to(D), boot, gamma, lam)
        a64, r64 = O.gae1d_nolp_misalign_f64(r.cpu().numpy(), v.cpu().numpy(), cu, boot.cpu().numpy(), gamma, lam)
        scale = max(1.0, float(np.abs(a64).max()))
        assert float(np.abs(adv.cpu().numpy() - a64).max()) <= 2e-7 * scale, (case, T, B)
        assert float(np.abs(ret.cpu().numpy() - r64).max()) <= 4e-7 * max(scale, float(np.abs(r64).max())), (case, T, B)


def test_intervals_random_views_and_dtypes():
    from areal_b200 import flatten_param as fp
    rng = np.random.RandomState(777)
    for case in range(60):
        es = int(rng.choice([1, 2, 4, 8]))
        tdt = {1: torch.uint8, 2: torch.int16, 4: torch.int32, 8: torch.int64}[es]
        npdt = {1: np.uint8, 2: np.int16, 4: np.int32, 8: np.int64}[es]
        n = int(rng.randint(1, 400_000))
        lead = int(rng.randint(0, 9))                       # flat view starts `lead` elements into its storage
        base = rng.randint(-100, 100, n + lead).astype(npdt)
        flat = torch.from_numpy(base).to(D)[lead:]
        k = int(rng.randint(0, 300))
        cuts = np.sort(rng.choice(n + 1, size=min(2 * k, n + 1), replace=False)) if k else np.array([], dtype=np.int64)
        iv = cuts[: 2 * (len(cuts) // 2)].reshape(-1, 2).astype(np.int64)      # disjoint, ascending
        if rng.rand() < 0.3 and len(iv):
            iv = np.concatenate([iv, iv[:1] * 0 + iv[0, 0]])[: len(iv)]          # keep shape; harmless
        if rng.rand() < 0.5:
            rng.shuffle(iv)
        if rng.rand() < 0.3 and len(iv):
            iv = np.concatenate([iv, np.array([[5, 5]]) if n >= 5 else np.zeros((0, 2), np.int64)])   # empty interval
        total = int((iv[:, 1] - iv[:, 0]).sum())
        tiv = torch.from_numpy(iv.reshape(-1, 2)).to(D)
        want = C.slice_intervals(base[lead:], iv)
        plead = int(rng.randint(0, 5))
        out = fp.slice_intervals_any(flat, tiv, total)
        assert np.array_equal(out.cpu().numpy(), want), (case, es, n, len(iv))
        new = rng.randint(-100, 100, total + plead).astype(npdt)
        dst = flat.clone()
        fp.set_intervals_any(torch.from_numpy(new).to(D)[plead:], dst, tiv)
        ref = base[lead:].copy()
        C.set_intervals(new[plead:], ref, iv)
        assert np.array_equal(dst.cpu().numpy(), ref), (case, es, n, len(iv))


def test_logits_random_shapes():
    from areal_b200 import functional as F
    rng = np.random.RandomState(99)
    for case in range(24):
        T = int(rng.randint(1, 70))
        V = int(rng.choice([rng.randint(2, 300), 8 * rng.randint(1, 3000), rng.randint(300, 40000)]))
        dtype = [torch.bfloat16, torch.float16, torch.float32][case % 3]
        temp = float(rng.choice([1.0, 0.6, 1.7]))
        extra = int(rng.choice([0, 0, 3, 8]))                # row stride > V: a column slice of a wider matrix
        g = torch.Generator().manual_seed(case)
        wide = (torch.randn(T, V + extra, generator=g) * 3).to(dtype)
        labels = torch.randint(0, V, (T,), generator=g)
        x = wide.to(D)[:, :V].requires_grad_()
        lp, ent = F.gather_logprobs_entropy(x, labels.to(D), temp)
        ref_lp, ref_ent, *_ = O.gather_logprobs_entropy(wide[:, :V].float().numpy(), labels.numpy(), temp)
        assert H.rel_err(lp.detach().cpu().numpy(), ref_lp) <= H.TOL_LOGITS, (case, T, V, dtype)
        assert H.rel_err(ent.detach().cpu().numpy(), ref_ent) <= H.TOL_LOGITS, (case, T, V, dtype)
        w = torch.randn(T, generator=g)
        (grad,) = torch.autograd.grad((lp * w.to(D)).sum(), x)
        ref_g = O.dlogits(wide[:, :V].float().numpy(), labels.numpy(), temp, w.numpy())
        got = grad.float().cpu().numpy().astype(np.float64)
        ulp = {torch.bfloat16: 2 ** -8, torch.float16: 2 ** -11, torch.float32: 2 ** -22}[dtype]
        scale = np.maximum(np.abs(ref_g).max(-1, keepdims=True), 1e-3 * np.abs(w.numpy())[:, None] / temp) + 1e-30
        assert np.all(np.abs(got - ref_g) <= ulp * np.abs(ref_g) + H.TOL_LOGITS * scale), (case, T, V, dtype)


def test_padded_gae_random_shapes():
    from areal_b200 import ppo_functional as P
    rng = np.random.RandomState(5)
    for case in range(24):
        B, S = int(rng.randint(1, 40)), int(rng.choice([1, 2, 7, 8, 64, 2048, 2049, 2056, 4100, rng.randint(3, 6000)]))
        r = rng.randn(B, S).astype(np.float32)
        v = rng.randn(B, S).astype(np.float32)
        m = (rng.rand(B, S) > rng.choice([0.0, 0.3, 0.95])).astype(np.float32)
        ne = rng.rand(B) > 0.5
        g, l = float(rng.choice([1.0, 0.99, 0.8])), float(rng.choice([1.0, 0.95, 0.5]))
        a, rt = P.gae_padded_masked(torch.from_numpy(r).to(D), torch.from_numpy(v).to(D), torch.from_numpy(m).to(D),
                                    torch.from_numpy(ne).to(D), g, l)
        ra, rr = O.gae_padded_masked(r, v, m, ne, g, l)
        scale = np.maximum(1.0, np.abs(ra).max(-1, keepdims=True))
        assert float((np.abs(a.cpu().numpy() - ra) / scale).max()) <= 5 * H.TOL_GAE, (case, B, S)
        assert float((np.abs(rt.cpu().numpy() - rr) / np.maximum(scale, np.abs(rr).max(-1, keepdims=True))).max()) <= 5 * H.TOL_GAE
