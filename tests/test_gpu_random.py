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


def test_fused_actor_random_shapes_single_and_two_pass():
    """fused_actor_loss with and without loss_scale (single-pass row-resident kernel / two-pass) on random
    T, V (multiples of 8, so rows are 16-byte aligned), temperatures and clip settings: loss vs the fp64
    oracle, the two paths against each other."""
    from areal_b200 import functional as F
    rng = np.random.RandomState(4242)
    for case in range(16):
        T = int(rng.randint(1, 400))
        V = 8 * int(rng.choice([rng.randint(1, 64), rng.randint(64, 4000), rng.randint(4000, 19000)]))
        temp = float(rng.choice([1.0, 0.8, 1.3]))
        g = torch.Generator().manual_seed(1000 + case)
        logits = (torch.randn(T, V, generator=g) * 2.5).bfloat16()
        labels = torch.randint(0, V, (T,), generator=g)
        lp_ref, *_ = O.gather_logprobs_entropy(logits.float().numpy(), labels.numpy(), temp)
        prox = (torch.from_numpy(lp_ref) + 0.05 * torch.randn(T, generator=g)).float()
        old = (torch.from_numpy(lp_ref) + 0.1 * torch.randn(T, generator=g)).float()
        adv = torch.randn(T, generator=g)
        mask = torch.rand(T, generator=g) > float(rng.choice([0.0, 0.3, 0.9]))
        cfg = dict(eps_clip=float(rng.choice([0.1, 0.2, 0.4])), eps_clip_higher=rng.choice([None, 0.5]),
                   c_clip=rng.choice([None, 3.0]), behav_imp_weight_cap=rng.choice([None, 2.0, 5.0]))
        cfg = {k: (None if v is None else float(v)) for k, v in cfg.items()}
        scale = float(rng.choice([1.0, 0.25, 0.01]))
        res = {}
        for path in ("single", "twopass"):
            x = logits.to(D).requires_grad_()
            loss, st = F.fused_actor_loss(x, labels.to(D), temp, prox.to(D), old.to(D), adv.to(D), mask.to(D),
                                          loss_scale=scale if path == "single" else None, **cfg)
            (loss * scale).backward()
            res[path] = (float(loss.detach()), st["logprobs"].cpu().numpy(), x.grad.float().cpu().numpy())
        want, _, coef = O.ppo_actor_loss(lp_ref, prox.numpy().astype(np.float64), old.numpy().astype(np.float64),
                                         adv.numpy().astype(np.float64), cfg["eps_clip"], mask.numpy(),
                                         eps_clip_higher=cfg["eps_clip_higher"], c_clip=cfg["c_clip"],
                                         behav_imp_weight_cap=cfg["behav_imp_weight_cap"], dtype=np.float64)
        for path in res:
            assert abs(res[path][0] - want) <= 2e-3 * max(1.0, abs(want)), (case, path, T, V, res[path][0], want)
            assert H.rel_err(res[path][1], lp_ref) <= H.TOL_LOGITS, (case, path)
        assert np.allclose(res["single"][2], res["twopass"][2], rtol=2 ** -7, atol=1e-12), (case, T, V)


def test_compute_advantages_random_configs():
    import types
    from areal_b200.actor import compute_advantages
    rng = np.random.RandomState(31337)
    for case in range(24):
        B, S = int(rng.randint(1, 50)), int(rng.choice([1, 2, 3, 8, 16, 40, 512, 1000, 1024, rng.randint(2, 3000)]))
        lens = rng.randint(0, S + 1, B)
        am = np.arange(S)[None, :] < lens[:, None]
        lm = am & (rng.rand(B, S) > 0.3)
        logp = -rng.rand(B, S).astype(np.float32) * 5
        ref = -rng.rand(B, S).astype(np.float32) * 5
        val = rng.randn(B, S).astype(np.float32)
        score = (rng.randn(B) * 4).astype(np.float32)
        kl = str(rng.choice(["k1", "k2", "k3"]))
        with_ref, with_val = bool(rng.rand() < 0.7), bool(rng.rand() < 0.7)
        cfgv = dict(reward_bias=float(rng.choice([0.0, -0.5])), reward_scaling=float(rng.choice([1.0, 2.0])),
                    reward_clip=float(rng.choice([20.0, 1.5])), kl_ctl=float(rng.choice([0.0, 0.1, 1.0])),
                    discount=float(rng.choice([1.0, 0.99])), gae_lambda=float(rng.choice([1.0, 0.9])),
                    mask_no_eos_with_zero=bool(rng.rand() < 0.5))
        mdt = [torch.bool, torch.float32, torch.int64][case % 3]
        data = dict(input_ids=torch.zeros(B, S, dtype=torch.long, device=D),
                    attention_mask=torch.from_numpy(am).to(D).to(mdt), loss_mask=torch.from_numpy(lm).to(D).to(mdt),
                    logprobs=torch.from_numpy(logp).to(D), rewards=torch.from_numpy(score).to(D))
        if with_ref:
            data["ref_logp"] = torch.from_numpy(ref).to(D)
        if with_val:
            data["values"] = torch.from_numpy(val).to(D)
        actor = types.SimpleNamespace(
            config=types.SimpleNamespace(use_decoupled_loss=True, recompute_logprob=True, overlong_reward_penalty=False),
            kl_estimator=types.SimpleNamespace(kl_estimator=kl, apply_clamp=True), adv_norm=None, reward_norm=None, **cfgv)
        compute_advantages(actor, data)
        r = O.compute_advantages(score, lm, logp, am, ref_logp=ref if with_ref else None, values=val if with_val else None,
                                 roll_logp=True, kl_estimator=kl, **cfgv)
        for k in ("loss_mask", "logprobs"):
            assert np.array_equal(data[k].cpu().numpy(), r[k]), (case, k)
        if kl == "k3":
            np.testing.assert_allclose(data["tot_rewards"].cpu().numpy(), r["tot_rewards"], rtol=1e-5, atol=1e-6)
        else:
            assert np.array_equal(data["kl_rewards"].cpu().numpy(), r["kl_rewards"]), case
            assert np.array_equal(data["tot_rewards"].cpu().numpy(), r["tot_rewards"]), case
        for k in ("advantages", "returns"):
            got, want = data[k].cpu().numpy(), r[k]
            scale = np.maximum(1.0, np.abs(want).max(-1, keepdims=True))
            assert float((np.abs(got - want) / scale).max()) <= 10 * H.TOL_GAE, (case, k, B, S)
