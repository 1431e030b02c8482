"""CPU: the oracle restatement (oracle/cpu_oracle.py, oracle/oracle.c) against the golden vectors
produced by the unmodified reference (oracle/gen_golden.py), and against the reference's own
known-answer test for the dual-clip loss (realhf/tests/data/test_dual_clip.py)."""
import numpy as np
import pytest

from oracle import c_oracle as C
from oracle import cpu_oracle as O
from tests import helpers as H


def test_gae1d_bit_exact_vs_reference():
    z = H.load("gae1d.npz")
    n = 0
    for k, c in H.cases(z):
        g, l = c["gl"]
        a, r = O.gae1d_nolp_misalign(c["rewards"], c["values"], c["cu"], c["boot"], float(g), float(l))
        assert np.array_equal(a, c["adv"]), k
        assert np.array_equal(r, c["ret"]), k
        a2, r2 = C.gae1d_nolp_misalign(c["rewards"], c["values"], c["cu"], c["boot"], float(g), float(l))
        assert np.array_equal(a2, c["adv"]) and np.array_equal(r2, c["ret"]), k
        a64, _ = O.gae1d_nolp_misalign_f64(c["rewards"], c["values"], c["cu"], c["boot"], float(g), float(l))
        assert H.rel_err(a64, c["adv"], floor=max(1.0, np.abs(c["adv"]).max(initial=0))) < H.TOL_GAE
        n += 1
    assert n == 49


def test_gae_padded_bit_exact_vs_reference():
    z = H.load("gae_padded.npz")
    for k, c in H.cases(z):
        g, l = c["gl"]
        a, r = O.gae_padded_masked(c["rewards"], c["values"], c["mask"], c["no_eos"], float(g), float(l))
        assert np.array_equal(a, c["adv"]) and np.array_equal(r, c["ret"]), k
        a2, r2 = C.gae_padded_masked(c["rewards"], c["values"], c["mask"], c["no_eos"], float(g), float(l))
        assert np.array_equal(a2, c["adv"]) and np.array_equal(r2, c["ret"]), k


def test_gae_2d_olp_consistent_with_1d():
    """No reference test exists for the 2D kernels (their own outputs: test_gae2d_golden_vs_oracle); check the restatement
    reduces to the pinned 1D recurrence when every row is one bootstrapped segment."""
    rng = np.random.RandomState(3)
    B, Hh = 4, 33
    r = rng.randn(B, Hh).astype(np.float32)
    v = rng.randn(B, Hh + 1).astype(np.float32)
    done = np.zeros((B, Hh + 1), dtype=bool)
    adv, ret = O.gae_2d_olp(r, v, done, np.zeros(0, np.int32), np.zeros(B + 1, np.int32),
                            np.zeros(0, bool), 0.9, 0.5)
    cu = (np.arange(B + 1) * Hh).astype(np.int32)
    a1, r1 = O.gae1d_nolp_misalign(r.reshape(-1), v.reshape(-1), cu, np.ones(B, bool), 0.9, 0.5)
    assert np.allclose(adv.reshape(-1), a1, atol=1e-6) and np.allclose(ret.reshape(-1), r1, atol=1e-6)


def test_logprobs_entropy_and_grad_vs_reference():
    import torch
    z = H.load("logprobs.npz")
    for k, c in H.cases(z):
        logits = H.bf16_from_bits(c["logits_bits"]).float().numpy()
        temp = float(c["temp"])
        lp, ent, lse, mn, mx = O.gather_logprobs_entropy(logits, c["labels"], temp)
        assert H.rel_err(lp, c["logp"]) < 1e-5, k
        assert H.rel_err(lp, c["logp_only"]) < 1e-5, k
        assert H.rel_err(ent, c["entropy"]) < 1e-5, k
        lp32, ent32, *_ = O.gather_logprobs_entropy(logits, c["labels"], temp, dtype=np.float32)
        assert H.rel_err(lp32, c["logp"]) < 1e-5 and H.rel_err(ent32, c["entropy"]) < 1e-4, k
        g = O.dlogits(logits, c["labels"], temp, c["g_lp"])
        ref = H.bf16_from_bits(c["grad_lp_bits"]).float().numpy()
        # the reference grad is rounded to bf16 (2^-9 relative)
        assert np.max(np.abs(g - ref) / np.maximum(np.abs(ref), 1e-3)) < 2 ** -7, k
        g2 = O.dlogits(logits, c["labels"], temp, c["g_lp"], c["g_H"])
        ref2 = H.bf16_from_bits(c["grad_both_bits"]).float().numpy()
        assert np.max(np.abs(g2 - ref2) / np.maximum(np.abs(ref2), 1e-3)) < 2 ** -7, k
        assert np.array_equal(mn, logits.min(-1)) and np.array_equal(mx, logits.max(-1))


def test_vocab_parallel_combine_matches_unsharded():
    rng = np.random.RandomState(0)
    T, V, n = 9, 1000, 8
    logits = (rng.randn(T, V) * 3).astype(np.float32)
    labels = rng.randint(0, V, T)
    lp, ent, lse, *_ = O.gather_logprobs_entropy(logits, labels, 0.7)
    parts = [O.vocab_parallel_partials(logits[:, r * V // n:(r + 1) * V // n], labels, r * V // n, 0.7)
             for r in range(n)]
    lp2, ent2, lse2 = O.vocab_parallel_combine(parts)
    assert np.allclose(lp, lp2, atol=1e-10) and np.allclose(ent, ent2, atol=1e-10) and np.allclose(lse, lse2)


def _check_stats(st, c, k):
    for name in ("loss", "importance_weight", "approx_kl", "behave_imp_weight", "behave_approx_kl"):
        if "stat_" + name in c:
            assert np.allclose(st[name], c["stat_" + name], rtol=1e-6, atol=1e-7), (k, name)
    for name in ("clip_mask", "dual_clip_mask", "behave_mask"):
        if "stat_" + name in c:
            assert np.array_equal(st[name], c["stat_" + name]), (k, name)


def test_actor_loss_vs_reference():
    z = H.load("actor_loss.npz")
    for k, c in H.cases(z):
        eps, eps_hi, c_clip, cap = (H.opt(x) for x in c["cfg"])
        loss, st, grad = O.ppo_actor_loss(c["lp"], c["prox"], c["old"], c["adv"], eps, c["mask"],
                                          eps_clip_higher=eps_hi, c_clip=c_clip, behav_imp_weight_cap=cap)
        assert abs(loss - float(c["loss"])) <= 1e-5 * max(1.0, abs(float(c["loss"]))), k
        assert np.allclose(grad, c["grad"], rtol=1e-5, atol=1e-8), k
        _check_stats(st, c, k)
    for k, c in H.cases(z, prefix="l", count_key="n_legacy"):
        eps, c_clip, cap = (H.opt(x) for x in c["cfg"])
        prox = c["prox"] if bool(c["has_prox"]) else None
        loss, st, grad = O.ppo_actor_loss(c["lp"], prox, c["old"], c["adv"], eps, c["mask"],
                                          c_clip=c_clip, behav_imp_weight_cap=cap)
        assert abs(loss - float(c["loss"])) <= 1e-5 * max(1.0, abs(float(c["loss"]))), k
        assert np.allclose(grad, c["grad"], rtol=1e-5, atol=1e-8), k
        _check_stats(st, c, k)
        assert ("behave_mask" in st) == bool(c["has_prox"])


@pytest.mark.parametrize("eps_clip", [0.01, 0.05, 0.1, 0.2, 0.5])
@pytest.mark.parametrize("c_clip", [None, 1.5, 2.0, 5.0])
@pytest.mark.parametrize("size", [1, 10, 100])
def test_dual_clip_known_answer(size, eps_clip, c_clip):
    """Same grid and independent formula as realhf/tests/data/test_dual_clip.py:53-63."""
    rng = np.random.RandomState(size * 7 + int(eps_clip * 100))
    old = -np.abs(rng.randn(size)).astype(np.float32)
    new = -np.abs(rng.randn(size)).astype(np.float32)
    adv = rng.randn(size).astype(np.float32)
    want = O.ppo_dual_clip_known_answer(new, old, adv, eps_clip, c_clip)
    got, _, _ = O.ppo_actor_loss(new, None, old, adv, eps_clip, np.ones(size, bool), c_clip=c_clip)
    assert np.isclose(got, want, rtol=1e-5, atol=1e-7)


def test_critic_loss_vs_reference():
    z = H.load("critic_loss.npz")
    for k, c in H.cases(z):
        mask = c["mask"] if bool(c["has_mask"]) else None
        loss, st, grad = O.ppo_critic_loss(c["value"], c["old"], c["target"], float(c["eps"]), mask,
                                           "huber" if bool(c["huber"]) else "mse")
        assert abs(loss - float(c["loss"])) <= 1e-5 * max(1.0, abs(float(c["loss"]))), k
        assert np.allclose(grad, c["grad"], rtol=1e-5, atol=1e-9), k
        assert np.array_equal(st["clip_mask"], c["stat_clip_mask"]), k
        assert np.allclose(st["loss"], c["stat_loss"], rtol=1e-6, atol=1e-7), k


def test_intervals_bit_exact_vs_reference():
    z = H.load("intervals.npz")
    for k, c in H.cases(z):
        iv = c["intervals"]
        assert np.array_equal(O.slice_intervals(c["src"], iv), c["packed"]), k
        assert np.array_equal(C.slice_intervals(c["src"], iv), c["packed"]), k
        d = c["dst"].copy()
        O.set_intervals(c["newvals"], d, iv)
        assert np.array_equal(d, c["dst_after"]), k
        d = c["dst"].copy()
        C.set_intervals(c["newvals"], d, iv)
        assert np.array_equal(d, c["dst_after"]), k


def test_merge_intervals_vs_reference_cpp():
    cases = [[], [(0, 4)], [(0, 4), (4, 9), (9, 10)], [(5, 7), (0, 5), (7, 9), (9, 9), (9, 12)],
             [(0, 2), (3, 5), (5, 8), (1, 3), (3, 4)]]
    rng = np.random.RandomState(1)
    pts = np.sort(rng.choice(10000, 400, replace=False))
    cases.append([(int(a), int(b)) for a, b in zip(pts[:-1], pts[1:]) if rng.rand() > 0.2])
    want = [[], [(0, 4)], [(0, 10)], [(5, 7), (0, 5), (7, 12)], [(0, 2), (3, 8), (1, 4)]]
    for i, iv in enumerate(cases):
        got = O.merge_intervals(iv)
        assert got == C.merge_intervals(iv) if iv else got == []
        if i < len(want):
            assert got == want[i]
    # the reference's own C++ (built from /root/reference by oracle/build_oracle.py) when present
    import glob, importlib.util, os
    so = glob.glob(os.path.join(H.ROOT, "oracle", "_ref", "interval_op*.so"))
    if so:
        spec = importlib.util.spec_from_file_location("interval_op", so[0])
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        for iv in cases:
            assert [tuple(x) for x in mod.merge_intervals(iv)] == O.merge_intervals(iv)


def test_cpu_baseline_port_matches_golden():
    """oracle/cpu_baseline.py (the timed CPU port of the reference path) against the reference's outputs."""
    import torch
    from oracle import cpu_baseline as CB
    z = H.load("logprobs.npz")
    for k, c in H.cases(z):
        logits = H.bf16_from_bits(c["logits_bits"])
        lp, ent = CB.logp_entropy_torch(logits, torch.from_numpy(c["labels"]), float(c["temp"]))
        assert np.allclose(lp.numpy(), c["logp"], rtol=1e-6, atol=1e-6), k
        assert np.allclose(ent.numpy(), c["entropy"], rtol=1e-5, atol=1e-6), k
    z = H.load("actor_loss.npz")
    for k, c in H.cases(z):
        eps, eps_hi, c_clip, cap = (H.opt(x) for x in c["cfg"])
        t = lambda a: torch.from_numpy(a)
        loss = CB.actor_loss_torch(t(c["lp"]), t(c["prox"]), t(c["old"]), t(c["adv"]), eps, t(c["mask"]), eps_hi, c_clip, cap)
        assert abs(float(loss) - float(c["loss"])) <= 1e-6 * max(1.0, abs(float(c["loss"]))), k


def test_cpu_baseline_gae_loop_and_interval_ports_match_golden():
    """The two timed CPU legs BASELINE.md section 5 names -- pygae1d_nolp_misalign's per-token torch loop and
    `_slice_intervals_py` / `_set_intervals_py` -- as restated in oracle/cpu_baseline.py, against the
    reference's own outputs: bit for bit."""
    import torch
    from oracle import cpu_baseline as CB
    z = H.load("gae1d.npz")
    n = 0
    for k, c in H.cases(z):
        if c["rewards"].size > 3000:           # the loop runs at ~10 k tokens/s: keep the CPU suite short
            continue
        g, l = c["gl"]
        a, r = CB.gae1d_torch_loop(torch.from_numpy(c["rewards"]), torch.from_numpy(c["values"]).clone(),
                                   torch.from_numpy(c["cu"]), torch.from_numpy(c["boot"].astype(np.float32)),
                                   float(g), float(l))
        assert np.array_equal(a.numpy(), c["adv"]) and np.array_equal(r.numpy(), c["ret"]), k
        n += 1
    assert n >= 10
    z = H.load("intervals.npz")
    for k, c in H.cases(z):
        iv = [(int(a), int(b)) for a, b in c["intervals"]]
        src = torch.from_numpy(c["src"].view(np.int16) if c["src"].dtype == np.uint16 else c["src"])
        got = CB.slice_intervals_torch(src, iv).numpy()
        assert np.array_equal(got.view(c["packed"].dtype), c["packed"]), k
        d = torch.from_numpy((c["dst"].view(np.int16) if c["dst"].dtype == np.uint16 else c["dst"]).copy())
        nv = torch.from_numpy(c["newvals"].view(np.int16) if c["newvals"].dtype == np.uint16 else c["newvals"])
        CB.set_intervals_torch(nv, d, iv)
        assert np.array_equal(d.numpy().view(c["dst_after"].dtype), c["dst_after"]), k


def test_legacy_packed_vs_reference():
    """realhf spelling of P2 (gather_packed_shifted_log_probs, build_*_indices, calc_entropy)."""
    z = H.load("legacy_packed.npz")
    for k, c in H.cases(z):
        bf16 = bool(c["is_bf16"])
        logits = H.bf16_from_bits(c["logits"]).float().numpy() if bf16 else c["logits"]
        T = logits.shape[0]
        assert np.array_equal(O.leave_one_indices(T, c["cu"]), c["leave_one"])
        assert np.array_equal(O.shift_one_indices(T, c["cu"]), c["shift_one"])
        lp = O.gather_packed_shifted_log_probs(logits, c["cu"], c["ids"])
        ent = O.calc_entropy_packed(logits, c["cu"])
        if bf16:   # the reference computes and returns bf16 here: one bf16 ulp; its bf16 softmax is coarser still
            assert np.all(np.abs(lp - c["logp"]) <= 2 ** -7 * np.abs(c["logp"]) + 1e-6), k
            assert np.all(np.abs(ent - c["entropy"]) <= 0.05 * np.abs(c["entropy"]) + 1e-3), k
        else:
            assert np.allclose(lp, c["logp"], rtol=1e-5, atol=1e-5), k
            assert np.allclose(ent, c["entropy"], rtol=1e-5, atol=1e-5), k


_LEVEL = {0: "batch", 1: "group", 2: None}


def test_normalization_vs_reference():
    """`Normalization` over the grid of the reference's own tests, and `masked_normalization`."""
    z = H.load("normalization.npz")
    n = 0
    for k, c in H.cases(z):
        ml, sl, loo, unb, gs = (int(v) for v in c["cfg"])
        mask = c["mask"] if bool(c["has_mask"]) else None
        y = O.normalization(c["x"], mask, _LEVEL[ml], _LEVEL[sl], bool(loo), bool(unb), 1e-5, gs)
        assert y.dtype == np.float32
        assert np.allclose(y, c["y"], rtol=2e-6, atol=2e-6), (k, c["cfg"], np.abs(y - c["y"]).max())
        n += 1
    assert n == 304
    for k, c in H.cases(z, prefix="m", count_key="n_masked_norm"):
        mask = c["mask"] if bool(c["has_mask"]) else None
        y = O.masked_normalization(c["x"], mask, bool(c["unbiased"]))
        assert np.allclose(y, c["y"], rtol=2e-6, atol=2e-6), k


def test_compute_advantages_oracle_vs_reference():
    """oracle compute_advantages == PPOActor.compute_advantages run unmodified (cases without a
    Normalization: the scaling / KL / reward placement / GAE are bit-exact; k3 within 2 ulp of exp)."""
    from tests import adv_cases as A
    n = 0
    for c, ins, outs in A.load():
        if c["reward_norm"]:
            continue
        score = ins["rewards"]
        if c["overlong_reward_penalty"]:
            score = O.overlong_penalty(score, ins["loss_mask"], c["overlong_tokens"], c["overlong_penalty_factor"],
                                       c["max_new_tokens"])
            assert np.array_equal(score, outs["rewards"])
        swap = (not c["use_decoupled_loss"]) and c["recompute_logprob"]
        r = O.compute_advantages(score, ins["loss_mask"], ins["prox_logp"] if swap else ins["logprobs"],
                                 ins["attention_mask"], ref_logp=ins.get("ref_logp"), values=ins.get("values"),
                                 roll_logp=not swap, reward_bias=c["reward_bias"], reward_scaling=c["reward_scaling"],
                                 reward_clip=c["reward_clip"], kl_ctl=c["kl_ctl"], kl_estimator=c["kl_estimator"],
                                 mask_no_eos_with_zero=c["mask_no_eos_with_zero"], discount=c["discount"],
                                 gae_lambda=c["gae_lambda"])
        for k in ("loss_mask", "logprobs"):
            assert np.array_equal(r[k], outs[k]), k
        if "ref_logp" in outs:
            assert np.array_equal(r["ref_logp"], outs["ref_logp"])
        if c["kl_estimator"] == "k3":
            np.testing.assert_allclose(r["kl_rewards"], outs["kl_rewards"], rtol=3e-7, atol=1e-7)
            np.testing.assert_allclose(r["tot_rewards"], outs["tot_rewards"], rtol=3e-7, atol=1e-7)
            A.assert_gae_close(r["returns"], outs["returns"], "returns")
        else:
            for k in ("kl_rewards", "tot_rewards", "returns"):
                assert np.array_equal(r[k], outs[k]), (k, c)
            if not c["adv_norm"]:
                assert np.array_equal(r["advantages"], outs["advantages"])
        n += 1
    assert n >= 7


def test_packed_rewards_oracle_bit_exact_vs_reference():
    z = H.load("packed_rewards.npz")
    for k in range(int(z["n_cases"])):
        kl_ctl, clip, mz, delta = [float(v) for v in z[f"c{k}_cfg"]]
        a, b = O.packed_rewards(kl_ctl, clip, z[f"c{k}_lp"], z[f"c{k}_rf"], z[f"c{k}_score"], z[f"c{k}_cu"],
                                z[f"c{k}_ne"], bool(mz))
        assert np.array_equal(a, z[f"c{k}_kl"]) and np.array_equal(b, z[f"c{k}_tot"])
        c, d = O.packed_rewards(kl_ctl, clip, z[f"c{k}_lp"], z[f"c{k}_rf"], z[f"c{k}_dense"], z[f"c{k}_cu"],
                                dense=True, reward_delta=bool(delta))
        assert np.array_equal(c, z[f"c{k}_dkl"]) and np.array_equal(d, z[f"c{k}_dtot"])


def test_running_mean_std_oracle_vs_reference():
    z = H.load("rms.npz")
    for k in range(int(z["n_cases"])):
        exp, hp = [bool(v) for v in z[f"c{k}_cfg"]]
        mod = O.RunningMeanStd(exponential=exp, beta=0.9, epsilon=1e-5, high_precision=hp)
        tol = 1e-13 if hp else 3e-6
        for step in range(3):
            mod.update(z[f"x{step}"], z[f"m{step}"] if f"m{step}" in z else None)
            mean, std = mod.mean_std()
            np.testing.assert_allclose(mean, z[f"c{k}_mean{step}"][0], rtol=tol, atol=tol)
            np.testing.assert_allclose(std, z[f"c{k}_std{step}"][0], rtol=tol, atol=tol)
            np.testing.assert_allclose(mod.normalize(z["probe"]), z[f"c{k}_norm{step}"], rtol=3e-6, atol=3e-6)
            np.testing.assert_allclose(mod.denormalize(z["probe"]), z[f"c{k}_denorm{step}"], rtol=3e-6, atol=3e-6)


def test_clamped_entropy_oracle_vs_reference():
    """oracle == gather_logprobs_clamped_entropy + autograd run unmodified (recipe/AEnt/functional.py)."""
    from tests import clamp_cases as CC
    for c in CC.load():
        x = H.bf16_from_bits(c["bits"]).float().numpy()
        lp, ent = O.gather_logprobs_clamped_entropy(x, c["labels"], c["clamp"], c["temp"])
        np.testing.assert_allclose(lp, c["logp"], rtol=2e-5, atol=2e-5)
        np.testing.assert_allclose(ent, c["ent"], rtol=2e-5, atol=2e-5)
        g = O.clamped_entropy_dlogits(x, c["labels"], c["clamp"], c["temp"], c["w_lp"], c["w_h"])
        CC.check_grad(g, c["grad"].astype(np.float64), x, c["clamp"], 2e-5, "oracle")


def test_lm_head_golden_vs_oracle():
    """The LM-head golden (reference run on hidden @ weight.T, oracle/gen_golden.py::gen_lm_head): the oracle applied to
    the exactly rounded logits reproduces the reference's log-probs, entropy, loss, and -- through the fp64 chain
    rule -- its gradients w.r.t. hidden state and weight up to their bf16 rounding."""
    import torch
    z = H.load("lm_head.npz")
    for k, c in H.cases(z):
        h = H.bf16_from_bits(c["h_bits"]).double().numpy()
        w = H.bf16_from_bits(c["w_bits"]).double().numpy()
        logits = torch.from_numpy(h @ w.T).float().bfloat16().float().numpy()
        temp, scale = float(c["temp"]), float(c["scale"])
        lp, ent, *_ = O.gather_logprobs_entropy(logits, c["labels"], temp)
        assert np.allclose(lp, c["logp"], rtol=1e-5, atol=1e-5) and np.allclose(ent, c["entropy"], rtol=1e-5, atol=1e-5)
        cfg = {name: (float(c["cfg_" + name]) if "cfg_" + name in c else None)
               for name in ("eps_clip", "eps_clip_higher", "c_clip", "behav_imp_weight_cap")}
        loss, st, coef = O.ppo_actor_loss(lp, c["prox"].astype(np.float64), c["old"].astype(np.float64),
                                          c["adv"].astype(np.float64), loss_mask=c["mask"], dtype=np.float64, **cfg)
        assert abs(loss - float(c["loss"])) <= 1e-5 * max(1.0, abs(float(c["loss"])))
        assert np.array_equal(st["clip_mask"], c["clip_mask"])
        dl = O.dlogits(logits, c["labels"], temp, scale * coef)
        for ref_bits, mine in ((c["dh_bits"], dl @ w), (c["dw_bits"], dl.T @ h)):
            ref = H.bf16_from_bits(ref_bits).double().numpy()
            assert np.all(np.abs(mine - ref) <= 2 ** -6 * np.abs(mine) + 2 ** -7 * np.abs(mine).max() + 1e-30), k


def test_vocab_parallel_golden_vs_oracle():
    """Golden from the reference's OWN vocabulary-sharded code on two gloo ranks (oracle/gen_golden_tp.py:
    `_VocabParallelEntropy` + the tree's copy of Megatron's vocab_parallel_cross_entropy, as
    gather_logprobs_entropy runs them under tensor parallelism, areal/utils/functional.py:91-98): the oracle's
    per-shard partials + combine reproduce its log-probs and entropy, the unsharded oracle gradient its per-shard
    gradients."""
    z = H.load("vocab_parallel.npz")
    world = int(z["world"])
    for k, c in H.cases(z):
        logits = H.bf16_from_bits(c["logits_bits"]).float().numpy()
        labels, temp = c["labels"], float(c["temp"])
        Vs = logits.shape[1] // world
        parts = np.stack([O.shard_partials(logits[:, r * Vs:(r + 1) * Vs], labels, r * Vs, temp) for r in range(world)])
        logp, ent, row_max, ln_s = O.combine_partials(parts, temp)
        assert np.allclose(logp, c["logp"], rtol=2e-5, atol=2e-5), k
        assert np.allclose(ent, c["entropy"], rtol=2e-5, atol=2e-5), k
        ref = O.dlogits(logits, labels, temp, c["g_lp"].astype(np.float64), c["g_h"].astype(np.float64))
        for r in range(world):
            got = H.bf16_from_bits(c[f"grad{r}_bits"]).double().numpy()
            want = ref[:, r * Vs:(r + 1) * Vs]
            # the reference forms p - 1 at the label in fp32 (a cancellation on peaked rows; exactly 0 when p rounds to 1):
            # one bf16 ulp of the value plus 1e-6 of the upstream gradient's size
            slack = 1e-6 * (np.abs(c["g_lp"]) + 12 * np.abs(c["g_h"]))[:, None] / temp
            assert np.all(np.abs(got - want) <= 2 ** -7 * np.abs(want) + slack + 1e-30), (k, r)


def test_gae2d_golden_vs_oracle():
    """tests/golden/gae2d.npz holds outputs of the reference's own `gae_2d_olp` / `gae_2d_nolp` CUDA kernels
    (csrc/cugae/gae.cu compiled from where it lies, run on a B200 by oracle/gen_golden_gpu.py): the restatement
    follows them to fp32 rounding, and the two entry points agree bit for bit (the `_nolp` host wrapper launches
    the `_olp` kernel, gae.cu:199)."""
    z = H.load("gae2d.npz")
    for k, c in H.cases(z):
        a, r = O.gae_2d_olp(c["rewards"], c["values"], c["done"], c["idx"], c["cu"], c["trunc"], float(c["gamma"]),
                            float(c["lam"]))
        scale = max(1.0, float(np.abs(c["adv_olp"]).max()))
        assert np.abs(a - c["adv_olp"]).max() <= 1e-5 * scale and np.abs(r - c["ret_olp"]).max() <= 1e-5 * scale, k
        assert np.array_equal(c["adv_olp"], c["adv_nolp"]) and np.array_equal(c["ret_olp"], c["ret_nolp"]), k
