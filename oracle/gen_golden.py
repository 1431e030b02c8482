"""TEST INFRASTRUCTURE ONLY -- generates the golden vectors under `tests/golden/` by
executing the UNMODIFIED reference functions (via `oracle/ref_loader.py`) on seeded
inputs.  Run in the build container (needs /root/reference):

    python oracle/gen_golden.py

Each .npz stores the inputs and the reference's outputs, so the GPU box (which has no
/root/reference) can check both the oracle restatement and the CUDA path against them.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader as R  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def bf16_bits(t: torch.Tensor) -> np.ndarray:
    return t.detach().contiguous().view(torch.int16).numpy().view(np.uint16)


def gen_gae1d(pf):
    """Grid of realhf/tests/cpp_extensions/test_cugae.py:16-56 (+ mixed bootstrap, which the
    reference test never exercises, + zero/one-length sequences)."""
    out = {}
    k = 0
    g = torch.Generator().manual_seed(1234)
    for max_seqlen in (32, 128, 512):
        for bs in (2, 4):
            for gamma in (0.9, 1.0):
                for lam in (0.5, 1.0):
                    for boot_mode in ("ones", "mixed"):
                        seqlens = torch.randint(1, max_seqlen, (bs,), dtype=torch.int32, generator=g)
                        T = int(seqlens.sum())
                        rewards = torch.randn(T, generator=g)
                        values = torch.randn(T + bs, generator=g)
                        boot = (torch.ones(bs, dtype=torch.bool) if boot_mode == "ones"
                                else torch.rand(bs, generator=g) > 0.5)
                        cu = torch.nn.functional.pad(seqlens.cumsum(0), (1, 0)).int()
                        adv, ret = pf.pygae1d_nolp_misalign(rewards, values.clone(), cu, boot, gamma, lam)
                        p = f"c{k}_"
                        out[p + "rewards"] = rewards.numpy()
                        out[p + "values"] = values.numpy()
                        out[p + "cu"] = cu.numpy()
                        out[p + "boot"] = boot.numpy()
                        out[p + "gl"] = np.array([gamma, lam])
                        out[p + "adv"] = adv.numpy()
                        out[p + "ret"] = ret.numpy()
                        k += 1
    # BASELINE.json configs[0] in miniature (gamma=1, lam=0.95, longer sequences)
    seqlens = torch.randint(128, 1500, (6,), dtype=torch.int32, generator=g)
    T = int(seqlens.sum())
    rewards = torch.randn(T, generator=g)
    values = torch.randn(T + 6, generator=g)
    boot = torch.rand(6, generator=g) > 0.5
    cu = torch.nn.functional.pad(seqlens.cumsum(0), (1, 0)).int()
    adv, ret = pf.pygae1d_nolp_misalign(rewards, values.clone(), cu, boot, 1.0, 0.95)
    p = f"c{k}_"
    out.update({p + "rewards": rewards.numpy(), p + "values": values.numpy(), p + "cu": cu.numpy(),
                p + "boot": boot.numpy(), p + "gl": np.array([1.0, 0.95]),
                p + "adv": adv.numpy(), p + "ret": ret.numpy()})
    k += 1
    out["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "gae1d.npz"), **out)
    print("gae1d cases:", k)


def gen_gae_padded():
    out = {}
    g = torch.Generator().manual_seed(77)
    k = 0
    for (B, S) in ((3, 17), (5, 130), (8, 700)):
        for (gamma, lam) in ((1.0, 1.0), (1.0, 0.95), (0.99, 0.95)):
            rewards = torch.randn(B, S, generator=g)
            values = torch.randn(B, S, generator=g)
            lens = torch.randint(2, S + 1, (B,), generator=g)
            prompt = (lens.float() * torch.rand(B, generator=g) * 0.6).long()
            pos = torch.arange(S)[None, :]
            lm = ((pos >= prompt[:, None]) & (pos < lens[:, None])).float()
            lm = lm * (torch.rand(B, S, generator=g) > 0.1).float()   # holes inside the response
            lm = torch.roll(lm, shifts=-1, dims=-1)                     # actor.py:110
            no_eos = lens == S
            adv, ret = R.run_areal_padded_gae(rewards, values, lm, no_eos, gamma, lam)
            p = f"c{k}_"
            out.update({p + "rewards": rewards.numpy(), p + "values": values.numpy(),
                        p + "mask": lm.numpy(), p + "no_eos": no_eos.numpy(),
                        p + "gl": np.array([gamma, lam]), p + "adv": adv.numpy(), p + "ret": ret.numpy()})
            k += 1
    out["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "gae_padded.npz"), **out)
    print("gae_padded cases:", k)


def gen_logprobs(af):
    out = {}
    g = torch.Generator().manual_seed(2024)
    k = 0
    for (T, V, scale, temp) in ((32, 4099, 2.0, 1.0), (16, 4099, 8.0, 0.7), (1, 151936, 2.0, 1.0),
                                (7, 1000, 2.0, 0.7), (5, 8, 3.0, 1.0)):
        logits = (torch.randn(T, V, generator=g) * scale).bfloat16()
        labels = torch.randint(0, V, (T,), generator=g)
        if T >= 16:
            logits[0, labels[0]] += 30.0            # spike at the label
            logits[1, (labels[1] + 1) % V] += 30.0  # spike elsewhere
        lg = logits.clone().requires_grad_()
        lp, H = af.gather_logprobs_entropy(lg, labels, temp)
        lp_only = af.gather_logprobs(logits, labels, temp)
        g_lp = torch.randn(T, generator=g)
        g_H = torch.randn(T, generator=g) * 0.1
        (grad_lp,) = torch.autograd.grad((lp * g_lp).sum(), lg, retain_graph=True)
        (grad_both,) = torch.autograd.grad((lp * g_lp).sum() + (H * g_H).sum(), lg)
        p = f"c{k}_"
        out.update({p + "logits_bits": bf16_bits(logits), p + "labels": labels.numpy(),
                    p + "temp": np.array(temp), p + "logp": lp.detach().numpy(),
                    p + "logp_only": lp_only.numpy(), p + "entropy": H.detach().numpy(),
                    p + "g_lp": g_lp.numpy(), p + "g_H": g_H.numpy(),
                    p + "grad_lp_bits": bf16_bits(grad_lp), p + "grad_both_bits": bf16_bits(grad_both)})
        k += 1
    out["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "logprobs.npz"), **out)
    print("logprobs cases:", k)


def gen_actor_loss(af, pf):
    out = {}
    g = torch.Generator().manual_seed(99)
    k = 0
    T = 777
    settings = [
        dict(eps_clip=0.2, eps_clip_higher=None, c_clip=None, behav_imp_weight_cap=None),
        dict(eps_clip=0.4, eps_clip_higher=None, c_clip=None, behav_imp_weight_cap=5.0),
        dict(eps_clip=0.2, eps_clip_higher=0.28, c_clip=3.0, behav_imp_weight_cap=None),
        dict(eps_clip=0.2, eps_clip_higher=0.28, c_clip=3.0, behav_imp_weight_cap=1.02),
        dict(eps_clip=0.05, eps_clip_higher=None, c_clip=1.5, behav_imp_weight_cap=5.0),
    ]
    for s in settings:
        for mask_mode in ("rand", "none_valid"):
            lp0 = -torch.rand(T, generator=g) * 5
            lp = (lp0 + 0.3 * torch.randn(T, generator=g)).requires_grad_()
            prox = lp0 + 0.02 * torch.randn(T, generator=g)
            old = lp0 + 0.05 * torch.randn(T, generator=g)
            adv = torch.randn(T, generator=g)
            adv[::50] = 0.0
            mask = (torch.rand(T, generator=g) > 0.3) if mask_mode == "rand" else torch.zeros(T, dtype=torch.bool)
            loss, st = af.ppo_actor_loss_fn(lp, prox, old, adv, loss_mask=mask, **s)
            (grad,) = torch.autograd.grad(loss, lp, allow_unused=True)
            if grad is None:
                grad = torch.zeros(T)
            p = f"c{k}_"
            out.update({p + "lp": lp.detach().numpy(), p + "prox": prox.numpy(), p + "old": old.numpy(),
                        p + "adv": adv.numpy(), p + "mask": mask.numpy(),
                        p + "cfg": np.array([s["eps_clip"],
                                             np.nan if s["eps_clip_higher"] is None else s["eps_clip_higher"],
                                             np.nan if s["c_clip"] is None else s["c_clip"],
                                             np.nan if s["behav_imp_weight_cap"] is None else s["behav_imp_weight_cap"]]),
                        p + "loss": loss.detach().numpy(), p + "grad": grad.numpy()})
            for name, val in st.items():
                out[p + "stat_" + name] = val.detach().numpy()
            k += 1
    out["n_cases"] = np.array(k)
    # legacy signature (realhf ppo_functional.actor_loss_fn), incl. proximal=None
    j = 0
    for (eps, c_clip, use_prox, cap) in ((0.2, None, False, None), (0.1, 2.0, False, None),
                                         (0.2, 3.0, True, 5.0), (0.5, None, True, None)):
        lp0 = -torch.rand(T, generator=g) * 5
        lp = (lp0 + 0.3 * torch.randn(T, generator=g)).requires_grad_()
        old = lp0 + 0.05 * torch.randn(T, generator=g)
        prox = lp0 + 0.02 * torch.randn(T, generator=g) if use_prox else None
        adv = torch.randn(T, generator=g)
        mask = torch.rand(T, generator=g) > 0.3
        loss, st = pf.actor_loss_fn(lp, old, adv, eps, loss_mask=mask, c_clip=c_clip,
                                    proximal_logprobs=prox, behav_imp_weight_cap=cap)
        (grad,) = torch.autograd.grad(loss, lp)
        p = f"l{j}_"
        out.update({p + "lp": lp.detach().numpy(), p + "old": old.numpy(), p + "adv": adv.numpy(),
                    p + "mask": mask.numpy(), p + "has_prox": np.array(use_prox),
                    p + "prox": (prox if use_prox else torch.zeros(T)).numpy(),
                    p + "cfg": np.array([eps, np.nan if c_clip is None else c_clip,
                                         np.nan if cap is None else cap]),
                    p + "loss": loss.detach().numpy(), p + "grad": grad.numpy()})
        for name, val in st.items():
            out[p + "stat_" + name] = val.detach().numpy()
        j += 1
    out["n_legacy"] = np.array(j)
    np.savez_compressed(os.path.join(OUT, "actor_loss.npz"), **out)
    print("actor_loss cases:", k, "legacy:", j)


def gen_critic_loss(af, pf):
    out = {}
    g = torch.Generator().manual_seed(5)
    T = 513
    k = 0
    for fn_name, fn in (("areal", af.ppo_critic_loss_fn), ("legacy", pf.critic_loss_fn)):
        for loss_type in ("mse", "huber"):
            for use_mask in (True, False):
                scale = 8.0 if loss_type == "huber" else 1.0
                old = torch.randn(T, generator=g) * scale
                val = (old + 0.4 * scale * torch.randn(T, generator=g)).requires_grad_()
                tgt = old + scale * torch.randn(T, generator=g)
                mask = (torch.rand(T, generator=g) > 0.3) if use_mask else None
                loss, st = fn(val, old, tgt, 0.5, loss_mask=mask, loss_fn_type=loss_type)
                (grad,) = torch.autograd.grad(loss, val)
                p = f"c{k}_"
                out.update({p + "value": val.detach().numpy(), p + "old": old.numpy(), p + "target": tgt.numpy(),
                            p + "has_mask": np.array(use_mask),
                            p + "mask": (mask if use_mask else torch.ones(T, dtype=torch.bool)).numpy(),
                            p + "huber": np.array(loss_type == "huber"), p + "eps": np.array(0.5),
                            p + "loss": loss.detach().numpy(), p + "grad": grad.numpy(),
                            p + "stat_clip_mask": st["clip_mask"].numpy(), p + "stat_loss": st["loss"].numpy()})
                k += 1
    out["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "critic_loss.npz"), **out)
    print("critic_loss cases:", k)


def gen_intervals(fp):
    """make_intervals pattern (test_interval_ops.py:20-31) at reduced buffer size + ragged sets."""
    out = {}
    rng = np.random.RandomState(0)
    k = 0
    cases = []
    for (maxsize, n) in ((20000, 1), (20000, 100), (20000, 1000), (20003, 2001), (16384, 500)):
        s = maxsize // n
        iv = np.array([(i * s, i * s + s // 2) for i in range(n)], dtype=np.int64)
        rng.shuffle(iv)
        cases.append((maxsize, iv))
    # ragged, odd-aligned, touching, and empty intervals
    starts = np.sort(rng.choice(19000, 300, replace=False)).astype(np.int64)
    sizes = rng.randint(0, 120, 300).astype(np.int64)
    ends = np.minimum(starts + sizes, np.append(starts[1:], 20000))
    iv = np.stack([starts, ends], 1)
    rng.shuffle(iv)
    cases.append((20000, iv))
    for ci, (maxsize, iv) in enumerate(cases):
        for dt_name, dt in (("bf16", torch.bfloat16), ("fp16", torch.float16), ("fp32", torch.float32)):
            if dt_name != "bf16" and ci not in (3, 5):
                continue
            g = torch.Generator().manual_seed(k)
            src = torch.randn(maxsize, generator=g).to(dt)
            packed = fp._slice_intervals_py(src, [tuple(x) for x in iv.tolist()])
            dst = torch.randn(maxsize, generator=g).to(dt)
            newvals = torch.randn(packed.shape[0], generator=g).to(dt)
            dst_after = dst.clone()
            fp._set_intervals_py(newvals, dst_after, [tuple(x) for x in iv.tolist()])
            p = f"c{k}_"

            def bits(t):
                if t.dtype == torch.float32:
                    return t.numpy().view(np.uint32)
                return t.contiguous().view(torch.int16).numpy().view(np.uint16)
            out.update({p + "dtype": np.array(dt_name), p + "intervals": iv, p + "src": bits(src),
                        p + "packed": bits(packed), p + "dst": bits(dst), p + "newvals": bits(newvals),
                        p + "dst_after": bits(dst_after)})
            k += 1
    out["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "intervals.npz"), **out)
    print("interval cases:", k)


def gen_legacy_masked_norm():
    """Legacy masked_normalization (realhf/impl/model/utils/functional.py:252-319), dim=None."""
    rf = R.load_realhf_functional()
    g = torch.Generator().manual_seed(606)
    z = {}
    k = 0
    for shape in ((1000,), (37, 129), (4, 5, 6)):
        x = torch.randn(*shape, generator=g) * 3 + 0.7
        m = torch.rand(*shape, generator=g) > 0.4
        for mask in (None, m):
            for unbiased in (False, True):
                out = rf.masked_normalization(x, mask, unbiased=unbiased, all_reduce=False)
                z.update({f"c{k}_x": x.numpy(), f"c{k}_mask": (mask if mask is not None else torch.ones_like(m)).numpy(),
                          f"c{k}_cfg": np.array([mask is not None, unbiased]), f"c{k}_out": out.numpy()})
                k += 1
    z["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "legacy_masked_norm.npz"), **z)
    print("legacy masked_normalization golden:", k, "cases")


def gen_legacy_packed():
    """realhf spelling of P2: gather_packed_shifted_log_probs / build_*_indices / calc_entropy."""
    rf = R.load_realhf_functional()
    out = {}
    g = torch.Generator().manual_seed(31)
    k = 0
    for (lens, V, dt) in (([5, 1, 9, 2], 300, torch.float32), ([40, 17, 3, 60], 1031, torch.bfloat16),
                          ([33, 31], 512, torch.float32)):
        T = sum(lens)
        cu = torch.nn.functional.pad(torch.tensor(lens).cumsum(0), (1, 0)).int()
        logits = (torch.randn(T, V, generator=g) * 2).to(dt)
        ids = torch.randint(0, V, (T,), generator=g)
        lp = rf.gather_packed_shifted_log_probs(logits, cu, ids)
        ent = R.run_realhf_calc_entropy(logits, cu)
        p = f"c{k}_"
        out.update({p + "logits": (bf16_bits(logits) if dt == torch.bfloat16 else logits.numpy()), p + "is_bf16": np.array(dt == torch.bfloat16),
                    p + "cu": cu.numpy(), p + "ids": ids.numpy(), p + "logp": lp.float().numpy(),
                    p + "entropy": ent.float().numpy(),
                    p + "leave_one": rf.build_leave_one_indices(logits, cu).numpy(),
                    p + "shift_one": rf.build_shift_one_indices(logits, cu).numpy()})
        k += 1
    out["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "legacy_packed.npz"), **out)
    print("legacy_packed cases:", k)


def gen_normalization(af):
    """`Normalization` (areal/utils/data.py:1073-1304) over the grid its own tests walk
    (areal/tests/test_adv_norm_config.py): mean/std level x leave-one-out x unbiased x mask x group size;
    plus `masked_normalization` (areal/utils/functional.py:130-168)."""
    Norm, NormConfig = R.load_areal_normalization()
    out = {}
    g = torch.Generator().manual_seed(4242)
    k = 0
    for (bs, S, gs) in ((8, 16, 4), (6, 33, 2), (5, 7, 1), (7, 12, 3)):
        for mean_level in ("batch", "group", None):
            for std_level in ("batch", "group", None):
                for loo in (False, True):
                    for unbiased in (False, True):
                        for mask_mode in ("none", "rand", "zero"):
                            if mask_mode == "zero" and not (mean_level == "batch" and std_level == "batch"):
                                continue
                            if (bs, S, gs) != (8, 16, 4) and (k % 3) != 0 and mask_mode == "none":
                                pass
                            cfg = NormConfig(mean_level=mean_level, mean_leave1out=loo, std_level=std_level,
                                             std_unbiased=unbiased, eps=1e-5, group_size=gs)
                            x = torch.randn(bs, S, generator=g) * 3 + 1.5
                            if mask_mode == "none":
                                m = None
                            elif mask_mode == "rand":
                                m = (torch.rand(bs, S, generator=g) > 0.35).float()
                                m[0] = 0.0            # a fully masked row
                            else:
                                m = torch.zeros(bs, S)
                            y = Norm(cfg)(x.clone(), None if m is None else m.clone())
                            p = f"c{k}_"
                            out.update({p + "x": x.numpy(), p + "has_mask": np.array(m is not None),
                                        p + "mask": (m if m is not None else torch.ones(bs, S)).numpy(),
                                        p + "cfg": np.array([{"batch": 0, "group": 1, None: 2}[mean_level],
                                                             {"batch": 0, "group": 1, None: 2}[std_level],
                                                             int(loo), int(unbiased), gs]),
                                        p + "y": y.numpy()})
                            k += 1
    out["n_cases"] = np.array(k)
    j = 0
    for (shape, use_mask, unbiased) in (((300,), True, False), ((7, 41), True, True), ((64,), False, False)):
        x = torch.randn(*shape, generator=g) * 2 - 0.7
        m = (torch.rand(*shape, generator=g) > 0.3) if use_mask else None
        y = af.masked_normalization(x, m, unbiased=unbiased)
        p = f"m{j}_"
        out.update({p + "x": x.numpy(), p + "has_mask": np.array(use_mask),
                    p + "mask": (m if use_mask else torch.ones(*shape, dtype=torch.bool)).numpy(),
                    p + "unbiased": np.array(unbiased), p + "y": y.numpy()})
        j += 1
    out["n_masked_norm"] = np.array(j)
    np.savez_compressed(os.path.join(OUT, "normalization.npz"), **out)
    print("normalization cases:", k, "masked_normalization:", j)


def gen_dynamic_sampling(af):
    """dynamic_sampling (areal/utils/functional.py:314-373): mixed dtypes / trailing dims, groups of 4."""
    g = torch.Generator().manual_seed(12)
    z = {}
    for ci, pattern in enumerate(("mixed", "all_kept", "none_kept")):
        B, S, G = 24, 11, 4
        rewards = torch.randint(0, 2, (B,), generator=g).float()
        if pattern == "all_kept":
            rewards = torch.arange(B).float()
        elif pattern == "none_kept":
            rewards = torch.ones(B)
        else:
            rewards[4:8] = 1.0          # one all-equal group, one all-zero group
            rewards[12:16] = 0.0
        data = dict(rewards=rewards, input_ids=torch.randint(0, 999, (B, S), generator=g),
                    loss_mask=torch.rand(B, S, generator=g) > 0.5, logprobs=torch.randn(B, S, generator=g),
                    hidden=torch.randn(B, S, 3, generator=g).bfloat16(), meta="keep", other=torch.randn(7, generator=g))
        out, stat = af.dynamic_sampling({k: (v.clone() if torch.is_tensor(v) else v) for k, v in data.items()}, G)
        for k in ("rewards", "input_ids", "loss_mask", "logprobs", "other"):
            z[f"c{ci}_in_{k}"] = data[k].numpy()
            z[f"c{ci}_out_{k}"] = out[k].numpy()
        z[f"c{ci}_in_hidden"] = bf16_bits(data["hidden"].reshape(-1)).reshape(B, S, 3)
        z[f"c{ci}_out_hidden"] = bf16_bits(out["hidden"].reshape(-1)).reshape(-1, S, 3)
        z[f"c{ci}_stat"] = np.array([stat["n_group_kept"], stat["n_group_filtered"]])
    np.savez_compressed(os.path.join(OUT, "dynamic_sampling.npz"), **z)
    print("dynamic_sampling golden ok")


def gen_pack():
    """`pack_tensor_dict` (areal/utils/data.py:266-315) on a padded batch with mixed dtypes / trailing dims."""
    pack = R.load_areal_pack_tensor_dict()
    g = torch.Generator().manual_seed(11)
    B, S = 7, 29
    lens = torch.tensor([29, 1, 13, 0, 20, 7, 29])
    am = (torch.arange(S)[None, :] < lens[:, None]).long()
    data = dict(attention_mask=am,
                input_ids=torch.randint(0, 50000, (B, S), generator=g),
                loss_mask=(torch.rand(B, S, generator=g) > 0.5),
                logprobs=torch.randn(B, S, generator=g),
                hidden=torch.randn(B, S, 3, generator=g).bfloat16(),
                rewards=torch.randn(B, generator=g),          # [B] only: left unchanged
                tag="keep")
    out = pack({k: (v.clone() if torch.is_tensor(v) else v) for k, v in data.items()})
    z = {"attention_mask": am.numpy(), "cu_seqlens": out["cu_seqlens"].numpy(), "max_seqlen": np.array(out["max_seqlen"])}
    for k in ("input_ids", "loss_mask", "logprobs", "rewards"):
        z["in_" + k] = data[k].numpy()
        z["out_" + k] = out[k].numpy()
    z["in_hidden"] = bf16_bits(data["hidden"].reshape(-1)).reshape(B, S, 3)
    z["out_hidden"] = bf16_bits(out["hidden"].reshape(-1)).reshape(-1, 3)
    np.savez_compressed(os.path.join(OUT, "pack.npz"), **z)
    print("pack golden: total", int(out["cu_seqlens"][-1]))


def adv_case_grid():
    """Configurations of PPOActor.compute_advantages covered by the golden file."""
    base = dict(reward_bias=0.0, reward_scaling=1.0, reward_clip=20.0, kl_ctl=0.1, kl_estimator="k1",
                discount=1.0, gae_lambda=1.0, mask_no_eos_with_zero=False, use_decoupled_loss=True,
                recompute_logprob=True, overlong_reward_penalty=False, overlong_tokens=8, overlong_penalty_factor=1.0,
                max_new_tokens=24, with_values=False, with_ref=True, adv_norm=False, reward_norm=False)
    grid = [dict(),
            dict(kl_estimator="k2", kl_ctl=0.5, discount=0.99, gae_lambda=0.95, with_values=True),
            dict(kl_estimator="k3", kl_ctl=0.05, reward_bias=-0.5, reward_scaling=3.0, reward_clip=2.5,
                 mask_no_eos_with_zero=True, with_values=True, discount=0.9, gae_lambda=0.8),
            dict(with_ref=False, kl_ctl=0.0, adv_norm=True),
            dict(use_decoupled_loss=False, recompute_logprob=True, with_values=True, discount=0.97, gae_lambda=0.9),
            dict(use_decoupled_loss=False, recompute_logprob=False, kl_estimator="k2"),
            dict(overlong_reward_penalty=True, overlong_tokens=6, overlong_penalty_factor=0.75, max_new_tokens=20,
                 reward_scaling=2.0, with_values=True, discount=0.95, gae_lambda=0.9),
            dict(reward_norm=True, adv_norm=True, kl_estimator="k3", with_values=True, gae_lambda=0.7)]
    return [dict(base, **g) for g in grid]


def adv_case_inputs(ci, B=12, S=37):
    g = torch.Generator().manual_seed(400 + ci)
    lens = torch.randint(2, S + 1, (B,), generator=g)
    lens[0], lens[1], lens[2] = S, 1, 2                      # no-eos row, shortest rows
    if ci % 2:
        lens[3] = S
    am = torch.arange(S)[None, :] < lens[:, None]
    prompt = torch.minimum(torch.randint(0, 9, (B,), generator=g), lens)
    lm = am & (torch.arange(S)[None, :] >= prompt[:, None]) & (torch.rand(B, S, generator=g) > 0.15)
    return dict(input_ids=torch.randint(0, 1000, (B, S), generator=g), attention_mask=am, loss_mask=lm.long(),
                logprobs=-torch.rand(B, S, generator=g) * 4, prox_logp=-torch.rand(B, S, generator=g) * 4,
                ref_logp=-torch.rand(B, S, generator=g) * 4, values=torch.randn(B, S, generator=g),
                rewards=torch.randn(B, generator=g) * 3)


def gen_advantages():
    """PPOActor.compute_advantages (areal/engine/ppo/actor.py:72-164) run unmodified on CPU over the grid."""
    import types
    fn, KLEstimator, Normalization, NormConfig = R.load_areal_compute_advantages()
    z = {}
    for ci, c in enumerate(adv_case_grid()):
        d = adv_case_inputs(ci)
        if not c["with_values"]:
            d.pop("values")
        if not c["with_ref"]:
            d.pop("ref_logp")
        for k, v in d.items():
            z[f"c{ci}_in_{k}"] = v.numpy().copy()
        cfg = types.SimpleNamespace(**{k: c[k] for k in ("use_decoupled_loss", "recompute_logprob", "overlong_reward_penalty",
                                                          "overlong_tokens", "overlong_penalty_factor", "max_new_tokens")})
        actor = types.SimpleNamespace(
            config=cfg, reward_bias=c["reward_bias"], reward_scaling=c["reward_scaling"], reward_clip=c["reward_clip"],
            kl_ctl=c["kl_ctl"], kl_estimator=KLEstimator(c["kl_estimator"]), discount=c["discount"],
            gae_lambda=c["gae_lambda"], mask_no_eos_with_zero=c["mask_no_eos_with_zero"],
            adv_norm=Normalization(NormConfig(mean_level="batch", std_level="batch")) if c["adv_norm"] else None,
            reward_norm=Normalization(NormConfig(mean_level="group", std_level="group", group_size=4)) if c["reward_norm"] else None)
        with torch.no_grad():
            fn(actor, d)
        for k in ("advantages", "returns", "kl_rewards", "tot_rewards", "loss_mask", "logprobs", "rewards", "ref_logp", "prox_logp"):
            if k in d:
                z[f"c{ci}_out_{k}"] = d[k].numpy().copy()
    z["n_cases"] = np.array(len(adv_case_grid()))
    np.savez_compressed(os.path.join(OUT, "advantages.npz"), **z)
    print("advantages golden:", len(adv_case_grid()), "cases")


def grpo_case(mb, T=300, V=977):
    g = torch.Generator().manual_seed(900 + mb)
    logits = (torch.randn(T, V, generator=g) * 2.5).bfloat16()
    ids = torch.randint(0, V, (T,), generator=g)
    old = -torch.rand(T, generator=g) * 3
    prox = old + torch.randn(T, generator=g) * 0.3
    adv = torch.randn(T, generator=g)
    mask = torch.rand(T, generator=g) > (0.3 if mb < 2 else 2.0)        # third micro-batch: nothing valid
    return dict(logits=logits, input_ids=ids, logprobs=old, prox_logp=prox, advantages=adv, loss_mask=mask)


GRPO_ARGS = [dict(temperature=1.0, eps_clip=0.2, eps_clip_higher=None, c_clip=None, behav_imp_weight_cap=None),
             dict(temperature=0.7, eps_clip=0.1, eps_clip_higher=0.3, c_clip=3.0, behav_imp_weight_cap=2.0)]


def gen_grpo_stats():
    """grpo_loss_fn + DistributedStatsTracker.export (actor.py:313-391, stats_tracker.py) run unmodified over
    three micro-batches (the last with an all-false loss mask) inside a scope; fp32 logits on CPU."""
    fn, st = R.load_areal_grpo_loss_fn()
    z = {}
    for ai, a in enumerate(GRPO_ARGS):
        tr = st.DistributedStatsTracker()
        st.denominator, st.stat = tr.denominator, tr.stat               # what grpo_loss_fn calls
        losses = []
        with tr.scope("ppo_actor"):
            for mb in range(3):
                c = grpo_case(mb)
                if ai == 0:
                    for k, v in c.items():
                        z[f"mb{mb}_{k}"] = bf16_bits(v) if v.dtype == torch.bfloat16 else v.numpy()
                d = {k: v for k, v in c.items() if k != "logits"}
                losses.append(float(fn(c["logits"].float(), d, **a)))
            tr.scalar(lr=1e-3)
            tr.scalar(lr=3e-3)
        out = tr.export()
        z[f"a{ai}_losses"] = np.array(losses)
        z[f"a{ai}_keys"] = np.array(sorted(out.keys()))
        z[f"a{ai}_vals"] = np.array([out[k] for k in sorted(out.keys())], dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "grpo_stats.npz"), **z)
    print("grpo_stats golden:", len(out), "keys; e.g.", list(out.items())[:3])


def gen_packed_rewards(pf):
    """get_packed_rewards / get_packed_reward_dense (realhf ppo_functional.py:228-289) run unmodified."""
    g = torch.Generator().manual_seed(321)
    z = {}
    k = 0
    for lens in ([5, 1, 17, 2, 300], [1], [64] * 9, list(range(1, 40))):
        cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32)
        B, Ts = len(lens), int(cu[-1])
        lp, rf = -torch.rand(Ts, generator=g) * 5, -torch.rand(Ts, generator=g) * 5
        score = torch.randn(B, generator=g) * 6
        dense = torch.randn(Ts + B, generator=g) * 6
        ne = torch.rand(B, generator=g) > 0.5
        for kl_ctl, clip, mz in ((0.1, 5.0, False), (0.02, 2.0, True)):
            a, b = pf.get_packed_rewards(kl_ctl, clip, lp, rf, score, cu, ne, mz)
            for delta in (False, True):
                c, d = pf.get_packed_reward_dense(kl_ctl, clip, lp, rf, dense, cu, ne.float(), delta)
                z.update({f"c{k}_cu": cu.numpy(), f"c{k}_lp": lp.numpy(), f"c{k}_rf": rf.numpy(), f"c{k}_score": score.numpy(),
                          f"c{k}_dense": dense.numpy(), f"c{k}_ne": ne.numpy(), f"c{k}_cfg": np.array([kl_ctl, clip, mz, delta]),
                          f"c{k}_kl": a.numpy(), f"c{k}_tot": b.numpy(), f"c{k}_dkl": c.numpy(), f"c{k}_dtot": d.numpy()})
                k += 1
    z["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "packed_rewards.npz"), **z)
    print("packed_rewards golden:", k, "cases")


def gen_rms():
    """ExponentialRunningMeanStd / MovingAverageRunningMeanStd (realhf/impl/model/modules/rms.py:19-183):
    three updates (unmasked, masked, masked 2-D), then mean_std / normalize / denormalize."""
    rms = R.load_realhf_rms()
    g = torch.Generator().manual_seed(808)
    xs = [torch.randn(5000, generator=g) * 3 + 1.5, torch.randn(777, generator=g) * 0.2 - 4.0,
          torch.randn(64, 100, generator=g) * 10]
    ms = [None, torch.rand(777, generator=g) > 0.4, torch.rand(64, 100, generator=g) > 0.7]
    probe = torch.randn(3000, generator=g) * 8
    z = {"probe": probe.numpy()}
    for i, (x, m) in enumerate(zip(xs, ms)):
        z[f"x{i}"] = x.numpy()
        if m is not None:
            z[f"m{i}"] = m.numpy()
    k = 0
    for cls in ("ExponentialRunningMeanStd", "MovingAverageRunningMeanStd"):
        for hp in (True, False):
            mod = getattr(rms, cls)(beta=0.9, epsilon=1e-5, high_precision=hp) if cls.startswith("Exp") else getattr(rms, cls)(high_precision=hp)
            for step, (x, m) in enumerate(zip(xs, ms)):
                mod.update(x, m)
                mean, std = mod.mean_std()
                z[f"c{k}_mean{step}"] = mean.double().numpy()
                z[f"c{k}_std{step}"] = std.double().numpy()
                z[f"c{k}_norm{step}"] = mod.normalize(probe).numpy()
                z[f"c{k}_denorm{step}"] = mod.denormalize(probe).numpy()
            z[f"c{k}_cfg"] = np.array([cls.startswith("Exp"), hp])
            k += 1
    z["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "rms.npz"), **z)
    print("rms golden:", k, "cases")


def gen_clamped_entropy():
    """gather_logprobs_clamped_entropy (recipe/AEnt/functional.py:39-52) and its autograd gradient, run
    unmodified on CPU.  bf16 logits (few distinct values: ties at the threshold are the rule)."""
    _, gather = R.load_aent_functional()
    g = torch.Generator().manual_seed(4711)
    z = {}
    k = 0
    for (T, V), clamp, temp in (((12, 1000), 0.1, 1.0), ((12, 1000), 0.5, 0.7), ((6, 4096), 0.9, 1.0),
                                ((3, 151936), 0.3, 1.0), ((5, 777), 0.0, 1.3), ((4, 64), 0.25, 1.0)):
        logits = (torch.randn(T, V, generator=g) * 3).bfloat16()
        labels = torch.randint(0, V, (T,), generator=g)
        w_lp, w_h = torch.randn(T, generator=g), torch.randn(T, generator=g)
        x = logits.float().requires_grad_()
        lp, ent = gather(x, labels, clamp, temp)
        ((lp * w_lp).sum() + (ent * w_h).sum()).backward()
        z.update({f"c{k}_logits_bits": bf16_bits(logits.reshape(-1)).reshape(T, V), f"c{k}_labels": labels.numpy(),
                  f"c{k}_cfg": np.array([clamp, temp]), f"c{k}_w_lp": w_lp.numpy(), f"c{k}_w_h": w_h.numpy(),
                  f"c{k}_logp": lp.detach().numpy(), f"c{k}_ent": ent.detach().numpy(), f"c{k}_grad": x.grad.numpy()})
        k += 1
    z["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "clamped_entropy.npz"), **z)
    print("clamped_entropy golden:", k, "cases")


def gen_aent_stats():
    """aent_grpo_loss_fn + tracker export (recipe/AEnt/actor.py:185-270) run unmodified over the micro-batches of
    gen_grpo_stats (inputs shared with grpo_stats.npz), entropy_clamp 0.25 and 0."""
    fn, st = R.load_aent_grpo_loss_fn()
    z = {}
    for ai, (coeff, clamp) in enumerate(((0.01, 0.25), (0.002, 0.0))):
        tr = st.DistributedStatsTracker()
        st.denominator, st.stat = tr.denominator, tr.stat
        losses = []
        with tr.scope("aent"):
            for mb in range(3):
                c = grpo_case(mb)
                d = {k: v for k, v in c.items() if k != "logits"}
                losses.append(float(fn(c["logits"].float(), d, entropy_coeff=coeff, entropy_clamp=clamp, **GRPO_ARGS[1])))
        out = tr.export()
        z[f"a{ai}_cfg"] = np.array([coeff, clamp])
        z[f"a{ai}_losses"] = np.array(losses)
        z[f"a{ai}_keys"] = np.array(sorted(out.keys()))
        z[f"a{ai}_vals"] = np.array([out[k] for k in sorted(out.keys())], dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "aent_stats.npz"), **z)
    print("aent_stats golden:", len(out), "keys, losses", losses)


def critic_case(mb, T=700):
    g = torch.Generator().manual_seed(1300 + mb)
    old = torch.randn(T, generator=g)
    return dict(value=(old + 0.3 * torch.randn(T, generator=g)).unsqueeze(-1), values=old,
                returns=old + torch.randn(T, generator=g), loss_mask=torch.rand(T, generator=g) > (0.4 if mb < 2 else 2.0))


def gen_critic_stats():
    """ppo_loss_fn of the critic + tracker export (areal/engine/ppo/critic.py:81-127), three micro-batches."""
    fn, st = R.load_areal_critic_loss_fn()
    z = {}
    tr = st.DistributedStatsTracker()
    st.denominator, st.stat = tr.denominator, tr.stat
    losses = []
    with tr.scope("ppo_critic"):
        for mb in range(3):
            c = critic_case(mb)
            for k, v in c.items():
                z[f"mb{mb}_{k}"] = v.numpy()
            losses.append(float(fn(c["value"], {k: v for k, v in c.items() if k != "value"}, 0.2)))
    out = tr.export()
    z["losses"] = np.array(losses)
    z["keys"] = np.array(sorted(out.keys()))
    z["vals"] = np.array([out[k] for k in sorted(out.keys())], dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "critic_stats.npz"), **z)
    print("critic_stats golden:", len(out), "keys")


def sft_case(mb, V=977):
    g = torch.Generator().manual_seed(1700 + mb)
    lens = torch.randint(3, 60, (9,), generator=g)
    cu = torch.zeros(10, dtype=torch.int32)
    cu[1:] = lens.cumsum(0)
    T = int(cu[-1])
    mask = torch.rand(T, generator=g) > 0.5
    mask[int(cu[3]):int(cu[4]) + 1] = False                      # one sequence without any target token
    return dict(logits=(torch.randn(T, V, generator=g) * 2.5).bfloat16(), input_ids=torch.randint(0, V, (T,), generator=g),
                cu_seqlens=cu, loss_mask=mask)


def gen_sft_stats():
    """compute_packed_sft_loss + tracker export (areal/engine/sft/lm_engine.py:58-110), two micro-batches."""
    fn, st = R.load_areal_sft_loss_fn()
    z = {}
    tr = st.DistributedStatsTracker()
    st.denominator, st.stat = tr.denominator, tr.stat
    losses = []
    with tr.scope("sft"):
        for mb in range(2):
            c = sft_case(mb)
            for k, v in c.items():
                z[f"mb{mb}_{k}"] = bf16_bits(v) if v.dtype == torch.bfloat16 else v.numpy()
            losses.append(float(fn(c["logits"].float(), {k: v for k, v in c.items() if k != "logits"})))
    out = tr.export()
    z["losses"] = np.array(losses)
    z["keys"] = np.array(sorted(out.keys()))
    z["vals"] = np.array([out[k] for k in sorted(out.keys())], dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "sft_stats.npz"), **z)
    print("sft_stats golden:", len(out), "keys", losses)


def gen_lm_head(af):
    """The step BEFORE the path: logits = hidden @ lm_head.weight.T as the engine materialises them
    (areal/engine/fsdp_engine.py:571-573), then the unmodified gather_logprobs_entropy + ppo_actor_loss_fn and
    `(loss * loss_scale).backward()` down to the hidden state and the weight.  Hidden states / weights sit on
    coarse binary grids so that every dot product is exact in fp32: the bf16 logits are the same bits on the CPU
    here and in any GEMM on the GPU."""
    z = {}
    k = 0
    for (T, Hd, V, temp, scale, cfg) in ((300, 64, 4104, 0.9, 0.37, dict(eps_clip=0.4, c_clip=3.0, behav_imp_weight_cap=5.0)),
                                         (130, 64, 8200, 1.0, 1.0, dict(eps_clip=0.2, eps_clip_higher=0.28))):
        g = torch.Generator().manual_seed(4200 + k)
        h = (torch.randint(-16, 17, (T, Hd), generator=g).float() / 8).bfloat16().requires_grad_()
        w = (torch.randint(-16, 17, (V, Hd), generator=g).float() / 32).bfloat16().requires_grad_()
        labels = torch.randint(0, V, (T,), generator=g)
        logits = torch.nn.functional.linear(h, w)
        exact = (h.detach().double() @ w.detach().double().t()).float().bfloat16()
        assert torch.equal(logits.detach(), exact), "CPU bf16 linear is not the exactly rounded product"
        lp, ent = af.gather_logprobs_entropy(logits, labels, temp)
        prox = (lp.detach() + 0.02 * torch.randn(T, generator=g)).float()
        old = (lp.detach() + 0.05 * torch.randn(T, generator=g)).float()
        adv = torch.randn(T, generator=g)
        mask = torch.rand(T, generator=g) > 0.3
        loss, stat = af.ppo_actor_loss_fn(logprobs=lp, proximal_logprobs=prox, old_logprobs=old, advantages=adv,
                                          loss_mask=mask, **cfg)
        (loss * scale).backward()
        pfx = f"c{k}_"
        z.update({pfx + "h_bits": bf16_bits(h.detach()), pfx + "w_bits": bf16_bits(w.detach()), pfx + "labels": labels.numpy(),
                  pfx + "temp": np.array(temp), pfx + "scale": np.array(scale), pfx + "prox": prox.numpy(),
                  pfx + "old": old.numpy(), pfx + "adv": adv.numpy(), pfx + "mask": mask.numpy(),
                  pfx + "loss": np.array(float(loss.detach())), pfx + "logp": lp.detach().numpy(),
                  pfx + "entropy": ent.detach().numpy(), pfx + "dh_bits": bf16_bits(h.grad), pfx + "dw_bits": bf16_bits(w.grad),
                  pfx + "importance_weight": stat["importance_weight"].numpy(), pfx + "clip_mask": stat["clip_mask"].numpy()})
        for name, val in cfg.items():
            z[pfx + "cfg_" + name] = np.array(val)
        k += 1
    z["n_cases"] = np.array(k)
    np.savez_compressed(os.path.join(OUT, "lm_head.npz"), **z)
    print("lm_head golden cases:", k)


def gen_nan_cases(af):
    """NaN inputs: torch.clamp / torch.max / torch.min hand a NaN on (CUDA's fminf / fmaxf would not), so a
    NaN log-prob, advantage, value or score must come out as NaN, not as a clamp bound.  Unmodified reference
    functions on the CPU: ppo_actor_loss_fn, ppo_critic_loss_fn, PPOActor.compute_advantages."""
    import types
    z = {}
    g = torch.Generator().manual_seed(77)
    T = 96
    nan = float("nan")
    k = 0
    for s in (dict(eps_clip=0.2, eps_clip_higher=None, c_clip=None, behav_imp_weight_cap=None),
              dict(eps_clip=0.2, eps_clip_higher=0.28, c_clip=3.0, behav_imp_weight_cap=5.0)):
        for where in ("lp", "prox", "old", "adv", "masked_lp"):
            lp0 = -torch.rand(T, generator=g) * 5
            lp = lp0 + 0.3 * torch.randn(T, generator=g)
            prox = lp0 + 0.02 * torch.randn(T, generator=g)
            old = lp0 + 0.05 * torch.randn(T, generator=g)
            adv = torch.randn(T, generator=g)
            mask = torch.rand(T, generator=g) > 0.3
            mask[[3, 40, 77]] = where != "masked_lp"
            {"lp": lp, "prox": prox, "old": old, "adv": adv, "masked_lp": lp}[where][[3, 40, 77]] = nan
            lp.requires_grad_()
            loss, st = af.ppo_actor_loss_fn(lp, prox, old, adv, loss_mask=mask, **s)
            (grad,) = torch.autograd.grad(loss, lp)
            p = f"a{k}_"
            z.update({p + "lp": lp.detach().numpy(), p + "prox": prox.numpy(), p + "old": old.numpy(),
                      p + "adv": adv.numpy(), p + "mask": mask.numpy(),
                      p + "cfg": np.array([s["eps_clip"],
                                           np.nan if s["eps_clip_higher"] is None else s["eps_clip_higher"],
                                           np.nan if s["c_clip"] is None else s["c_clip"],
                                           np.nan if s["behav_imp_weight_cap"] is None else s["behav_imp_weight_cap"]]),
                      p + "loss": loss.detach().numpy(), p + "grad": grad.numpy()})
            for name, val in st.items():
                z[p + "stat_" + name] = val.detach().numpy()
            k += 1
    z["n_actor"] = np.array(k)
    k = 0
    for loss_type in ("mse", "huber"):
        for where in ("value", "old", "target"):
            old = torch.randn(T, generator=g)
            val = old + 0.4 * torch.randn(T, generator=g)
            tgt = old + torch.randn(T, generator=g)
            mask = torch.rand(T, generator=g) > 0.3
            mask[[5, 50]] = True
            {"value": val, "old": old, "target": tgt}[where][[5, 50]] = nan
            val.requires_grad_()
            loss, st = af.ppo_critic_loss_fn(val, old, tgt, 0.5, loss_mask=mask, loss_fn_type=loss_type)
            (grad,) = torch.autograd.grad(loss, val)
            p = f"v{k}_"
            z.update({p + "value": val.detach().numpy(), p + "old": old.numpy(), p + "target": tgt.numpy(),
                      p + "mask": mask.numpy(), p + "huber": np.array(loss_type == "huber"),
                      p + "loss": loss.detach().numpy(), p + "grad": grad.numpy(),
                      p + "stat_clip_mask": st["clip_mask"].numpy(), p + "stat_loss": st["loss"].numpy()})
            k += 1
    z["n_critic"] = np.array(k)
    # compute_advantages: a NaN behaviour / reference log-prob through the clamped KL estimator, a NaN score
    # through the reward clip (grid cases 0: k1, 1: k2, 2: k3 with scaling + clip)
    fn, KLEstimator, Normalization, NormConfig = R.load_areal_compute_advantages()
    for ci in (0, 1, 2):
        c = adv_case_grid()[ci]
        d = adv_case_inputs(50 + ci)
        d["logprobs"][4, 6] = nan
        d["ref_logp"][6, 3] = nan
        d["rewards"][8] = nan
        d["loss_mask"][4, 5:8] = 1
        d["loss_mask"][6, 2:5] = 1
        d["attention_mask"][4, :9] = True
        d["attention_mask"][6, :6] = True
        # every planted NaN sits on a token that takes part in the loss (rolled mask 1), the score of row 8
        # included (it lands on column seqlen - 2): the reference masks by multiplying with 0, which would
        # spread a NaN from a MASKED token too (NaN * 0), an artefact the kernels do not reproduce
        d["loss_mask"][8, int(d["attention_mask"][8].sum()) - 1] = 1
        if not c["with_values"]:
            d.pop("values")
        for kk, v in d.items():
            z[f"c{ci}_in_{kk}"] = v.numpy().copy()
        cfg = types.SimpleNamespace(**{kk: c[kk] for kk in ("use_decoupled_loss", "recompute_logprob", "overlong_reward_penalty",
                                                            "overlong_tokens", "overlong_penalty_factor", "max_new_tokens")})
        actor = types.SimpleNamespace(
            config=cfg, reward_bias=c["reward_bias"], reward_scaling=c["reward_scaling"], reward_clip=c["reward_clip"],
            kl_ctl=c["kl_ctl"], kl_estimator=KLEstimator(c["kl_estimator"]), discount=c["discount"],
            gae_lambda=c["gae_lambda"], mask_no_eos_with_zero=c["mask_no_eos_with_zero"], adv_norm=None, reward_norm=None)
        with torch.no_grad():
            fn(actor, d)
        for kk in ("advantages", "returns", "kl_rewards", "tot_rewards", "loss_mask", "logprobs", "ref_logp"):
            if kk in d:
                z[f"c{ci}_out_{kk}"] = d[kk].numpy().copy()
    z["n_adv"] = np.array(3)
    np.savez_compressed(os.path.join(OUT, "nan_cases.npz"), **z)
    print("nan cases: actor", int(z["n_actor"]), "critic", int(z["n_critic"]), "advantages 3")


def main():
    os.makedirs(OUT, exist_ok=True)
    assert R.available(), "needs /root/reference"
    torch.set_num_threads(8)
    pf = R.load_realhf_ppo_functional()
    af = R.load_areal_functional()
    fp = R.load_realhf_flatten_param()
    if len(sys.argv) > 2 and sys.argv[1] == "--only":       # regenerate one fixture: --only nan_cases
        fn = globals()["gen_" + sys.argv[2]]
        args = {"nan_cases": (af,), "gae1d": (pf,), "logprobs": (af,), "actor_loss": (af, pf), "critic_loss": (af, pf),
                "intervals": (fp,), "normalization": (af,), "dynamic_sampling": (af,), "packed_rewards": (pf,),
                "lm_head": (af,)}.get(sys.argv[2], ())
        fn(*args)
        return
    gen_gae1d(pf)
    gen_gae_padded()
    gen_logprobs(af)
    gen_actor_loss(af, pf)
    gen_critic_loss(af, pf)
    gen_intervals(fp)
    gen_legacy_packed()
    gen_legacy_masked_norm()
    gen_normalization(af)
    gen_pack()
    gen_dynamic_sampling(af)
    gen_advantages()
    gen_grpo_stats()
    gen_packed_rewards(pf)
    gen_rms()
    gen_clamped_entropy()
    gen_aent_stats()
    gen_critic_stats()
    gen_sft_stats()
    gen_lm_head(af)
    gen_nan_cases(af)


if __name__ == "__main__":
    main()
