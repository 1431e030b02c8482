#!/usr/bin/env python
"""Per-kernel measurements for DESIGN.md / profiles/ (NOT the driver's bench contract -- that is
bench.py).  Times every kernel family alone with CUDA events after warm-up, inputs larger than L2 or L2
flushed between iterations, and -- where present -- the reference's own CUDA kernels rebuilt for
sm_100a (oracle/_ref/*.so, timing baseline only) and the reference-style torch path.

    python bench_kernels.py [--what gae,padded,advantages,intervals,logits,loss,ref] [--out file.json]
"""
import argparse
import glob
import importlib.util
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from areal_b200 import flatten_param as fp  # noqa: E402
from areal_b200 import functional as F  # noqa: E402
from areal_b200 import ppo_functional as P  # noqa: E402

DEV = torch.device("cuda:0")
PEAK = 6539.9
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
_flush = None


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    _flush.zero_()


def timeit(fn, iters=20, warmup=5, flush=False):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush:
            flush_l2()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def load_ref(name):
    so = glob.glob(os.path.join(ROOT, "oracle", "_ref", name + "*.so"))
    if not so:
        return None
    spec = importlib.util.spec_from_file_location(name, so[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def row(name, ms, bytes_, **kw):
    gbs = bytes_ / (ms * 1e-3) / 1e9
    d = dict(kernel=name, ms=round(ms, 5), algorithmic_bytes=int(bytes_), gbs=round(gbs, 1),
             frac_of_measured_peak=round(gbs / PEAK, 4), **kw)
    print(json.dumps(d), flush=True)
    return d


def make_packed(T, rng, lo=128, hi=4097):
    lens, left = [], T
    while left > 0:
        n = min(left, int(rng.randint(lo, hi)))
        lens.append(n)
        left -= n
    return np.array(lens)


def bench_gae(out, ref):
    rng = np.random.RandomState(0)
    for logT in (15, 17, 20, 22, 24, 26, 27, 28):
        T = 1 << logT
        lens = make_packed(T, rng)
        B = len(lens)
        cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(DEV)
        r = torch.randn(T, device=DEV)
        v = torch.randn(T + B, device=DEV)
        boot = torch.rand(B, device=DEV) > 0.5
        small = (16 * T) < (200 << 20)
        ms = timeit(lambda: P.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 0.95), flush=small)
        out.append(row("gae1d", ms, 16 * T + 9 * B + 4, T=T, B=B, tokens_per_s=T / (ms * 1e-3), l2_flushed=small))
        if not small:
            # device time per call: four calls enqueued back to back, so the ~13 us the Python wrapper spends
            # before its first launch (the GPU is idle when a lone call starts) overlaps the previous call
            def four():
                for _ in range(4):
                    P.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 0.95)
            ms4 = timeit(four) / 4
            out.append(row("gae1d (4 calls back to back, per call)", ms4, 16 * T + 9 * B + 4, T=T, B=B,
                           tokens_per_s=T / (ms4 * 1e-3), l2_flushed=False))
        if ref is not None and logT <= 24:
            msr = timeit(lambda: ref.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 0.95), iters=5, warmup=2, flush=small)
            out.append(row("reference gae_1d_nolp_misalign (sm_100a rebuild)", msr, 16 * T + 9 * B + 4, T=T, B=B,
                           tokens_per_s=T / (msr * 1e-3), speedup_ours=msr / ms))
        del r, v


def bench_padded(out):
    rng = np.random.RandomState(1)
    for B, S in ((16, 32768), (512, 4096), (4096, 4096)):
        r = torch.randn(B, S, device=DEV)
        v = torch.randn(B, S, device=DEV)
        m = (torch.rand(B, S, device=DEV) > 0.2).float()
        ne = torch.rand(B, device=DEV) > 0.5
        small = 20 * B * S < (200 << 20)
        ms = timeit(lambda: P.gae_padded_masked(r, v, m, ne, 1.0, 0.95), flush=small)
        out.append(row("gae_padded_masked", ms, 20 * B * S, B=B, S=S, tokens_per_s=B * S / (ms * 1e-3)))
        if B * S <= 16 * 32768:
            # the reference: PPOActor.compute_advantages' python loop, restated with torch ops on the GPU
            def ref_loop():
                last = torch.zeros(B, device=DEV)
                nv = v[:, S - 1] * ne
                adv_rev = [last]
                for t in reversed(range(S - 1)):
                    delta = r[:, t] + 1.0 * nv - v[:, t]
                    new = delta + 0.95 * last
                    mk = m[:, t]
                    nv = nv * (1 - mk) + v[:, t] * mk
                    last = last * (1 - mk) + new * mk
                    adv_rev.append(last)
                return torch.stack(adv_rev[::-1], dim=1)
            msr = timeit(ref_loop, iters=1, warmup=0)
            out.append(row("reference padded GAE python loop (torch CUDA ops)", msr, 20 * B * S, B=B, S=S,
                           tokens_per_s=B * S / (msr * 1e-3), speedup_ours=msr / ms))


def bench_intervals(out, ref):
    rng = np.random.RandomState(0)
    n_elem = int(os.environ.get("INTERVAL_BUFFER_ELEMS", 7_600_000_000))   # cfg 5: 7B-sized flat bf16 buffer
    src = torch.empty(n_elem, dtype=torch.bfloat16, device=DEV)
    src.view(torch.int16).random_(-30000, 30000)
    for n in (1000, 4000, 10000, 100000):
        s = n_elem // n
        iv = np.stack([np.arange(n, dtype=np.int64) * s, np.arange(n, dtype=np.int64) * s + s // 2], 1)
        rng.shuffle(iv)
        total = n * (s // 2)
        ivc = torch.from_numpy(iv).to(DEV)
        ms = timeit(lambda: fp.slice_intervals_bf16(src, ivc, total, s // 2), iters=8, warmup=2)
        out.append(row("slice_intervals", ms, 4 * total + 16 * n, n_intervals=n, interval_elems=s // 2))
        packed = fp.slice_intervals_bf16(src, ivc, total, s // 2)
        ms = timeit(lambda: fp.set_intervals_bf16(packed, src, ivc, s // 2), iters=8, warmup=2)
        out.append(row("set_intervals", ms, 4 * total + 16 * n, n_intervals=n, interval_elems=s // 2))
        if ref is not None and n >= 1000:
            msr = timeit(lambda: ref.slice_intervals_bf16(src, ivc, total, s // 2), iters=3, warmup=1)
            out.append(row("reference slice_intervals_bf16 (sm_100a rebuild)", msr, 4 * total + 16 * n,
                           n_intervals=n, speedup_ours=msr / ms))
        del packed
    # realistic: TP 1->2 reshard of a Qwen2.5-7B-shaped model: per layer 2 x 3584 row-split intervals
    rows = []
    base = 0
    for layer in range(28):
        for width in (3584, 18944):          # o_proj [3584,3584], down_proj [3584,18944] split on dim 1
            for rr in range(3584):
                rows.append((base + rr * width + width // 2, base + (rr + 1) * width))
            base += 3584 * width
    iv = np.array(rows, dtype=np.int64)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    ivc = torch.from_numpy(iv).to(DEV)
    ms = timeit(lambda: fp.slice_intervals_bf16(src, ivc, total, 9472), iters=8, warmup=2)
    out.append(row("slice_intervals (7B TP reshard rows)", ms, 4 * total + 16 * len(rows), n_intervals=len(rows)))
    if ref is not None:
        msr = timeit(lambda: ref.slice_intervals_bf16(src, ivc, total, 9472), iters=3, warmup=1)
        out.append(row("reference slice_intervals_bf16 (7B TP reshard rows)", msr, 4 * total + 16 * len(rows),
                       n_intervals=len(rows), speedup_ours=msr / ms))
    # misaligned (odd element offsets): 2-byte relative alignment
    iv = np.stack([np.arange(100000, dtype=np.int64) * 20000 + 1, np.arange(100000, dtype=np.int64) * 20000 + 9000], 1)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    ivc = torch.from_numpy(iv).to(DEV)
    ms = timeit(lambda: fp.slice_intervals_bf16(src, ivc, total, 9000), iters=8, warmup=2)
    out.append(row("slice_intervals (odd offsets, 2-byte relative alignment)", ms, 4 * total + 16 * 100000,
                   n_intervals=100000))
    del src


def bench_logits(out):
    T, V = 16384, 151936
    g = torch.Generator(device=DEV).manual_seed(0)
    logits = torch.empty(T, V, dtype=torch.bfloat16, device=DEV)
    for i in range(0, T, 2048):
        logits[i:i + 2048] = (torch.randn(2048, V, generator=g, device=DEV) * 2).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g, device=DEV)
    ms = timeit(lambda: F.gather_logprobs(logits, labels, 1.0))
    out.append(row("logits_fwd (logp only)", ms, 2 * T * V, tokens_per_s=T / (ms * 1e-3)))
    ms = timeit(lambda: F.gather_logprobs_entropy(logits, labels, 1.0))
    out.append(row("logits_fwd (logp+entropy)", ms, 2 * T * V, tokens_per_s=T / (ms * 1e-3)))
    ms = timeit(lambda: F.logits_stats(logits, labels, 1.0))
    out.append(row("logits_fwd (logp+entropy+min/max)", ms, 2 * T * V, tokens_per_s=T / (ms * 1e-3)))
    x = logits.requires_grad_()
    lp, ent = F.gather_logprobs_entropy(x, labels, 1.0)
    w = torch.randn(T, device=DEV)
    for frac in (1.0, 0.8, 0.5):
        ww = w.clone()
        ww[torch.rand(T, device=DEV) > frac] = 0
        act = float((ww != 0).float().mean())
        ms = timeit(lambda: torch.autograd.grad((lp * ww).sum(), x, retain_graph=True), iters=10, warmup=3)
        out.append(row("logits_bwd", ms, 2 * T * V * (1 + act), active_rows=act, tokens_per_s=T / (ms * 1e-3)))
    wh = torch.randn(T, device=DEV)
    ms = timeit(lambda: torch.autograd.grad((lp * w).sum() + (ent * wh).sum(), x, retain_graph=True), iters=10, warmup=3)
    out.append(row("logits_bwd (with entropy gradient)", ms, 4 * T * V, tokens_per_s=T / (ms * 1e-3)))
    del lp, ent
    # the reference path on the GPU: eager (what torch.compile fuses), 1024-row chunks, fwd and fwd+bwd
    def ref_fwd(xx):
        lps, ents = [], []
        for i in range(0, T, 1024):
            lsm = torch.nn.functional.log_softmax(xx[i:i + 1024].float() / 1.0, dim=-1)
            ents.append(-torch.sum(lsm.exp() * lsm, dim=-1))
            lps.append(lsm.gather(-1, labels[i:i + 1024].unsqueeze(-1)).squeeze(-1))
        return torch.cat(lps), torch.cat(ents)
    with torch.no_grad():
        ms = timeit(lambda: ref_fwd(logits), iters=5, warmup=2)
    out.append(row("reference eager gather_logprobs_entropy fwd (torch CUDA)", ms, 2 * T * V, tokens_per_s=T / (ms * 1e-3)))
    def ref_fb():
        x.grad = None
        a, b = ref_fwd(x)
        (a * w).sum().backward()
    ms = timeit(ref_fb, iters=3, warmup=1)
    out.append(row("reference eager fwd+bwd (torch CUDA)", ms, 6 * T * V, tokens_per_s=T / (ms * 1e-3)))
    try:
        cfwd = torch.compile(lambda xx, ll: (lambda lsm: (lsm.gather(-1, ll.unsqueeze(-1)).squeeze(-1),
                                                           -torch.sum(lsm.exp() * lsm, dim=-1)))(
            torch.nn.functional.log_softmax(xx.float() / 1.0, dim=-1)))
        def comp_fwd():
            return [cfwd(logits[i:i + 1024], labels[i:i + 1024]) for i in range(0, T, 1024)]
        with torch.no_grad():
            ms = timeit(comp_fwd, iters=5, warmup=3)
        out.append(row("reference torch.compile gather_logprobs_entropy fwd", ms, 2 * T * V, tokens_per_s=T / (ms * 1e-3)))
    except Exception as e:
        print("torch.compile reference unavailable:", repr(e)[:200])
    x.grad = None
    # fp32 logits (Megatron path) and misaligned rows take the generic kernel
    Tf = 2048
    lf = torch.randn(Tf, V, device=DEV)
    ms = timeit(lambda: F.gather_logprobs_entropy(lf, labels[:Tf], 1.0))
    out.append(row("logits_fwd TMA kernel (fp32 logits)", ms, 4 * Tf * V, tokens_per_s=Tf / (ms * 1e-3)))
    odd = logits.detach()[:Tf, 1:V - 7]
    ms = timeit(lambda: F.gather_logprobs_entropy(odd, labels[:Tf] % (V - 8), 1.0))
    out.append(row("logits_fwd generic kernel (bf16, misaligned rows)", ms, 2 * Tf * (V - 8), tokens_per_s=Tf / (ms * 1e-3)))


def bench_loss(out):
    for T in (16384, 131072, 1 << 22):
        lp = -torch.rand(T, device=DEV) * 5
        prox = lp + 0.02 * torch.randn(T, device=DEV)
        old = lp + 0.05 * torch.randn(T, device=DEV)
        adv = torch.randn(T, device=DEV)
        mask = torch.rand(T, device=DEV) > 0.3
        ms = timeit(lambda: F.ppo_actor_loss_fn(lp, prox, old, adv, 0.4, mask, behav_imp_weight_cap=5.0), flush=T < (1 << 22))
        out.append(row("ppo_actor_loss", ms, 45 * T, T=T, tokens_per_s=T / (ms * 1e-3)))
        v = torch.randn(T, device=DEV)
        ms = timeit(lambda: F.ppo_critic_loss_fn(v, v + 0.1, v - 0.2, 0.5, mask), flush=T < (1 << 22))
        out.append(row("ppo_critic_loss", ms, 22 * T, T=T, tokens_per_s=T / (ms * 1e-3)))


def bench_advantages(out):
    """compute_advantages (reward construction + GAE, one launch) vs the reference's op sequence for the
    reward construction alone (actor.py:96-138 restated with torch ops on the GPU; its GAE loop is timed
    in bench_padded)."""
    import types
    from areal_b200.actor import compute_advantages
    for B, S in ((512, 4096), (2048, 8192)):
        lens = torch.randint(2, S + 1, (B,), device=DEV)
        am = torch.arange(S, device=DEV)[None, :] < lens[:, None]
        lm = am & (torch.rand(B, S, device=DEV) > 0.2)
        base = dict(input_ids=torch.zeros(B, S, dtype=torch.long, device=DEV), attention_mask=am, loss_mask=lm,
                    logprobs=-torch.rand(B, S, device=DEV), ref_logp=-torch.rand(B, S, device=DEV),
                    values=torch.randn(B, S, device=DEV), rewards=torch.randn(B, device=DEV))
        actor = types.SimpleNamespace(
            config=types.SimpleNamespace(use_decoupled_loss=True, recompute_logprob=True, overlong_reward_penalty=False),
            reward_bias=0.0, reward_scaling=1.0, reward_clip=20.0, kl_ctl=0.1,
            kl_estimator=types.SimpleNamespace(kl_estimator="k1", apply_clamp=True), discount=1.0, gae_lambda=0.95,
            mask_no_eos_with_zero=False, adv_norm=None, reward_norm=None)
        # bytes: masks 1+1, logp 4, ref 4 (+4 in-place write), values 4; six f32 outputs
        nbytes = (2 + 4 + 8 + 4 + 24) * B * S
        ms = timeit(lambda: compute_advantages(actor, dict(base)), flush=nbytes < (200 << 20))
        out.append(row("advantages_padded (compute_advantages, one launch)", ms, nbytes, B=B, S=S,
                       tokens_per_s=B * S / (ms * 1e-3)))

        def ref_rewards():
            rs = (base["rewards"] + 0.0) * 1.0
            rs = torch.clip(rs, max=20.0, min=-20.0)
            m = torch.roll(base["loss_mask"].float(), shifts=-1, dims=-1)
            old = torch.roll(base["logprobs"], shifts=-1, dims=-1)
            ref = base["ref_logp"]
            ref *= m
            old *= m
            seqlens = base["attention_mask"].sum(-1).long()
            no_eos = seqlens == S
            rw = -0.1 * (old.float() - ref.float()).clamp(min=-10, max=10)
            klr = rw.clone()
            bi = torch.arange(B, device=DEV)
            rw[bi, seqlens - 1] = 0
            rw[bi, torch.clip(seqlens - 2, min=0)] += rs
            return rw, klr, no_eos
        msr = timeit(ref_rewards, iters=10, warmup=3)
        out.append(row("reference reward construction only (torch ops, actor.py:96-138)", msr, nbytes, B=B, S=S,
                       tokens_per_s=B * S / (msr * 1e-3)))


def bench_aent(out):
    """AEnt clamped entropy: log-prob + clamped entropy forward and its backward at the bench shape; the
    reference's path (recipe/AEnt/functional.py:15-52: CPU top-k on a copy of the logits, [T,V] mask and
    temporaries) restated with torch ops on a 256-row sample."""
    from areal_b200 import aent_functional as A
    T, V = 16384, 151936
    g = torch.Generator(device=DEV).manual_seed(0)
    x = torch.empty(T, V, dtype=torch.bfloat16, device=DEV)
    for i in range(0, T, 2048):
        x[i:i + 2048] = (torch.randn(2048, V, generator=g, device=DEV) * 2).bfloat16()
    lab = torch.randint(0, V, (T,), generator=g, device=DEV)
    w = torch.randn(T, generator=g, device=DEV)
    xr = x.requires_grad_()
    ms = timeit(lambda: A.gather_logprobs_clamped_entropy(xr, lab, 0.3, 1.0), iters=10, warmup=3)
    out.append(row("clamped entropy fwd (logp + clamped H, clamp 0.3)", ms, 2.0 * T * V, T=T, V=V,
                   tokens_per_s=T / (ms * 1e-3)))

    def fb():
        xr.grad = None
        lp, h = A.gather_logprobs_clamped_entropy(xr, lab, 0.3, 1.0)
        ((lp * w).sum() + (h * w).sum()).backward()
    ms2 = timeit(fb, iters=10, warmup=3)
    out.append(row("clamped entropy fwd+bwd", ms2, 6.0 * T * V, T=T, V=V, tokens_per_s=T / (ms2 * 1e-3)))
    Ts = 256
    xs = x[:Ts].detach()

    def ref():
        z = xs.float()
        lc = z.cpu()
        k = int(V * 0.3)
        _, idx = torch.topk(lc, k=k, dim=-1, largest=False)
        rm = torch.zeros_like(lc, dtype=torch.bool)
        rm[torch.arange(Ts).unsqueeze(1), idx] = True
        rm = rm.to(DEV)
        cl = z.masked_fill(rm, -torch.inf)
        lpz = torch.where(rm, 0.0, torch.log_softmax(cl, dim=-1))
        return -torch.sum(torch.softmax(cl, dim=-1) * lpz, dim=-1)
    msr = timeit(ref, iters=3, warmup=1)
    out.append(row("reference clamped entropy forward (CPU top-k path), 256 rows", msr, 2.0 * Ts * V, T=Ts, V=V,
                   tokens_per_s=Ts / (msr * 1e-3), speedup_ours=(msr / Ts) / (ms / T)))


def bench_cfg2(out):
    """BASELINE.json configs[2] per-GPU step: 32k packed tokens, packed GAE + single-pass actor loss with
    dlogits + critic loss with its gradient (batch-sharded: every GPU runs exactly this)."""
    T, V = 32768, 151936
    rng = np.random.RandomState(2)
    lens = make_packed(T, rng)
    B = len(lens)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(DEV)
    g = torch.Generator(device=DEV).manual_seed(2)
    x = torch.empty(T, V, dtype=torch.bfloat16, device=DEV)
    for i in range(0, T, 2048):
        x[i:i + 2048] = (torch.randn(2048, V, generator=g, device=DEV) * 2).bfloat16()
    lab = torch.randint(0, V, (T,), generator=g, device=DEV)
    with torch.no_grad():
        lp = F.gather_logprobs(x, lab, 1.0)
    prox = lp + 0.02 * torch.randn(T, generator=g, device=DEV)
    old = lp + 0.05 * torch.randn(T, generator=g, device=DEV)
    mask = torch.rand(T, generator=g, device=DEV) > 0.2
    rew = torch.randn(T, generator=g, device=DEV)
    val = torch.randn(T + B, generator=g, device=DEV)
    boot = torch.rand(B, generator=g, device=DEV) > 0.5
    newv = torch.randn(T, generator=g, device=DEV, requires_grad=True)
    scale = torch.full((1,), 0.5, device=DEV)
    xr = x.requires_grad_()

    def step():
        xr.grad = None
        newv.grad = None
        adv, ret = P.cugae1d_nolp_misalign_func(rew, val, cu, boot, 1.0, 0.95)
        loss, _ = F.fused_actor_loss(xr, lab, 1.0, prox, old, adv, mask, eps_clip=0.2, behav_imp_weight_cap=5.0,
                                     loss_scale=scale)
        (loss * scale[0]).backward()
        closs, _ = F.ppo_critic_loss_fn(newv, val[:T], ret, 0.2, mask)
        closs.backward()
    ms = timeit(step, iters=10, warmup=3)
    out.append(row("configs[2] per-GPU step (32k tokens: GAE + actor single pass + critic)", ms, 4.0 * T * V, T=T, V=V,
                   tokens_per_s=T / (ms * 1e-3)))


def bench_lmhead(out):
    """SURVEY 8f-4 at the configs[1] shape (Qwen2.5-1.5B: H=1536, V=151936, T=16384): LM head + actor loss +
    all gradients, over row blocks (no [T,V] tensor) against the materialised-logits path."""
    from areal_b200.lm_head import fused_lm_head_actor_loss
    T, Hd, V = 16384, 1536, 151936
    g = torch.Generator(device=DEV).manual_seed(3)
    h = (torch.randn(T, Hd, generator=g, device=DEV)).bfloat16().requires_grad_()
    w = (torch.randn(V, Hd, generator=g, device=DEV) * (2.0 / Hd ** 0.5)).bfloat16().requires_grad_()
    lab = torch.randint(0, V, (T,), generator=g, device=DEV)
    with torch.no_grad():
        lp = torch.cat([F.gather_logprobs(torch.mm(h[i:i + 2048], w.t()), lab[i:i + 2048], 1.0)
                        for i in range(0, T, 2048)])
    prox = lp + 0.02 * torch.randn(T, generator=g, device=DEV)
    old = lp + 0.05 * torch.randn(T, generator=g, device=DEV)
    adv = torch.randn(T, generator=g, device=DEV)
    mask = torch.rand(T, generator=g, device=DEV) > 0.2
    scale = torch.full((1,), 0.5, device=DEV)
    kw = dict(eps_clip=0.2, behav_imp_weight_cap=5.0)
    flops = 3 * 2.0 * T * Hd * V

    def run(fn, name, **extra):
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()
        base = torch.cuda.memory_allocated()
        ms = timeit(fn, iters=5, warmup=2)
        peak = torch.cuda.max_memory_allocated() - base
        out.append(row(name, ms, 4.0 * T * V, T=T, H=Hd, V=V, tokens_per_s=T / (ms * 1e-3),
                       gemm_tflops=round(flops / (ms * 1e-3) / 1e12, 1), peak_extra_bytes=int(peak), **extra))

    def materialised():
        h.grad = None
        w.grad = None
        logits = torch.mm(h, w.t())
        loss, _ = F.fused_actor_loss(logits, lab, 1.0, prox, old, adv, mask, loss_scale=scale, **kw)
        (loss * scale[0]).backward()
    run(materialised, "lm head + actor loss, materialised [T,V] logits (torch.mm + single pass + autograd GEMMs)")
    for chunk, wd in ((4096, torch.float32), (2048, torch.float32), (4096, torch.bfloat16), (1024, torch.bfloat16)):
        def blocks():
            h.grad = None
            w.grad = None
            loss, _ = fused_lm_head_actor_loss(h, w, lab, 1.0, prox, old, adv, mask, loss_scale=scale,
                                               chunk_rows=chunk, wgrad_dtype=wd, **kw)
            (loss * scale[0]).backward()
        run(blocks, "lm head + actor loss over row blocks (no [T,V] tensor)", chunk_rows=chunk,
            wgrad_accumulator=str(wd).replace("torch.", ""))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="gae,padded,advantages,aent,cfg2,loss,logits,intervals")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    what = set(args.what.split(","))
    out = []
    ref_gae = load_ref("ref_cugae") if "ref" in what or True else None
    ref_iop = load_ref("ref_interval_op_cuda")
    # every section runs under the clock sampler of bench.py (SM clock + throttle reasons through NVML) and is
    # followed by a {"section", "clocks"} line, so that no fraction printed here is an unclocked claim
    from bench import Clocked
    sections = [("gae", lambda: bench_gae(out, ref_gae)), ("padded", lambda: bench_padded(out)),
                ("advantages", lambda: bench_advantages(out)), ("cfg2", lambda: bench_cfg2(out)),
                ("aent", lambda: bench_aent(out)), ("lmhead", lambda: bench_lmhead(out)),
                ("loss", lambda: bench_loss(out)), ("logits", lambda: bench_logits(out)),
                ("intervals", lambda: bench_intervals(out, ref_iop))]
    for name, fn in sections:
        if name not in what:
            continue
        with Clocked(0) as c:
            fn()
            torch.cuda.synchronize()
        d = dict(section=name, clocks=c.clocks)
        print(json.dumps(d), flush=True)
        out.append(d)
    if args.out:
        json.dump(out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
