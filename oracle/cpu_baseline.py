"""TEST / BENCH INFRASTRUCTURE ONLY -- the reference's CPU path for one hot-path step, restated with
the same torch-CPU operations the reference executes (eager, no torch.compile), so that it can be
TIMED on the GPU box's host cores (`bench.py` cpu_baseline / `--impl reference`).  /root/reference
cannot travel to the GPU box, so this port stands in for it (kind = "port"); its arithmetic is
pinned by tests/test_oracle_golden.py::test_cpu_baseline_port_matches_golden.

One step = packed GAE over the batch (pygae1d_nolp_misalign's per-token Python loop,
realhf/impl/model/utils/ppo_functional.py:292-323) + log-prob / entropy from the logits
(areal/utils/functional.py:28-34, chunks of 1024 rows :104-114) + decoupled PPO actor loss
(:171-235) + loss.backward() down to dlogits.
"""
from __future__ import annotations

import time
from typing import Optional

import numpy as np
import torch

from . import cpu_oracle as O


def logp_entropy_torch(logits: torch.Tensor, labels: torch.Tensor, temperature: float, chunk: int = 1024):
    lps, ents = [], []
    for i in range(0, logits.shape[0], chunk):
        z = logits[i:i + chunk].float() / temperature
        lsm = torch.nn.functional.log_softmax(z, dim=-1)
        ents.append(-torch.sum(lsm.exp() * lsm, dim=-1))
        lps.append(lsm.gather(-1, labels[i:i + chunk].unsqueeze(-1)).squeeze(-1))
    return torch.cat(lps), torch.cat(ents)


def actor_loss_torch(lp, prox, old, adv, eps_clip, mask, eps_clip_higher=None, c_clip=None, cap=None):
    count = mask.count_nonzero() or 1
    ratio = torch.where(mask, torch.exp(lp - prox), 0)
    hi = eps_clip if eps_clip_higher is None else eps_clip_higher
    l1 = -adv * ratio
    l2 = -adv * torch.clamp(ratio, 1.0 - eps_clip, 1.0 + hi)
    pg = torch.max(l1, l2)
    if c_clip is not None:
        pg = torch.min(pg, torch.sign(adv) * c_clip * adv)
    bk = prox - old
    w = bk.exp()
    bmask = (w <= cap).logical_and(mask) if cap is not None else mask
    pg = pg * torch.where(bmask, w, 0.0)
    return torch.where(mask, pg, 0).sum() / count


def step(logits, labels, temperature, prox, old, adv, mask, cfg, gae_inputs=None):
    """One CPU step; returns (loss, dlogits, adv_gae)."""
    gae_out = None
    if gae_inputs is not None:
        r, v, cu, boot, gamma, lam = gae_inputs
        # the reference's own loop of 0-dim torch ops (not the numpy restatement, which is faster)
        gae_out = gae1d_torch_loop(torch.as_tensor(r), torch.as_tensor(v).clone(), torch.as_tensor(cu),
                                   torch.as_tensor(np.asarray(boot, dtype=np.float32)), gamma, lam)
    x = logits.detach().requires_grad_()
    lp, ent = logp_entropy_torch(x, labels, temperature)
    loss = actor_loss_torch(lp, prox, old, adv, cfg["eps_clip"], mask, cfg.get("eps_clip_higher"),
                            cfg.get("c_clip"), cfg.get("behav_imp_weight_cap"))
    loss.backward()
    return loss.detach(), x.grad, gae_out, lp.detach(), ent.detach()


def time_steps(sample_rows: int, V: int, steps: int, warmup: int, cfg, temperature=1.0, seed=0,
               threads: Optional[int] = None, with_gae=True):
    """Times `steps` CPU steps on a `sample_rows`-token sample of the workload.  Returns
    (seconds_per_step, threads_used)."""
    if threads:
        torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(seed)
    logits = (torch.randn(sample_rows, V, generator=g) * 2.0).bfloat16()
    labels = torch.randint(0, V, (sample_rows,), generator=g)
    lp0 = -torch.rand(sample_rows, generator=g) * 5
    prox = lp0 + 0.02 * torch.randn(sample_rows, generator=g)
    old = lp0 + 0.05 * torch.randn(sample_rows, generator=g)
    adv = torch.randn(sample_rows, generator=g)
    mask = torch.rand(sample_rows, generator=g) > 0.15
    gae_inputs = None
    if with_gae:
        rng = np.random.RandomState(seed)
        lens = []
        left = sample_rows
        while left > 0:
            n = min(left, int(rng.randint(128, 4097)))
            lens.append(n)
            left -= n
        cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        gae_inputs = (rng.randn(sample_rows).astype(np.float32),
                      rng.randn(sample_rows + len(lens)).astype(np.float32), cu,
                      rng.rand(len(lens)) > 0.5, 1.0, 0.95)
    for _ in range(warmup):
        step(logits, labels, temperature, prox, old, adv, mask, cfg, gae_inputs)
    t0 = time.perf_counter()
    for _ in range(steps):
        step(logits, labels, temperature, prox, old, adv, mask, cfg, gae_inputs)
    dt = (time.perf_counter() - t0) / max(steps, 1)
    return dt, torch.get_num_threads()


def time_steps_on_gpu(rows: int, V: int, steps: int, warmup: int, cfg, temperature=1.0, seed=0, device="cuda:0"):
    """The SAME restated torch operations (logp_entropy_torch + actor_loss_torch + backward; no GAE: the reference's
    GPU GAE is its own CUDA extension, which cannot travel) executed eagerly on the GPU: what the reference's
    functions cost on this device without torch.compile.  Reported beside the CPU baseline, never a target and
    never the product path.  Returns seconds per step (CUDA events)."""
    g = torch.Generator(device=device).manual_seed(seed)
    logits = torch.empty(rows, V, dtype=torch.bfloat16, device=device)
    for i in range(0, rows, 1024):
        n = min(1024, rows - i)
        logits[i:i + n] = (torch.randn(n, V, generator=g, device=device) * 2.0).bfloat16()
    labels = torch.randint(0, V, (rows,), generator=g, device=device)
    lp0 = -torch.rand(rows, generator=g, device=device) * 5
    prox = lp0 + 0.02 * torch.randn(rows, generator=g, device=device)
    old = lp0 + 0.05 * torch.randn(rows, generator=g, device=device)
    adv = torch.randn(rows, generator=g, device=device)
    mask = torch.rand(rows, generator=g, device=device) > 0.15

    def one():
        x = logits.detach().requires_grad_()
        lp, ent = logp_entropy_torch(x, labels, temperature)
        loss = actor_loss_torch(lp, prox, old, adv, cfg["eps_clip"], mask, cfg.get("eps_clip_higher"),
                                cfg.get("c_clip"), cfg.get("behav_imp_weight_cap"))
        loss.backward()
        return x.grad
    for _ in range(warmup):
        one()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        one()
    e1.record()
    torch.cuda.synchronize(device)
    peak = torch.cuda.max_memory_allocated(device)
    return e0.elapsed_time(e1) * 1e-3 / max(steps, 1), peak


# --------------------------------------------------------------------------------------------------
# BASELINE.json configs[0]: the reference's CPU GAE exactly as it executes -- a Python double loop
# with one 0-dim torch op per token and per term (realhf/impl/model/utils/ppo_functional.py:292-323)
# --------------------------------------------------------------------------------------------------
def gae1d_torch_loop(rewards: torch.Tensor, values: torch.Tensor, cu_seqlens_: torch.Tensor, bootstrap: torch.Tensor,
                     gamma: float, lam: float):
    """Same tensor operations, in the same order, as `pygae1d_nolp_misalign` -- including the in-place
    `nextvalues *= bootstrap[i]` that writes through a view into `values` (:313-315); pass a clone when the
    caller's tensor must survive."""
    cu = cu_seqlens_.clone()
    cu[1:] += torch.ones_like(cu_seqlens_[1:]).cumsum(0)                       # :300-301
    bs = cu_seqlens_.shape[0] - 1
    assert values.shape[0] == rewards.shape[0] + bs                            # :304
    adv_rev, ret_rev = [], []
    for i in reversed(range(bs)):                                              # :307
        v_off = cu[i]
        r_off, r_end = cu_seqlens_[i], cu_seqlens_[i + 1]
        assert cu[i + 1] - v_off - 1 == r_end - r_off                          # :310
        last = 0
        for t in reversed(range(r_end - r_off)):                               # :312
            nv = values[v_off + t + 1]
            if t == r_end - r_off - 1:
                nv *= bootstrap[i]                                             # :314-315 (in place, through the view)
            delta = rewards[r_off + t] + gamma * nv - values[v_off + t]        # :316
            last = delta + gamma * lam * last                                  # :317
            adv_rev.append(last)
            ret_rev.append(last + values[v_off + t])                           # :319
    return torch.stack(adv_rev[::-1]), torch.stack(ret_rev[::-1])              # :321-322


def cfg0_batch(seed: int = 0, n_seqs: int = 64, lo: int = 128, hi: int = 4096):
    """BASELINE.json configs[0]: 64 sequences of U[128, 4096] tokens, bootstrap mixed."""
    rng = np.random.RandomState(seed)
    lens = rng.randint(lo, hi + 1, n_seqs)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    T = int(cu[-1])
    return (torch.from_numpy(rng.randn(T).astype(np.float32)), torch.from_numpy(rng.randn(T + n_seqs).astype(np.float32)),
            torch.from_numpy(cu), torch.from_numpy((rng.rand(n_seqs) > 0.5).astype(np.float32)))


def time_gae_cfg0(seed: int = 0, n_seqs: int = 64, gamma: float = 1.0, lam: float = 0.95, max_seconds: float = 30.0):
    """Times the loop above on configs[0] (single Python thread, as the reference runs it).  The batch is cut
    to whole sequences that fit `max_seconds` at the rate of the first sequence, so that the call is
    bounded on a slow host; returns (tokens, seconds, n_seqs_timed)."""
    torch.set_num_threads(1)
    r, v, cu, boot = cfg0_batch(seed, n_seqs)
    # rate probe on the first sequence
    n0 = int(cu[1])
    t0 = time.perf_counter()
    gae1d_torch_loop(r[:n0], v[:n0 + 1].clone(), cu[:2], boot[:1], gamma, lam)
    rate = n0 / max(time.perf_counter() - t0, 1e-9)
    budget = int(rate * max_seconds)
    k = int(np.searchsorted(cu.numpy()[1:], budget, side="right"))
    k = max(1, min(k, n_seqs))
    T = int(cu[k])
    t0 = time.perf_counter()
    gae1d_torch_loop(r[:T], v[:T + k].clone(), cu[:k + 1], boot[:k], gamma, lam)
    return T, time.perf_counter() - t0, k


# --------------------------------------------------------------------------------------------------
# interval ops on the CPU: `_slice_intervals_py` / `_set_intervals_py`
# (realhf/impl/model/nn/flatten_param.py:65-90), the path the reference takes without its extension
# --------------------------------------------------------------------------------------------------
def slice_intervals_torch(src: torch.Tensor, intervals):
    assert len(src.shape) == 1                                                 # :67
    assert all([x[0] >= 0 for x in intervals])
    assert all([x[1] <= src.shape[0] for x in intervals])
    slices = []
    for i, j in intervals:                                                     # :72-73
        slices.append(src[i:j])
    return torch.cat(slices, dim=0)                                            # :74


def set_intervals_torch(src: torch.Tensor, dst: torch.Tensor, intervals):
    assert len(dst.shape) == len(src.shape) == 1                               # :83
    offset = 0
    for i, j in intervals:                                                     # :85-89
        assert i >= 0
        assert j <= dst.shape[0], (j, dst.shape[0])
        dst[i:j] = src[offset: offset + j - i]
        offset += j - i
    assert offset == src.shape[0]                                              # :90


def time_intervals_cpu(n_elems: int = 100_000_000, n_intervals: int = 10_000, seed: int = 0, iters: int = 3,
                       threads: Optional[int] = None):
    """The reference test's pattern (realhf/tests/cpp_extensions/test_interval_ops.py:20-31: interval i =
    (i*s, i*s + s//2), shuffled) on a bf16 buffer of `n_elems`; returns (bytes moved per op, seconds per
    slice, seconds per set, threads)."""
    if threads:
        torch.set_num_threads(threads)
    rng = np.random.RandomState(seed)
    s = n_elems // n_intervals
    iv = np.stack([np.arange(n_intervals, dtype=np.int64) * s, np.arange(n_intervals, dtype=np.int64) * s + s // 2], 1)
    rng.shuffle(iv)
    ivl = [(int(a), int(b)) for a, b in iv]
    src = torch.empty(n_elems, dtype=torch.bfloat16)
    src.view(torch.int16).random_(-30000, 30000)
    packed = slice_intervals_torch(src, ivl)
    t0 = time.perf_counter()
    for _ in range(iters):
        packed = slice_intervals_torch(src, ivl)
    t_slice = (time.perf_counter() - t0) / iters
    t0 = time.perf_counter()
    for _ in range(iters):
        set_intervals_torch(packed, src, ivl)
    t_set = (time.perf_counter() - t0) / iters
    nbytes = 2 * 2 * n_intervals * (s // 2)
    return nbytes, t_slice, t_set, torch.get_num_threads()
