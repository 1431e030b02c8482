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
        gae_out = O.gae1d_nolp_misalign(r, v, cu, boot, gamma, lam)
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
