"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's algorithms for the
RL post-processing hot path.  Nothing under `areal_b200/`
imports this module; only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` do, and only as the checker
or as the timed CPU baseline -- never as the product path.

Parity status: PINNED.  Every function below is checked in
`tests/test_oracle_golden.py` against golden vectors produced by executing the
UNMODIFIED reference functions in the build container
(`oracle/gen_golden.py` -> `tests/golden/*.npz`), and -- where the reference has its
own known-answer test -- against that test's independent formula
(`realhf/tests/data/test_dual_clip.py:8-50`).  Two parts have no CPU-runnable reference
function and are pinned by running the reference's own code another way:
`gae_2d_olp` (no Python caller and no test in the reference; restated from
`csrc/cugae/gae.cu:62-89`) against outputs of the reference's CUDA kernels compiled from
where they lie and run on a B200 (`oracle/gen_golden_gpu.py` -> `tests/golden/gae2d.npz`),
and `shard_partials` / `combine_partials` against the reference's vocabulary-sharded code on
two gloo ranks (`oracle/gen_golden_tp.py` -> `tests/golden/vocab_parallel.npz`; megatron-core
itself is third-party and absent, the tree carries a copy of its cross entropy).

All file:line citations are relative to the reference root.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

f32 = np.float32


# --------------------------------------------------------------------------- #
# P1  packed GAE
# --------------------------------------------------------------------------- #
def gae1d_nolp_misalign(rewards, values, cu_seqlens, bootstrap, gamma: float, lam: float):
    """fp32, sequential, same operation order as `pygae1d_nolp_misalign`
    (realhf/impl/model/utils/ppo_functional.py:292-323): per sequence i, t from last
    to first, nv = values[cu[i]+i+t+1] (times bootstrap[i] at the last step, :313-315),
    delta = r + gamma*nv - v (:316), A = delta + gamma*lam*A (:317), ret = A + v (:319).
    torch multiplies a 0-dim fp32 tensor by a Python float in fp32 (the scalar is cast
    to fp32), and `gamma * lam` is a Python double product (:317) -- both reproduced.
    Unlike the reference this does NOT write through to `values`."""
    r = np.asarray(rewards, dtype=f32)
    v = np.asarray(values, dtype=f32)
    cu = np.asarray(cu_seqlens, dtype=np.int64)
    bs = cu.shape[0] - 1
    boot = np.asarray(bootstrap).astype(f32)
    assert v.shape[0] == r.shape[0] + bs
    g = f32(gamma)
    gl = f32(gamma * lam)
    adv = np.zeros_like(r)
    ret = np.zeros_like(r)
    for i in range(bs):
        rs, re = int(cu[i]), int(cu[i + 1])
        vs = rs + i
        last = f32(0.0)
        n = re - rs
        for t in range(n - 1, -1, -1):
            nv = v[vs + t + 1]
            if t == n - 1:
                nv = f32(nv * boot[i])
            delta = f32(f32(r[rs + t] + f32(g * nv)) - v[vs + t])
            last = f32(delta + f32(gl * last))
            adv[rs + t] = last
            ret[rs + t] = f32(last + v[vs + t])
    return adv, ret


def gae1d_nolp_misalign_f64(rewards, values, cu_seqlens, bootstrap, gamma: float, lam: float):
    """fp64 arbiter of the same recurrence: delta is formed exactly as the fp32
    reference forms it, the recurrence itself is carried in fp64 with the fp32-rounded
    coefficient.  Used to report both implementations' distance from the exact result
    (SURVEY section 7 'fp32 reassociation in scans')."""
    r = np.asarray(rewards, dtype=f32)
    v = np.asarray(values, dtype=f32)
    cu = np.asarray(cu_seqlens, dtype=np.int64)
    bs = cu.shape[0] - 1
    boot = np.asarray(bootstrap).astype(f32)
    g = f32(gamma)
    gl = float(f32(gamma * lam))
    adv = np.zeros(r.shape[0], dtype=np.float64)
    for i in range(bs):
        rs, re = int(cu[i]), int(cu[i + 1])
        vs = rs + i
        n = re - rs
        if n == 0:
            continue
        nv = v[vs + 1: vs + n + 1].copy()
        nv[-1] = f32(nv[-1] * boot[i])
        delta = ((r[rs:re] + g * nv).astype(f32) - v[vs: vs + n]).astype(f32).astype(np.float64)
        last = 0.0
        for t in range(n - 1, -1, -1):
            last = delta[t] + gl * last
            adv[rs + t] = last
    ret = (adv.astype(f32) + v_at_rewards(v, cu)).astype(f32)
    return adv, ret


def v_at_rewards(values, cu_seqlens):
    """values[cu[i]+i+t] laid out like rewards (the 'misalign' indexing, ppo_functional.py:301-311)."""
    cu = np.asarray(cu_seqlens, dtype=np.int64)
    out = np.empty(int(cu[-1]), dtype=np.asarray(values).dtype)
    for i in range(cu.shape[0] - 1):
        rs, re = int(cu[i]), int(cu[i + 1])
        out[rs:re] = values[rs + i: re + i]
    return out


def gae_padded_masked(rewards, values, loss_mask, seq_no_eos_mask, gamma: float, lam: float):
    """areal-lite's padded [B,S] masked GAE, `PPOActor.compute_advantages`
    (areal/engine/ppo/actor.py:131-152): nv0 = values[:,S-1]*no_eos (:135); for t=S-2..0:
    delta = r_t + gamma*nv - v_t (:137); new = delta + gamma*lam*last (:138); m = mask_t;
    nv = nv*(1-m) + v_t*m (:142); last = last*(1-m) + new*m (:143); adv_t = last;
    adv[:,S-1] = 0 (:131-133); returns = adv + values (:147).  fp32, same op order."""
    r = np.asarray(rewards, dtype=f32)
    v = np.asarray(values, dtype=f32)
    m_all = np.asarray(loss_mask, dtype=f32)
    B, S = r.shape
    g = f32(gamma)
    gl = f32(gamma * lam)
    adv = np.zeros((B, S), dtype=f32)
    last = np.zeros(B, dtype=f32)
    nv = (v[:, S - 1] * np.asarray(seq_no_eos_mask).astype(f32)).astype(f32)
    one = f32(1.0)
    for t in range(S - 2, -1, -1):
        delta = ((r[:, t] + g * nv).astype(f32) - v[:, t]).astype(f32)
        new = (delta + (gl * last).astype(f32)).astype(f32)
        m = m_all[:, t]
        nv = ((nv * (one - m)).astype(f32) + (v[:, t] * m).astype(f32)).astype(f32)
        last = ((last * (one - m)).astype(f32) + (new * m).astype(f32)).astype(f32)
        adv[:, t] = last
    ret = (adv + v).astype(f32)
    return adv, ret


def gae_2d_olp(rewards, values, done, done_y_indices, cu_num_dones, is_truncate,
               gamma: float, lam: float):
    """Segment-table 2D GAE as it EXECUTES in the reference (`gae_kernel_2d_olp`,
    csrc/cugae/gae.cu:62-89; the `_nolp` host wrapper launches this same kernel,
    gae.cu:199).  Thread (b, seg) walks [t_start, t_end) backwards; nex_v = 0 when
    t is the segment tail and done[b,t+1]; a truncated non-final segment zeroes its
    tail GAE (:85).  No Python caller / test upstream: pinned against outputs of the kernel
    itself (tests/golden/gae2d.npz)."""
    r = np.asarray(rewards, dtype=f32)
    v = np.asarray(values, dtype=f32)
    d = np.asarray(done).astype(bool)
    idx = np.asarray(done_y_indices, dtype=np.int64)
    cu = np.asarray(cu_num_dones, dtype=np.int64)
    trunc = np.asarray(is_truncate).astype(bool)
    B, H = r.shape
    g = f32(gamma)
    gl = f32(f32(gamma) * f32(lam))  # device code multiplies two floats (gae.cu:84)
    adv = np.zeros((B, H), dtype=f32)
    ret = np.zeros((B, H), dtype=f32)
    for b in range(B):
        breadth = int(cu[b + 1] - cu[b]) + 1
        for s in range(breadth):
            t_start = 0 if s == 0 else int(idx[cu[b] + s - 1])
            t_end = H if s == breadth - 1 else int(idx[cu[b] + s])
            abandon = (s != breadth - 1) and bool(trunc[cu[b] + s])
            last = f32(0.0)
            for t in range(t_end - 1, t_start - 1, -1):
                nex_v = f32(0.0) if (t == t_end - 1 and d[b, t + 1]) else v[b, t + 1]
                cur_v = v[b, t]
                delta = f32(f32(r[b, t] + f32(g * nex_v)) - cur_v)
                last = f32(delta + f32(gl * last))
                if t == t_end - 1 and abandon:
                    last = f32(0.0)
                adv[b, t] = last
                ret[b, t] = f32(last + cur_v)
    return adv, ret


# --------------------------------------------------------------------------- #
# P2  log-prob / entropy from [T,V] logits
# --------------------------------------------------------------------------- #
def gather_logprobs_entropy(logits, labels, temperature: float = 1.0, dtype=np.float64):
    """`_gather_logprobs_entropy` (areal/utils/functional.py:28-34): z = logits.float()/temp;
    lp = log_softmax(z); entropy = -sum(exp(lp)*lp); logp = lp[label].  `dtype=np.float32`
    restates the reference's fp32 arithmetic; float64 (default) is the arbiter.  Also
    returns lse (natural log, of z) and raw row min / max (areal/engine/ppo/actor.py:375-376)."""
    x = np.asarray(logits)
    raw = x.astype(np.float32)
    z = raw.astype(dtype) / dtype(temperature)
    m = z.max(axis=-1, keepdims=True)
    e = np.exp(z - m)
    s = e.sum(axis=-1, keepdims=True)
    lse = (m + np.log(s))
    lp = z - lse
    ent = -(np.exp(lp) * lp).sum(axis=-1)
    lab = np.asarray(labels, dtype=np.int64)
    logp = np.take_along_axis(lp, lab[:, None], axis=-1)[:, 0]
    return logp, ent, lse[:, 0], raw.min(axis=-1), raw.max(axis=-1)


def dlogits(logits, labels, temperature, g_logp, g_entropy=None, dtype=np.float64):
    """Gradient of sum_t g_logp[t]*logp[t] + g_entropy[t]*H[t] w.r.t. the raw logits:
    d logp_y / d z_v = 1[v=y] - p_v ;  d H / d z_v = -p_v (lp_v + H)
    (same as `_VocabParallelEntropy.backward`, areal/utils/mcore/functional.py:41-52:
    -p*(z - sum p z)); times 1/temperature.  Autograd of functional.py:20-34 yields this."""
    x = np.asarray(logits).astype(np.float32).astype(dtype)
    z = x / dtype(temperature)
    m = z.max(axis=-1, keepdims=True)
    e = np.exp(z - m)
    s = e.sum(axis=-1, keepdims=True)
    lp = z - (m + np.log(s))
    p = np.exp(lp)
    g = -p * np.asarray(g_logp, dtype=dtype)[:, None]
    lab = np.asarray(labels, dtype=np.int64)
    g[np.arange(x.shape[0]), lab] += np.asarray(g_logp, dtype=dtype)
    if g_entropy is not None:
        H = -(p * lp).sum(axis=-1, keepdims=True)
        g += np.asarray(g_entropy, dtype=dtype)[:, None] * (-p * (lp + H))
    return g / dtype(temperature)


def calc_entropy_legacy(logits, dtype=np.float64):
    """Legacy `calc_entropy` (realhf/impl/model/interface/ppo_interface.py:52-58):
    p = softmax(logits); -sum p*log(p + 1e-7)."""
    z = np.asarray(logits).astype(np.float32).astype(dtype)
    m = z.max(axis=-1, keepdims=True)
    e = np.exp(z - m)
    p = e / e.sum(axis=-1, keepdims=True)
    return -(p * np.log(p + dtype(1e-7))).sum(axis=-1)


def vocab_parallel_partials(logits_shard, labels, vocab_start: int, temperature: float):
    """Per-shard partials of the vocab-sharded log-softmax: row max m (of z),
    s = sum exp(z-m), u = sum exp(z-m)*z, z_label if the label is local else 0.
    Follows `_VocabParallelEntropy.forward` (areal/utils/mcore/functional.py:15-34) and
    `_VocabParallelCrossEntropy` (realhf/impl/model/parallelism/tensor_parallel/modules.py:1063-1118)."""
    z = np.asarray(logits_shard).astype(np.float32).astype(np.float64) / temperature
    m = z.max(axis=-1)
    e = np.exp(z - m[:, None])
    s = e.sum(axis=-1)
    u = (e * z).sum(axis=-1)
    lab = np.asarray(labels, dtype=np.int64) - vocab_start
    local = (lab >= 0) & (lab < z.shape[1])
    zl = np.where(local, z[np.arange(z.shape[0]), np.clip(lab, 0, z.shape[1] - 1)], 0.0)
    return m, s, u, zl


def vocab_parallel_combine(parts: Sequence[Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]]):
    """all_reduce(MAX) of m, rescale, all_reduce(SUM) of (s,u,z_label) -> logp, entropy, lse
    (mcore/functional.py:16-31: entropy = max + log(sum_exp) - sum softmax*logits)."""
    M = np.max(np.stack([p[0] for p in parts]), axis=0)
    S = sum(p[1] * np.exp(p[0] - M) for p in parts)
    U = sum(p[2] * np.exp(p[0] - M) for p in parts)
    ZL = sum(p[3] for p in parts)
    lse = M + np.log(S)
    return ZL - lse, lse - U / S, lse


# --------------------------------------------------------------------------- #
# P3  PPO losses
# --------------------------------------------------------------------------- #
def ppo_actor_loss(logprobs, proximal_logprobs, old_logprobs, advantages, eps_clip: float,
                   loss_mask, eps_clip_higher: Optional[float] = None,
                   c_clip: Optional[float] = None, behav_imp_weight_cap: Optional[float] = None,
                   dtype=np.float32) -> Tuple[float, Dict[str, np.ndarray], np.ndarray]:
    """`ppo_actor_loss_fn` (areal/utils/functional.py:171-235); the legacy `actor_loss_fn`
    (realhf/impl/model/utils/ppo_functional.py:51-149) is the same math with
    eps_clip_higher=None and, when proximal_logprobs is None, denorm=old and no behaviour
    weight (pass `proximal_logprobs=None`).  Returns (loss, stats, dloss/dlogprobs).

    Gradient (closed form): mask * w * sel * (-A * r) / count, sel = 1 iff the
    unclipped branch of torch.max is active (L1 > L2, or L1 == L2 with the clamp passing the
    gradient: for an exact tie torch.max splits 1/2-1/2 and the clamp branch contributes its
    own 1/2 only when r is inside [lo, hi]) and the dual clip is not active (torch.min:
    L < L3 -> 1, L == L3 -> 1/2)."""
    lp = np.asarray(logprobs, dtype=dtype)
    old = np.asarray(old_logprobs, dtype=dtype)
    A = np.asarray(advantages, dtype=dtype)
    mask = np.asarray(loss_mask).astype(bool)
    decoupled = proximal_logprobs is not None
    prox = np.asarray(proximal_logprobs, dtype=dtype) if decoupled else old
    count = int(mask.sum()) or 1                                    # :190
    with np.errstate(over="ignore", invalid="ignore"):
        ratio = np.where(mask, np.exp(lp - prox), dtype(0))         # :191
    lo = dtype(1.0 - eps_clip)
    hi = dtype(1.0 + (eps_clip if eps_clip_higher is None else eps_clip_higher))
    clipped = np.clip(ratio, lo, hi)                                # :193-197
    l1 = -A * ratio
    l2 = -A * clipped
    clip_mask = l1 < l2                                             # :201
    pg = np.maximum(l1, l2)                                         # :202
    # d pg / d ratio (torch.max tie -> 1/2 each; clamp passes grad iff lo <= r <= hi)
    inside = (ratio >= lo) & (ratio <= hi)
    # backward of torch.max(a, b): grad, halved on a tie, zeroed only for the strictly smaller side (so a NaN
    # operand leaves the full gradient on both sides)
    w1 = np.where(l1 < l2, 0.0, np.where(l1 == l2, 0.5, 1.0))
    w2 = np.where(l1 > l2, 0.0, np.where(l1 == l2, 0.5, 1.0))
    dpg = -A * (w1 + np.where(inside, w2, 0.0))
    if c_clip is not None:
        assert c_clip > 1.0
        l3 = np.sign(A) * dtype(c_clip) * A                         # :205
        dual_mask = l3 < pg                                         # :206
        wmin = np.where(pg > l3, 0.0, np.where(pg == l3, 0.5, 1.0))
        pg = np.minimum(pg, l3)                                     # :207
        dpg = dpg * wmin
    else:
        dual_mask = np.zeros_like(clip_mask)
    stats: Dict[str, np.ndarray] = {}
    if decoupled:
        bk = prox - old                                             # :210
        with np.errstate(over="ignore"):
            w = np.exp(bk)
        bmask = ((w <= dtype(behav_imp_weight_cap)) & mask) if behav_imp_weight_cap is not None else mask
        bk = np.where(bmask, bk, dtype(0))
        w = np.where(bmask, w, dtype(0))
        pg = pg * w                                                 # :219
        dpg = dpg * w
        stats.update(behave_imp_weight=w.astype(dtype), behave_approx_kl=bk.astype(dtype),
                     behave_mask=bmask)
    logging_loss = pg.astype(dtype)
    loss = np.where(mask, pg, dtype(0)).astype(np.float64).sum() / count   # :221
    with np.errstate(over="ignore", invalid="ignore"):
        # torch.where hands the masked side a zero gradient, exp's backward multiplies it by exp(lp - prox):
        # 0 for finite values, NaN for NaN / inf ones (what autograd returns)
        grad = np.where(mask, dpg, 0.0) * np.exp(lp - prox) / count
    stats.update(loss=logging_loss, importance_weight=ratio.astype(dtype),
                 approx_kl=(lp - prox).astype(dtype), clip_mask=clip_mask & mask,
                 dual_clip_mask=dual_mask & mask)
    return float(loss), stats, grad.astype(dtype)


def ppo_critic_loss(value, old_value, target_value, value_eps_clip: float, loss_mask=None,
                    loss_fn_type: str = "mse", dtype=np.float32):
    """`ppo_critic_loss_fn` (areal/utils/functional.py:247-311) == legacy `critic_loss_fn`
    (realhf/impl/model/utils/ppo_functional.py:161-225).  Returns (loss, stats, dloss/dvalue).
    Mean over mask.count_nonzero() with NO `or 1` (:306) -> nan/inf for an all-false mask."""
    v = np.asarray(value, dtype=dtype)
    old = np.asarray(old_value, dtype=dtype)
    tgt = np.asarray(target_value, dtype=dtype)

    def f(x, y):
        d = x - y
        if loss_fn_type == "mse":
            return dtype(0.5) * d * d, d
        if loss_fn_type == "huber":
            delta = dtype(10.0)
            ad = np.abs(d)
            return (np.where(ad < delta, dtype(0.5) * ad * ad, delta * (ad - dtype(0.5) * delta)),
                    np.where(ad < delta, d, delta * np.sign(d)))
        raise NotImplementedError(loss_fn_type)

    eps = dtype(value_eps_clip)
    diff = v - old
    vc = old + np.clip(diff, -eps, eps)
    lo_, dlo = f(v, tgt)
    lc_, dlc = f(vc, tgt)
    inside = (diff >= -eps) & (diff <= eps)
    loss_el = np.maximum(lo_, lc_)
    w1 = np.where(lo_ < lc_, 0.0, np.where(lo_ == lc_, 0.5, 1.0))
    w2 = np.where(lo_ > lc_, 0.0, np.where(lo_ == lc_, 0.5, 1.0))
    g_el = w1 * dlo + np.where(inside, w2 * dlc, 0.0)
    clip_mask = lc_ > lo_
    if loss_mask is not None:
        mask = np.asarray(loss_mask).astype(bool)
        clip_mask = clip_mask & mask
        cnt = int(mask.sum())
        with np.errstate(divide="ignore", invalid="ignore"):
            loss = np.float64(np.where(mask, loss_el, dtype(0)).astype(np.float64).sum()) / np.float64(cnt)
            grad = np.where(mask, g_el, 0.0) / np.float64(cnt)
    else:
        loss = loss_el.astype(np.float64).mean()
        grad = g_el / loss_el.size
    return float(loss), dict(clip_mask=clip_mask, loss=loss_el.astype(dtype)), grad.astype(dtype)


def ppo_dual_clip_known_answer(logp_new, logp_old, adv, clip_ratio, dual_clip):
    """Independent dual-clip formula the reference tests `actor_loss_fn` against
    (realhf/tests/data/test_dual_clip.py:8-50, PPOxFamily): used here the same way."""
    ratio = np.exp(logp_new - logp_old)
    surr1 = ratio * adv
    surr2 = np.clip(ratio, 1 - clip_ratio, 1 + clip_ratio) * adv
    clip1 = np.minimum(surr1, surr2)
    if dual_clip is not None:
        clip2 = np.maximum(clip1, dual_clip * adv)
        return float(-(np.where(adv < 0, clip2, clip1)).mean())
    return float(-clip1.mean())


# --------------------------------------------------------------------------- #
# P4  interval ops
# --------------------------------------------------------------------------- #
def slice_intervals(src: np.ndarray, intervals) -> np.ndarray:
    """`_slice_intervals_py` (realhf/impl/model/nn/flatten_param.py:65-75): concatenation of
    src[s:e] in the GIVEN order."""
    iv = np.asarray(intervals, dtype=np.int64).reshape(-1, 2)
    assert (iv[:, 0] >= 0).all() and (iv[:, 1] <= src.shape[0]).all()
    if iv.shape[0] == 0:
        return src[:0].copy()
    return np.concatenate([src[s:e] for s, e in iv])


def set_intervals(src: np.ndarray, dst: np.ndarray, intervals) -> None:
    """`_set_intervals_py` (flatten_param.py:78-90): dst[s:e] = src[off:off+e-s], off running
    in the given order; asserts the packed source is consumed exactly (:90)."""
    off = 0
    for s, e in np.asarray(intervals, dtype=np.int64).reshape(-1, 2):
        assert s >= 0 and e <= dst.shape[0]
        dst[s:e] = src[off: off + e - s]
        off += e - s
    assert off == src.shape[0]


def merge_intervals(intervals: List[Tuple[int, int]]) -> List[Tuple[int, int]]:
    """`merge_intervals` (csrc/interval_op/interval_op.cpp:4-25): merge (a,b),(b,c)->(a,c)
    for NEIGHBOURS in input order only; no sorting, no overlap handling."""
    out: List[Tuple[int, int]] = []
    for a, b in intervals:
        if out and out[-1][1] == a:
            out[-1] = (out[-1][0], b)
        else:
            out.append((int(a), int(b)))
    return out


def make_intervals(maxsize: int, n_intervals: int, rng: np.random.RandomState):
    """Workload generator of the reference test (realhf/tests/cpp_extensions/test_interval_ops.py:20-31):
    interval i = (i*s, i*s + s//2), s = maxsize // n, then shuffled."""
    assert maxsize // n_intervals > 1
    s = maxsize // n_intervals
    iv = np.stack([np.arange(n_intervals, dtype=np.int64) * s,
                   np.arange(n_intervals, dtype=np.int64) * s + s // 2], axis=1)
    rng.shuffle(iv)
    return iv, n_intervals * (s // 2), s // 2


def shard_partials(logits_shard, labels, vocab_start: int, temperature: float) -> np.ndarray:
    """Per-shard partials in the layout of `areal_b200_logits_partial_fwd` ([4,T] rows: raw max m,
    S = sum 2^t, W = sum 2^t * t with t = (x - m) * log2(e)/temperature, raw label logit or 0) -- the
    same information as vocab_parallel_partials() above (reference: areal/utils/mcore/functional.py:15-34),
    shifted by the shard max so that every term stays small."""
    x = np.asarray(logits_shard).astype(np.float32).astype(np.float64)
    c2 = math.log2(math.e) / temperature
    m = x.max(axis=-1)
    t = (x - m[:, None]) * c2
    e = np.exp2(t)
    lab = np.asarray(labels, dtype=np.int64) - vocab_start
    local = (lab >= 0) & (lab < x.shape[1])
    xl = np.where(local, x[np.arange(x.shape[0]), np.clip(lab, 0, x.shape[1] - 1)], 0.0)
    return np.stack([m, e.sum(-1), (e * t).sum(-1), xl])


def combine_partials(all_parts: np.ndarray, temperature: float):
    """[R,4,T] -> (logp, entropy, row_max, ln_sumexp): the math of `areal_b200_vocab_combine`
    (all_reduce(MAX) + rescale + all_reduce(SUM) in the reference, mcore/functional.py:16-31)."""
    p = np.asarray(all_parts, dtype=np.float64)
    c2 = math.log2(math.e) / temperature
    M = p[:, 0].max(axis=0)
    d = (p[:, 0] - M) * c2
    f = np.exp2(d)
    S = (p[:, 1] * f).sum(0)
    W = (f * (p[:, 2] + d * p[:, 1])).sum(0)
    XL = p[:, 3].sum(0)
    ln_s = np.log(S)
    return (XL - M) / temperature - ln_s, ln_s - math.log(2.0) * W / S, M, ln_s


# --------------------------------------------------------------------------- #
# P2, legacy (realhf) spelling
# --------------------------------------------------------------------------- #
def leave_one_indices(total: int, cu_seqlens) -> np.ndarray:
    """`build_leave_one_indices` (realhf/impl/model/utils/functional.py:129-166): every position but
    the last of each sequence."""
    cu = np.asarray(cu_seqlens, dtype=np.int64)
    keep = np.ones(total, dtype=bool)
    keep[cu[1:] - 1] = False
    return np.nonzero(keep)[0]


def shift_one_indices(total: int, cu_seqlens) -> np.ndarray:
    """`build_shift_one_indices` (functional.py:91-126): every position but the first of each sequence."""
    cu = np.asarray(cu_seqlens, dtype=np.int64)
    keep = np.ones(total, dtype=bool)
    keep[cu[:-1]] = False
    return np.nonzero(keep)[0]


def gather_packed_shifted_log_probs(logits, cu_seqlens, ids):
    """`gather_packed_shifted_log_probs` (functional.py:189-234): labels shifted left by one (last
    padded with 0), log_softmax gathered, last token of each sequence dropped.  fp64."""
    ids = np.asarray(ids, dtype=np.int64)
    lab = np.concatenate([ids[1:], [0]])
    lp, *_ = gather_logprobs_entropy(logits, lab, 1.0)
    return lp[leave_one_indices(len(ids), cu_seqlens)]


def calc_entropy_packed(logits, cu_seqlens):
    """`calc_entropy` (realhf/impl/model/interface/ppo_interface.py:52-58)."""
    return calc_entropy_legacy(logits)[leave_one_indices(np.asarray(logits).shape[0], cu_seqlens)]


# --------------------------------------------------------------------------- #
# next row: advantage / reward normalisation
# --------------------------------------------------------------------------- #
def masked_normalization(x, mask=None, unbiased=False, eps=1e-5):
    """`masked_normalization` (areal/utils/functional.py:130-168), dim=None, high_precision, single
    process: fp64 moments of x*mask, var = E[x^2] - mean^2, output ((x*mask - mean)/(sqrt(var)+eps)).float()."""
    xd = np.asarray(x, dtype=np.float64)
    if mask is None:
        factor = float(xd.size)
    else:
        md = np.asarray(mask).astype(np.float64)
        xd = xd * md
        factor = md.sum()
    mean = xd.sum() / factor
    var = (xd ** 2).sum() / factor - mean ** 2
    if unbiased:
        var *= factor / (factor - 1)
    return ((xd - mean) / (np.sqrt(var) + eps)).astype(np.float32)


def normalization(x, loss_mask, mean_level, std_level, mean_leave1out, std_unbiased, eps, group_size):
    """`Normalization.__call__` (areal/utils/data.py:1105-1204) with `_compute_mean` (:1206-1262) and
    `_compute_std` (:1264-1304), single process, high_precision.  The dtype path is part of the result:
    batch-level statistics stay fp64 until the final `.float()`, group-level ones are stored into
    `zeros_like(x)` (fp32) first (:1131, :1173)."""
    x32 = np.asarray(x, dtype=np.float32)
    bs = x32.shape[0]
    m32 = None if loss_mask is None else np.asarray(loss_mask, dtype=np.float32)
    if m32 is not None and m32.sum() == 0:
        return x32.copy()                                                   # :1115-1116

    def compute_mean(xx, mm, loo):
        xd = xx.astype(np.float64)
        if mm is None:
            factor = float(xd.size)
            xm = xd
        else:
            md = mm.astype(np.float64)
            xm = xd * md
            factor = md.sum()
        xs = xm.sum()
        if loo:
            if factor <= 1:
                return np.zeros((1,) * xd.ndim)
            if mm is None:
                return (xs - xd) / (factor - 1)
            return np.where(md > 0, (xs - xm) / np.maximum(factor - md, 1.0), xs / factor)
        if factor == 0:
            return np.zeros((1,) * xd.ndim)
        return np.full((1,) * xd.ndim, xs / factor)

    def compute_std(xx, mm, mean, unbiased):
        xd = xx.astype(np.float64)
        mean = np.asarray(mean, dtype=np.float64)
        if mm is None:
            factor = float(xd.size)
            ss = ((xd - mean) ** 2).sum()
        else:
            md = mm.astype(np.float64)
            factor = md.sum()
            ss = ((xd * md - mean * md) ** 2).sum()
        if unbiased:
            return 1.0 if factor <= 1 else np.sqrt(ss / (factor - 1))
        return 1.0 if factor == 0 else np.sqrt(ss / factor)

    ng = bs // group_size if group_size else 0
    if mean_level == "batch":
        mean = np.broadcast_to(compute_mean(x32, m32, mean_leave1out), x32.shape).astype(np.float64)
    elif mean_level == "group":
        mean = np.zeros_like(x32)
        for i in range(ng):
            s = slice(i * group_size, (i + 1) * group_size)
            if group_size == 1 and mean_leave1out:
                gm = np.zeros((1, x32.shape[1]))
            else:
                gm = compute_mean(x32[s], None if m32 is None else m32[s], mean_leave1out)
            mean[s] = np.broadcast_to(gm, x32[s].shape).astype(np.float32)
    else:
        mean = np.zeros_like(x32)
    xc = x32 - mean                      # fp64 for batch level, fp32 otherwise (type promotion)
    if m32 is not None:
        xc = xc * m32
    if std_level == "batch":
        std = np.float64(compute_std(x32, m32, mean, std_unbiased))
    elif std_level == "group":
        std = np.zeros_like(x32)
        for i in range(ng):
            s = slice(i * group_size, (i + 1) * group_size)
            if group_size == 1 and std_unbiased:
                gs_ = 1.0
            else:
                gs_ = compute_std(x32[s], None if m32 is None else m32[s], mean[s], std_unbiased)
            std[s] = np.float32(gs_)
    else:
        std = np.ones_like(x32)
        eps = 0.0
    if isinstance(std, np.ndarray) and std.dtype == np.float32 and xc.dtype == np.float32:
        return (xc / (std + np.float32(eps))).astype(np.float32)
    return (xc / (std + eps)).astype(np.float32)


# --------------------------------------------------------------------------------------------------
# PPOActor.compute_advantages: reward construction + padded GAE (areal/engine/ppo/actor.py:92-152)
# --------------------------------------------------------------------------------------------------
def approx_kl(log_probs, log_probs_base, kind: str = "k1", apply_clamp: bool = True):
    """KLEstimator._compute_approx_kl (areal/utils/data.py:1328-1364), float32 step by step."""
    f = np.float32
    lr = log_probs.astype(f) - log_probs_base.astype(f)                     # :1344
    if kind == "k2":
        lr = (lr * lr) / f(2.0)                                             # :1353
    elif kind == "k3":
        lr = -lr                                                            # :1359
        lr = (np.exp(lr).astype(f) - f(1)) - lr                             # :1360
    elif kind != "k1":
        raise ValueError(f"Invalid KL estimator: {kind}. Valid choices: k1, k2, k3")
    if apply_clamp:
        lr = np.clip(lr, f(-10), f(10))                                     # :1363
    return lr.astype(f)


def overlong_penalty(score, loss_mask, overlong_tokens: int, factor: float, max_response_length: int):
    """reward_overlong_penalty (areal/utils/functional.py:376-398); returns the new scores."""
    f = np.float32
    resp = loss_mask.astype(np.int64).sum(-1)                               # :384
    exceed = resp - (max_response_length - overlong_tokens)                 # :389-390
    a = ((-exceed).astype(f) / f(overlong_tokens)) * f(factor)              # :392
    return (score.astype(f) + np.where(a > 0, f(0), a).astype(f)).astype(f)  # min(a, 0), :391-394


def compute_advantages(score, loss_mask, logprobs, attention_mask, *, ref_logp=None, values=None, roll_logp=True,
                       reward_bias=0.0, reward_scaling=1.0, reward_clip=20.0, apply_scaling=True, kl_ctl=0.1,
                       kl_estimator="k1", kl_clamp=True, mask_no_eos_with_zero=False, discount=1.0, gae_lambda=1.0):
    """actor.py:96-152 between the optional penalties / normalisations.  `loss_mask` is NOT rolled.
    Returns dict(advantages, returns, kl_rewards, tot_rewards, loss_mask, logprobs, ref_logp)."""
    f = np.float32
    B, S = logprobs.shape
    sc = score.astype(f)
    if apply_scaling:
        sc = (sc + f(reward_bias)) * f(reward_scaling)                      # :98
        sc = np.clip(sc, f(-reward_clip), f(reward_clip))                   # :99-101
    lm = np.roll(loss_mask.astype(f), -1, axis=-1)                          # :109-110
    old = (np.roll(logprobs.astype(f), -1, axis=-1) if roll_logp else logprobs.astype(f)) * lm   # :114-116, :122
    ref = (np.zeros_like(old) if ref_logp is None else ref_logp.astype(f)) * lm                  # :120-121
    seqlens = attention_mask.astype(np.int64).sum(-1)                       # :126
    no_eos = seqlens == S                                                   # :127
    kl_rewards = (f(-kl_ctl) * approx_kl(old, ref, kl_estimator, kl_clamp)).astype(f)             # :128
    rewards = kl_rewards.copy()
    rows = np.arange(B)
    rewards[rows, seqlens - 1] = 0                                          # :131 (negative index wraps)
    idx = np.clip(seqlens - 2, 0, None)                                     # :132
    add = np.where(no_eos, f(0), sc) if mask_no_eos_with_zero else sc       # :133-138
    rewards[rows, idx] = rewards[rows, idx] + add.astype(f)
    vals = np.zeros_like(rewards) if values is None else values.astype(f)   # :141-144
    adv, ret = gae_padded_masked(rewards, vals, lm, no_eos, discount, gae_lambda)
    return dict(advantages=adv, returns=ret, kl_rewards=kl_rewards, tot_rewards=rewards, loss_mask=lm, logprobs=old,
                ref_logp=ref)


def packed_rewards(kl_ctl, clip_reward_value, log_probs, ref_log_probs, score, short1cu_seqlens, seq_no_eos_mask=None,
                   mask_no_eos_with_zero=False, dense=False, reward_delta=False):
    """get_packed_rewards / get_packed_reward_dense (realhf/impl/model/utils/ppo_functional.py:228-289)."""
    f = np.float32
    cu = np.asarray(short1cu_seqlens, dtype=np.int64)
    tot = (f(-kl_ctl) * (log_probs.astype(f) - ref_log_probs.astype(f))).astype(f)      # :241 / :267
    kl = tot.copy()
    sc = np.clip(score.astype(f), f(-clip_reward_value), f(clip_reward_value))         # :243 / :269
    if not dense:
        add = np.where(seq_no_eos_mask, f(0), sc) if mask_no_eos_with_zero else sc      # :245-250
        idx = cu[1:] - 1
        tot[idx] = tot[idx] + add.astype(f)
        return kl, tot
    off = off2 = 0
    for n in (cu[1:] - cu[:-1]):                                                        # :272-288
        n = int(n)
        if reward_delta:
            tot[off:off + n] = tot[off:off + n] + (sc[off2 + 1:off2 + n + 1] - sc[off2:off2 + n]).astype(f)
        else:
            tot[off:off + n] = tot[off:off + n] + sc[off2 + 1:off2 + n + 1]
        off += n
        off2 += n + 1
    return kl, tot


# --------------------------------------------------------------------------------------------------
# legacy running value statistics (realhf/impl/model/modules/rms.py)
# --------------------------------------------------------------------------------------------------
class RunningMeanStd:
    """ExponentialRunningMeanStd (rms.py:19-109, exponential=True) / MovingAverageRunningMeanStd
    (:112-183) in numpy, module dtype float64 (high_precision) or float32."""

    def __init__(self, exponential=True, beta=0.999, epsilon=1e-5, high_precision=True):
        self.dt = np.float64 if high_precision else np.float32
        self.exp, self.beta, self.eps = exponential, self.dt(beta), self.dt(epsilon)
        self.mean = self.dt(0)
        self.mean_sq = self.dt(0)
        self.aux = self.dt(0)          # debiasing term / accumulated denominator

    def update(self, x, mask=None):
        dt = self.dt
        x = np.asarray(x).astype(dt)
        if mask is None:
            factor = dt(x.size)                                            # :57-58
        else:
            m = np.asarray(mask).astype(dt)
            x = x * m                                                      # :60
            factor = m.sum(dtype=dt)
        s, sq = x.sum(dtype=dt), np.square(x).sum(dtype=dt)                # :63-64
        if self.exp:
            one = dt(1.0)
            self.mean = self.beta * self.mean + (s / factor) * (one - self.beta)        # :77-79
            self.mean_sq = self.beta * self.mean_sq + (sq / factor) * (one - self.beta)  # :80-82
            self.aux = self.beta * self.aux + one - self.beta                            # :83-85
        else:
            self.mean = (self.aux * self.mean + s) / (self.aux + factor)                # :164-166
            self.mean_sq = (self.aux * self.mean_sq + sq) / (self.aux + factor)          # :167-169
            self.aux = self.aux + factor

    def mean_std(self):
        dt = self.dt
        if self.exp:
            d = max(self.aux, self.eps)                                    # :88-89
            mean, msq = self.mean / d, self.mean_sq / d
        else:
            mean, msq = self.mean, self.mean_sq
        return dt(mean), dt(np.sqrt(max(msq - mean * mean, dt(1e-2))))      # :90-91 / :175-177

    def normalize(self, x):
        mean, std = self.mean_std()
        return np.clip((np.asarray(x).astype(self.dt) - mean) / std, -5, 5).astype(np.float32)   # :95-101

    def denormalize(self, x):
        mean, std = self.mean_std()
        return (np.asarray(x).astype(self.dt) * std + mean).astype(np.float32)                   # :104-108


# --------------------------------------------------------------------------------------------------
# AEnt clamped entropy (recipe/AEnt/functional.py:15-52)
# --------------------------------------------------------------------------------------------------
def clamped_weights(logits_row, k: int):
    """Weight of each element in the clamped softmax: 0 for the k smallest, 1 for the rest; the elements
    equal to the k-th smallest value share the kept count equally (top-k's tie-break is unspecified,
    recipe/AEnt/functional.py:22)."""
    x = np.asarray(logits_row, dtype=np.float64)
    w = np.ones_like(x)
    if k <= 0:
        return w
    tau = np.partition(x, k - 1)[k - 1]
    less = x < tau
    ties = x == tau
    removed_at_tau = k - int(less.sum())
    w[less] = 0.0
    w[ties] = (int(ties.sum()) - removed_at_tau) / int(ties.sum())
    return w


def gather_logprobs_clamped_entropy(logits, labels, entropy_clamp: float, temperature: float = 1.0):
    """(log-prob of the labels under the full softmax, clamped entropy) in float64 (:39-52); the masked
    elements contribute p*0 to the entropy sum (:30-33)."""
    z = np.asarray(logits, dtype=np.float64) / temperature
    T, V = z.shape
    k = int(V * entropy_clamp)                                             # :21
    m = z.max(-1, keepdims=True)
    lse = m[:, 0] + np.log(np.exp(z - m).sum(-1))
    logp = z[np.arange(T), labels] - lse
    ent = np.empty(T)
    for t in range(T):
        w = clamped_weights(z[t], k)
        e = w * np.exp(z[t] - m[t])
        q = e / e.sum()
        lq = (z[t] - m[t]) - np.log(e.sum())
        ent[t] = -(q * np.where(w > 0, lq, 0.0)).sum()
    return logp, ent


def clamped_entropy_dlogits(logits, labels, entropy_clamp, temperature, g_logp, g_ent):
    """d(sum g_logp*logp + g_ent*H_clamped)/dlogits in float64, ties sharing their weight."""
    z = np.asarray(logits, dtype=np.float64) / temperature
    T, V = z.shape
    k = int(V * entropy_clamp)
    out = np.zeros_like(z)
    for t in range(T):
        m = z[t].max()
        p = np.exp(z[t] - m)
        p /= p.sum()
        g = -g_logp[t] * p
        g[labels[t]] += g_logp[t]
        w = clamped_weights(z[t], k)
        e = w * np.exp(z[t] - m)
        q = e / e.sum()
        lq = (z[t] - m) - np.log(e.sum())
        h = -(q * np.where(w > 0, lq, 0.0)).sum()
        g += g_ent[t] * np.where(w > 0, -q * (lq + h), 0.0)
        out[t] = g / temperature
    return out
