#!/usr/bin/env python
"""Headline benchmark of the RL post-processing hot path (BASELINE.json `metric`).

One STEP = what one PPO/GRPO micro-batch costs on this path, for a packed batch of T tokens:
  (1) packed GAE over the batch                                  (areal_b200_gae1d_nolp_misalign)
  (2) `loss = loss_fn(logits, mb)` -- log-prob + entropy + decoupled PPO actor loss + the tracker's
      statistics from [T, V] bf16 logits                          (areal_b200.actor.grpo_loss_fn)
  (3) `loss *= loss_scale; loss.backward()` -> dlogits [T, V] bf16
(2) and (3) run inside a stand-in for the reference engine's `train_batch` loop that repeats its
arithmetic line by line (areal/engine/fsdp_engine.py:518-526, 571-586) around pre-materialised logits, so
the step goes through the same calls the unmodified engine makes.

Modes (`--mode`):
  hook     (default) `areal_b200.engine_hooks.install_loss_scale_hook(engine)` is installed, the scale is
           announced before the forward and (2)+(3) are ONE kernel that reads the logits once
           (areal_b200_fused_actor_fwd_bwd, 4TV bytes);
  twopass  nothing installed beyond the drop-in functions: forward kernel, then backward kernel (6TV bytes);
  kernel   the round-1 line: `fused_actor_loss(..., loss_scale=s)` called directly, no engine loop, no
           statistics launch.
All three are timed in every run; `value` is the `--mode` one, the others sit beside it.

N = 1 workload: BASELINE.json configs[1] (Qwen2.5-1.5B GRPO math shape, T = 16,384 packed tokens,
V = 151,936).  N > 1 (torchrun, one rank per GPU): the batch is sharded by packed sequence, every rank
holds its own T tokens (weak scaling); the exchanges are the engine's all-reduce of the loss weight and
the hook's copy of it.  Beside the headline the line carries clocked blocks for the other BASELINE
configs: `gae_sweep`, `intervals` (configs[4]), `cfg2_step` (configs[2], one GPU's share), `kernels`
(padded GAE, advantages, AEnt, LM head) at N = 1; `multi_gpu_parity`, `vocab_sharded` (configs[3]) and
`cfg2_strong` (configs[2] split over N) at N > 1.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--mode hook|twopass|kernel]

`--impl reference` times the reference's own CPU path (oracle port, see oracle/cpu_baseline.py) on the
box's host cores on a bounded sample of the same workload.
"""
import argparse
import ctypes
import functools
import json
import math
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "tokens/s fused logprob+entropy+PPO loss and packed GAE; % of HBM roofline"
T_TOKENS = 16384
VOCAB = 151936
CFG = dict(eps_clip=0.4, eps_clip_higher=None, c_clip=None, behav_imp_weight_cap=5.0)   # examples/math/gsm8k_grpo.yaml:44,57-65
TEMPERATURE = 1.0
GAMMA, LAM = 1.0, 0.95


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while a timed region runs."""

    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples = []
        self.stop_flag = False
        self.max_mhz = None
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag:
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((time.perf_counter(), mhz, reasons))
            except Exception:
                pass
            time.sleep(self.period)

    def summary(self, t0, t1):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"], "samples": 0}
        inside = [s for s in self.samples if t0 <= s[0] <= t1] or self.samples
        mhz = sorted(s[1] for s in inside)
        bits = 0
        for s in inside:
            bits |= s[2]
        names = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
                 0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
                 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}
        reasons = [n for b, n in names.items() if bits & b and n != "gpu_idle"]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(inside)}


class Clocked:
    """`with Clocked(index) as c: ...` then `c.clocks`: the ClockSampler summary of the block."""

    def __init__(self, index):
        self.sampler = ClockSampler(index, period=float(os.environ.get("BENCH_CLOCK_PERIOD", "0.02")))
        self.clocks = None

    def __enter__(self):
        if os.environ.get("BENCH_NO_SAMPLER") != "1":
            self.sampler.start()
        self.t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        self.sampler.stop_flag = True
        self.clocks = self.sampler.summary(self.t0, time.perf_counter())
        return False


def packed_lengths(T, rng, lo=128, hi=4097):
    lens, left = [], T
    while left > 0:
        n = min(left, int(rng.randint(lo, hi)))
        lens.append(n)
        left -= n
    return lens


def fill_logits(torch, out, gen, scale=2.0, rows=2048):
    T, V = out.shape
    for i in range(0, T, rows):
        n = min(rows, T - i)
        out[i:i + n] = (torch.randn(n, V, generator=gen, device=out.device) * scale).bfloat16()
    return out


def make_batch(torch, dev, T, V, seed):
    """Synthetic packed batch of the cfg-2 shape: sequences of U[128,4096] tokens packed to
    T, prompt prefix (5-25 % of each sequence) masked out, logits = randn * 2 in bf16."""
    import numpy as np
    rng = np.random.RandomState(seed)
    lens = packed_lengths(T, rng)
    B = len(lens)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    mask = np.ones(T, dtype=bool)
    for i in range(B):
        mask[cu[i]:cu[i] + int(lens[i] * rng.uniform(0.05, 0.25))] = False
    g = torch.Generator(device=dev).manual_seed(seed)
    logits = fill_logits(torch, torch.empty(T, V, dtype=torch.bfloat16, device=dev), g)
    labels = torch.randint(0, V, (T,), generator=g, device=dev)
    batch = dict(
        logits=logits, labels=labels,
        loss_mask=torch.from_numpy(mask).to(dev),
        advantages=torch.randn(T, generator=g, device=dev),
        rewards=torch.randn(T, generator=g, device=dev),
        values=torch.randn(T + B, generator=g, device=dev),
        cu_seqlens=torch.from_numpy(cu).to(dev),
        bootstrap=torch.from_numpy(rng.rand(B) > 0.5).to(dev),
        n_seqs=B, lens=lens,
    )
    return batch


class EngineLoop:
    """Stand-in for the reference train engine around PRE-MATERIALISED logits: `train_batch` repeats
    areal/engine/fsdp_engine.py:518-526 (total loss weight, all-reduced over the data-parallel group) and
    :571-586 (loss_fn, loss_scale, `loss *= loss_scale`, backward) line by line; the model forward is
    replaced by handing out the logits leaf.  The engine's `assert total_loss_weight != 0` (a host sync
    that belongs to the engine, not to this path) is left out."""

    def __init__(self, torch, dist, dev, world):
        self.torch, self.dist = torch, dist
        self.device = dev
        self.dp_group = dist.group.WORLD if world > 1 else None
        import types
        self.parallel_helper = types.SimpleNamespace(dp_size=world, sp_size=1)
        self.after_loss_fn = None

    def train_batch(self, input_, loss_fn, loss_weight_fn):
        torch = self.torch
        total_loss_weight = (torch.stack([loss_weight_fn(mb) for mb in input_]).sum().detach().clone()
                             .to(dtype=torch.float32, device=self.device))                         # :518-524
        if self.parallel_helper.dp_size > 1:
            self.dist.all_reduce(total_loss_weight, group=self.dp_group)                            # :526
        loss = None
        for mb in input_:
            logits = mb["_logits"]                                                                 # :571-573
            loss = loss_fn(logits, mb)                                                             # :578
            if self.after_loss_fn is not None:
                self.after_loss_fn()
            loss_scale = loss_weight_fn(mb) / total_loss_weight                                    # :580
            loss_scale *= self.parallel_helper.dp_size                                             # :583
            loss *= loss_scale                                                                     # :585
            loss.backward()                                                                        # :586
        return loss


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from areal_b200 import _lib
    from areal_b200 import actor as A
    from areal_b200 import engine_hooks as EH
    from areal_b200 import functional as F
    from areal_b200 import ppo_functional as P
    from areal_b200 import stats_tracker as ST

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    _lib.lib()   # fail loudly if the CUDA extension is missing
    peak, peak_src = measured_peaks()
    only = set(args.only.split(",")) if args.only else None

    def want(name):
        return only is None or name in only

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    T, V = T_TOKENS, VOCAB
    b = make_batch(torch, dev, T, V, seed=rank)
    # proximal / behaviour log-probs around the true ones
    with torch.no_grad():
        lp_true = F.gather_logprobs(b["logits"], b["labels"], TEMPERATURE)
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        b["prox_logp"] = lp_true + 0.02 * torch.randn(T, generator=g, device=dev)
        b["old_logp"] = lp_true + 0.05 * torch.randn(T, generator=g, device=dev)
    logits = b["logits"].requires_grad_()
    kernel_scale = torch.full((1,), 0.5, dtype=torch.float32, device=dev)   # `kernel` mode: the announced scalar
    weight_fn = lambda x: x["loss_mask"].count_nonzero()                     # areal/engine/ppo/actor.py:269
    mb = dict(_logits=logits, input_ids=b["labels"], rolled_input_ids=b["labels"], logprobs=b["old_logp"],
              prox_logp=b["prox_logp"], advantages=b["advantages"], loss_mask=b["loss_mask"])

    # ---- event taps around OUR dominant kernels (no torch kernel launches in between: the wrappers only
    # allocate outputs before the C-ABI call), recorded on the launch stream inside the timed region
    taps = {"on": False, "fwd": [], "bwd": []}
    real_fused, real_bwd = F.fused_actor_loss, F._logits_backward

    def tapped(real, key):
        @functools.wraps(real)
        def wrapper(*a, **k):
            if not taps["on"]:
                return real(*a, **k)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = real(*a, **k)
            e1.record()
            taps[key].append((e0, e1))
            return out
        return wrapper
    F.fused_actor_loss = tapped(real_fused, "fwd")
    F._logits_backward = tapped(real_bwd, "bwd")

    def make_step(mode):
        tracker = ST.DistributedStatsTracker()
        engine = EngineLoop(torch, dist, dev, world)
        if mode == "hook":
            EH.install_loss_scale_hook(engine)
        loss_fn = functools.partial(A.grpo_loss_fn, temperature=TEMPERATURE, tracker=tracker, **CFG)

        def step(ev=None):
            if ev is not None:
                ev[0].record()
            adv, ret = P.cugae1d_nolp_misalign_func(b["rewards"], b["values"], b["cu_seqlens"], b["bootstrap"], GAMMA, LAM)
            if ev is not None:
                ev[1].record()
            logits.grad = None
            if mode == "kernel":
                loss, st = F.fused_actor_loss(logits, b["labels"], TEMPERATURE, b["prox_logp"], b["old_logp"],
                                              b["advantages"], b["loss_mask"], loss_scale=kernel_scale, **CFG)
                if ev is not None:
                    ev[2].record()
                (loss * kernel_scale[0]).backward()
            else:
                engine.after_loss_fn = (lambda: ev[2].record()) if ev is not None else None
                loss = engine.train_batch([mb], loss_fn, weight_fn)
            if ev is not None:
                ev[3].record()
            return loss
        return step, tracker, engine

    def timed_loop(mode, steps, warmup, main):
        """W warm-up steps, then `steps` timed steps between barrier+synchronize, CUDA events on the launch
        stream; returns per-step time (max over ranks), per-segment spreads, kernel taps, host enqueue time,
        clocks."""
        step, tracker, engine = make_step(mode)
        pool = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(steps)]
        for _ in range(warmup):
            step()
        tracker.export()                       # drop the warm-up's totals (one host sync, outside the timing)
        taps["fwd"].clear()
        taps["bwd"].clear()
        clk = Clocked(local_rank)
        sync_all()
        with clk:
            e_start = torch.cuda.Event(enable_timing=True)
            e_stop = torch.cuda.Event(enable_timing=True)
            if main:
                torch.cuda.profiler.start()      # ncu --profile-from-start off captures exactly the timed region
            taps["on"] = True
            e_start.record()
            h0 = time.perf_counter()
            for i in range(steps):
                loss = step(pool[i])
            e_stop.record()
            host_ms = (time.perf_counter() - h0) * 1e3 / steps     # CPU enqueue time per step
            sync_all()
            taps["on"] = False
            if main:
                torch.cuda.profiler.stop()
        ms_total = e_start.elapsed_time(e_stop)
        if world > 1:
            mt = torch.tensor([ms_total], device=dev)
            dist.all_reduce(mt, op=dist.ReduceOp.MAX)
            ms_total = float(mt.item())

        def spread(v):
            v = sorted(v)
            return {"min": v[0], "median": v[len(v) // 2], "mean": float(np.mean(v)), "max": v[-1]} if v else None
        out = {"ms_step": ms_total / steps, "host_ms": host_ms, "clocks": clk.clocks, "loss": float(loss.detach())}
        out["spreads"] = {"gae": spread([e[0].elapsed_time(e[1]) for e in pool]),
                          "loss_fn": spread([e[1].elapsed_time(e[2]) for e in pool]),
                          "scale_and_backward": spread([e[2].elapsed_time(e[3]) for e in pool])}
        out["kernel_fwd"] = spread([a.elapsed_time(c) for a, c in taps["fwd"]])
        out["kernel_bwd"] = spread([a.elapsed_time(c) for a, c in taps["bwd"]])
        stats = tracker.export(reduce_group=dist.group.WORLD if world > 1 else None)
        out["tracker"] = {k: stats[k] for k in ("actor_loss/avg", "entropy/avg", "importance_weight/avg", "n_valid_tokens",
                                                "clip_ratio/avg") if k in stats}
        ann = getattr(engine.train_batch, "last_announcer", None)
        out["announced"] = None if ann is None else int(ann.n_announced)
        return out

    def graph_replay(steps):
        """The `kernel`-mode step captured once in a CUDA graph and replayed: what a trainer that graphs its
        post-processing pays per step (no per-launch host work).  N=1 only; reported beside `value`."""
        step, _, _ = make_step("kernel")
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3):
                step()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            loss = step()
        for _ in range(3):
            gr.replay()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            gr.replay()
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / steps
        return {"ms_per_step": ms, "tokens_per_s": T / (ms * 1e-3), "steps": steps, "loss": float(loss.detach())}

    W = max(args.warmup, 3)
    results = {}
    if want("headline"):
        results[args.mode] = timed_loop(args.mode, args.steps, W, True)
        for m in ("hook", "twopass", "kernel"):
            if m not in results:
                results[m] = timed_loop(m, max(3, min(args.steps, 20)), 3, False)
    line = {
        "metric": METRIC, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": W,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "accumulate_dtype": "f32 (GAE scan f64)", "data": "synthetic", "mode": args.mode,
    }
    if want("headline"):
        main = results[args.mode]
        ms_step = main["ms_step"]
        coef_rows = int(torch.count_nonzero(logits.grad[:, :8].float().abs().sum(-1)).item())
        active = coef_rows / T

        def roof(kernel, bytes_, ms, **kw):
            gbs = bytes_ / (ms * 1e-3) / 1e9
            # `traffic` = DRAM bytes of one launch from an `ncu --set full` capture of THIS kernel at this shape;
            # profiles/ncu_traffic.json is regenerated with every capture (profiles/README.md), never pasted
            tr = ncu_traffic(kernel, T, V)
            return dict({"bound": "hbm", "kernel": kernel, "achieved": gbs, "peak": peak, "unit": "GB/s",
                         "frac": gbs / peak, "traffic": tr[0] if tr else None,
                         "traffic_source": tr[1] if tr else None, "peak_source": peak_src, "ms": ms,
                         "algorithmic_bytes": bytes_}, **kw)

        def mode_block(name, r):
            d = {"ms_per_step": r["ms_step"], "tokens_per_s": world * T / (r["ms_step"] * 1e-3),
                 "host_enqueue_ms_per_step": r["host_ms"], "segments_ms": r["spreads"], "loss": r["loss"],
                 "tracker": r["tracker"]}
            if name in ("hook", "kernel"):
                d["roofline"] = roof("fused_rowres_kernel", 4.0 * T * V + 64.0 * T, r["kernel_fwd"]["median"],
                                     kernel_ms_spread=r["kernel_fwd"],
                                     note="one read of the logits (2TV) + dlogits written (2TV); the two-pass split moves 6TV")
                if name == "hook":
                    d["scales_announced_per_run"] = r["announced"]
            else:
                d["fwd"] = roof("logits_fwd_tma_kernel", 2.0 * T * V + 64.0 * T, r["kernel_fwd"]["median"],
                                kernel_ms_spread=r["kernel_fwd"])
                d["bwd"] = roof("logits_bwd_tma_kernel", 2.0 * T * V * (1.0 + active), r["kernel_bwd"]["median"],
                                active_row_frac=active, kernel_ms_spread=r["kernel_bwd"],
                                note="bytes = 2TV written + 2TV*active read (rows whose gradient is zero are not read); "
                                     "the reference's split API needs 4TV")
            return d
        blocks = {m: mode_block(m, r) for m, r in results.items()}
        api = {"hook": "areal_b200.actor.grpo_loss_fn called by the engine loop with "
                       "areal_b200.engine_hooks.install_loss_scale_hook installed (unmodified engine, single pass)",
               "twopass": "areal_b200.actor.grpo_loss_fn called by the engine loop, nothing else installed (two passes)",
               "kernel": "areal_b200.functional.fused_actor_loss(loss_scale=s) called directly (no engine loop, no statistics)"}
        line.update({
            "value": world * T / (ms_step * 1e-3), "ms_per_step": ms_step,
            "config": {"workload": "BASELINE.json configs[1]: Qwen2.5-1.5B GRPO math shape, T=16384 packed tokens "
                                   "per GPU, V=151936 bf16 logits; step = packed GAE + loss_fn(logits, mb) "
                                   "[logprob+entropy+decoupled PPO loss+tracker statistics] + loss*=loss_scale + "
                                   "backward to dlogits, inside the reference engine's train_batch arithmetic; API: "
                                   + api[args.mode],
                       "api": api[args.mode], "tokens_per_gpu": T, "vocab": V,
                       "parallelism": f"dp{world} (batch-sharded by packed sequence)",
                       "l2": "inputs (4.98 GB logits per GPU) larger than the 126 MB L2; no flush needed",
                       "loss_cfg": {k: v for k, v in CFG.items()}, "temperature": TEMPERATURE},
            "roofline": blocks[args.mode].get("roofline") or blocks[args.mode]["bwd"],
            "engine_hook_single_pass": blocks["hook"], "two_pass": blocks["twopass"], "kernel_only_single_pass": blocks["kernel"],
            "roofline_gae": {"bound": "launch-latency", "kernel": "gae1d_prep_kernel + gae1d_kernel",
                             "ms": main["spreads"]["gae"]["median"],
                             "achieved": (16.0 * T + 9.0 * b["n_seqs"] + 4) / (main["spreads"]["gae"]["median"] * 1e-3) / 1e9,
                             "unit": "GB/s", "algorithmic_bytes": 16.0 * T + 9.0 * b["n_seqs"] + 4,
                             "note": "eager call incl. the wrapper's host time; GPU-time figures are in gae_sweep"},
            # our kernels per step: gae1d_prep + gae1d + masked_stats + (fused_rowres + rescale_if_needed |
            # logits_fwd + logits_bwd); `kernel` mode has no masked_stats
            "gpu_launches": {"hook": 5, "twopass": 5, "kernel": 4}[args.mode] * args.steps,
            "clocks": main["clocks"], "host_enqueue_ms_per_step": main["host_ms"], "loss": main["loss"],
        })
        if world == 1 and os.environ.get("BENCH_NO_GRAPH") != "1":
            try:
                line["graph_replay_kernel_mode"] = graph_replay(max(3, min(args.steps, 30)))
            except Exception as exc:            # reported, never fatal: the eager loop above is the headline
                line["graph_replay_kernel_mode"] = {"error": str(exc)[:200]}

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timed region
    if want("e2e") and not args.skip_e2e:
        try:
            line["e2e"] = run_e2e(torch, dist, dev, b, world, max(3, min(args.steps, 5)), args.mode)
        except Exception as ex:  # pinned allocation can fail on a small host: say so instead of faking
            line["e2e"] = {"value": None, "unit": "tokens/s", "error": repr(ex)[:300]}

    # free the headline batch before the side blocks (the interval buffer alone is 15.2 GB)
    F.fused_actor_loss, F._logits_backward = real_fused, real_bwd
    logits.grad = None
    del logits, mb
    b.clear()
    torch.cuda.empty_cache()

    from bench_blocks import Blocks
    blk = Blocks(torch, dist, dev, world, rank, local_rank, peak, Clocked, EngineLoop)
    if world == 1:
        for name in ("e2e_from_hidden", "gae_sweep", "intervals", "cfg2_step", "kernels"):
            if want(name) and not args.skip_blocks:
                line[name] = blk.run(name)
    else:
        for name in ("multi_gpu_parity", "vocab_sharded", "cfg2_strong"):
            if want(name) and not args.skip_blocks:
                line[name] = blk.run(name)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if world == 1 and not args.skip_cpu and want("cpu"):
        line["cpu_baseline"] = cpu_baseline()
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def ncu_traffic(kernel, T, V):
    """(dram bytes per launch, capture file) for `kernel` at shape (T, V) from profiles/ncu_traffic.json -- written by
    profiles/ncu_summary.py from the `ncu --set full` capture of the same commit -- or None."""
    try:
        tab = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        e = tab.get(f"{kernel}:{T}x{V}")
        return (float(e["dram_bytes"]), e["source"]) if e else None
    except Exception:
        return None


def host_mem_policy(mode, local_rank):
    """Experiment knob BENCH_PIN_POLICY for the pinned host buffers of the e2e leg: "interleave" spreads the pages over
    all NUMA nodes of the host, "local" binds them to the node the GPU hangs off (sysfs), anything else leaves the
    default (first touch).  Returns a description for the JSON line.  set_mempolicy(2) through libc's syscall()."""
    import ctypes
    import glob
    nodes = sorted(int(p.rsplit("node", 1)[1]) for p in glob.glob("/sys/devices/system/node/node[0-9]*"))
    info = {"policy": mode or "default", "numa_nodes": len(nodes)}
    if not mode or mode == "default" or len(nodes) < 2:
        return info
    try:
        libc = ctypes.CDLL(None, use_errno=True)
        mask = ctypes.c_ulong(0)
        if mode == "interleave":
            for n in nodes:
                mask.value |= 1 << n
            pol = 3                                    # MPOL_INTERLEAVE
        else:
            import torch
            pr = torch.cuda.get_device_properties(local_rank)
            node = 0
            try:
                bdf = "%04x:%02x:%02x.0" % (getattr(pr, "pci_domain_id", 0), pr.pci_bus_id, pr.pci_device_id)
                node = max(0, int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read()))
                info["bdf"] = bdf
            except Exception as ex:
                info["bdf_error"] = repr(ex)[:80]
            mask.value = 1 << node
            pol = 2                                    # MPOL_BIND
            info["node"] = node
        rc = libc.syscall(238, pol, ctypes.byref(mask), ctypes.c_ulong(64))   # x86_64: __NR_set_mempolicy
        info["rc"] = int(rc)
        if rc != 0:
            info["errno"] = ctypes.get_errno()
    except Exception as ex:
        info["error"] = repr(ex)[:120]
    return info


def run_e2e(torch, dist, dev, b, world, steps, mode):
    """Same step, but the inputs start in pinned HOST memory every step and the results are read back:
    H2D of logits + per-token vectors, GAE + loss_fn + backward on the device (through the engine loop of
    `mode`), D2H of the loss and the per-token advantages / returns.  Also measured, the same way and on all
    ranks at once: the bare pinned host->device copy of the same bytes (`h2d_ceiling_gbs`), which is what
    bounds this number."""
    from areal_b200 import ppo_functional as P
    T, V = b["logits"].shape
    keys = ("logits", "labels", "loss_mask", "advantages", "rewards", "values", "cu_seqlens", "bootstrap",
            "prox_logp", "old_logp")
    pin_policy = host_mem_policy(os.environ.get("BENCH_PIN_POLICY", ""), dev.index or 0)
    host = {}
    for k in keys:
        t = b[k].detach()
        h = torch.empty(t.shape, dtype=t.dtype, device="cpu", pin_memory=True)
        h.copy_(t)
        host[k] = h
    # two device-side input sets: step i's host->device copies run on a copy stream while step i-1
    # computes (what a trainer streaming micro-batches from host memory does); every byte of every
    # step is still copied inside the timed region
    dst2 = [{k: torch.empty_like(b[k].detach()) for k in host} for _ in range(2)]
    out_host = {k: torch.empty(T, dtype=torch.float32, pin_memory=True) for k in ("adv", "ret")}
    loss_host = torch.empty((), dtype=torch.float32, pin_memory=True)
    h2d = sum(h.numel() * h.element_size() for h in host.values())
    d2h = 4 + sum(h.numel() * 4 for h in out_host.values())
    copy_stream = torch.cuda.Stream(dev)
    main_stream = torch.cuda.current_stream(dev)
    copied = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    for e in consumed:
        e.record(main_stream)
    chunks = max(1, int(os.environ.get("BENCH_H2D_CHUNKS", "4")))

    def copy_in(dst, k):
        """one tensor host->device; the 4.98 GB logits go as `chunks` row blocks (several copies in flight)"""
        if k == "logits" and chunks > 1:
            n = host[k].shape[0]
            step_ = (n + chunks - 1) // chunks
            for i in range(0, n, step_):
                dst[k][i:i + step_].copy_(host[k][i:i + step_], non_blocking=True)
        else:
            dst[k].copy_(host[k], non_blocking=True)

    def upload(i):
        s = i & 1
        copy_stream.wait_event(consumed[s])           # the compute that last read this set is done
        with torch.cuda.stream(copy_stream):
            for k in host:
                copy_in(dst2[s], k)
            copied[s].record(copy_stream)

    import functools as ft
    from areal_b200 import actor as A, engine_hooks as EH, functional as F, stats_tracker as ST
    tracker = ST.DistributedStatsTracker()
    engine = EngineLoop(torch, dist, dev, world)
    if mode == "hook":
        EH.install_loss_scale_hook(engine)
    loss_fn = ft.partial(A.grpo_loss_fn, temperature=TEMPERATURE, tracker=tracker, **CFG)
    weight_fn = lambda x: x["loss_mask"].count_nonzero()
    kscale = torch.full((1,), 0.5, dtype=torch.float32, device=dev)

    def compute(i):
        s = i & 1
        dst = dst2[s]
        main_stream.wait_event(copied[s])
        x = dst["logits"].requires_grad_()
        adv, ret = P.cugae1d_nolp_misalign_func(dst["rewards"], dst["values"], dst["cu_seqlens"], dst["bootstrap"], GAMMA, LAM)
        if mode == "kernel":
            loss, _ = F.fused_actor_loss(x, dst["labels"], TEMPERATURE, dst["prox_logp"], dst["old_logp"],
                                         dst["advantages"], dst["loss_mask"], loss_scale=kscale, **CFG)
            (loss * kscale[0]).backward()
        else:
            m = dict(_logits=x, input_ids=dst["labels"], rolled_input_ids=dst["labels"], logprobs=dst["old_logp"],
                     prox_logp=dst["prox_logp"], advantages=dst["advantages"], loss_mask=dst["loss_mask"])
            loss = engine.train_batch([m], loss_fn, weight_fn)
        loss_host.copy_(loss.detach(), non_blocking=True)
        out_host["adv"].copy_(adv, non_blocking=True)
        out_host["ret"].copy_(ret, non_blocking=True)
        x.grad = None
        dst["logits"] = x.detach()
        consumed[s].record(main_stream)

    def run(n):
        upload(0)
        for i in range(n):
            if i + 1 < n:
                upload(i + 1)
            compute(i)

    def timed(fn, n):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        fn(n)
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / n
        if world > 1:
            mt = torch.tensor([ms], device=dev)
            dist.all_reduce(mt, op=dist.ReduceOp.MAX)
            ms = float(mt.item())
        return ms

    def bare_copies(n):
        for i in range(n):
            for k in host:
                copy_in(dst2[i & 1], k)

    run(2)
    ms = timed(run, steps)
    tracker.export()
    ms_copy = timed(bare_copies, steps)
    ceiling = h2d / (ms_copy * 1e-3) / 1e9
    achieved = h2d / (ms * 1e-3) / 1e9
    return {"value": world * T / (ms * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "ms_per_step": ms, "steps": steps, "mode": mode,
            "h2d_gbs_per_gpu": achieved, "h2d_ceiling_gbs_per_gpu": ceiling,
            "h2d_ceiling_gbs_all_gpus": ceiling * world, "e2e_frac_of_ceiling": achieved / ceiling,
            "h2d_chunks_in_flight": chunks, "host_mem_policy": pin_policy,
            "note": "bounded by the 4.98 GB pinned host->device copy of the logits per GPU and step; `h2d_ceiling_*` is "
                    "the same bytes copied with nothing else running, on all ranks at once (max over ranks); copies of "
                    "step i+1 overlap the kernels of step i (two device-side input sets, copy stream)"}


def cpu_baseline(sample_rows=1024, steps=2, warmup=1):
    """The oracle port of the reference's CPU path, timed on this box's host cores on a bounded sample, plus the
    two CPU legs BASELINE.md section 5 names on their own: configs[0] (pygae1d_nolp_misalign's per-token torch
    loop) and the interval copies in Python."""
    from oracle import cpu_baseline as CB
    threads = os.cpu_count() or 1
    dt, used = CB.time_steps(sample_rows, VOCAB, steps, warmup, CFG, TEMPERATURE, threads=threads)
    out = {"value": sample_rows / dt, "unit": "tokens/s", "cores": used, "kind": "port",
           "sample": f"{sample_rows} tokens x V={VOCAB} bf16 (1/{T_TOKENS // sample_rows} of the workload), "
                     f"{steps} steps after {warmup} warm-up; torch CPU eager log_softmax/entropy/PPO loss + backward "
                     f"on {used} threads, GAE as the reference's per-token loop of 0-dim torch ops (1 thread)",
           "seconds_per_step": dt}
    try:
        nbytes, t_slice, t_set, thr = CB.time_intervals_cpu(100_000_000, 10_000, threads=threads)
        out["intervals_cpu"] = {"slice_gbs": nbytes / t_slice / 1e9, "set_gbs": nbytes / t_set / 1e9, "cores": thr,
                                "sample": "_slice_intervals_py / _set_intervals_py (flatten_param.py:65-90) restated, 1e8 "
                                          "bf16 elements, 10,000 intervals of the reference test's pattern "
                                          "(test_interval_ops.py:20-31), 3 repetitions"}
        tokens, secs, nseq = CB.time_gae_cfg0(max_seconds=20.0)
        out["gae_cfg0"] = {"value": tokens / secs, "unit": "tokens/s", "cores": 1, "tokens": tokens, "seconds": secs,
                           "sample": f"BASELINE.json configs[0]: the first {nseq} of 64 sequences of U[128,4096] tokens "
                                     "(whole sequences that fit ~20 s), gamma=1.0, lam=0.95, pygae1d_nolp_misalign's "
                                     "loop of 0-dim torch ops (ppo_functional.py:292-323) restated, single Python thread"}
    except Exception as ex:
        out["extra_error"] = repr(ex)[:200]
    try:
        import torch
        if torch.cuda.is_available():
            # the same restated reference ops, eager, on this GPU (reported: SURVEY 8d "GPU reference baseline")
            torch.cuda.empty_cache()
            torch.cuda.reset_peak_memory_stats(0)
            secs, peak = CB.time_steps_on_gpu(T_TOKENS, VOCAB, 3, 1, CFG, TEMPERATURE)
            out["torch_gpu_port"] = {"value": T_TOKENS / secs, "unit": "tokens/s", "ms_per_step": secs * 1e3,
                                     "peak_memory_gb": peak / 1e9,
                                     "sample": f"the same restated torch ops (chunked fp32 log_softmax / entropy, PPO loss, "
                                               f"backward to bf16 dlogits; no GAE) run eagerly on cuda:0 at the full "
                                               f"T={T_TOKENS}, V={VOCAB}: 3 steps after 1 warm-up, CUDA events"}
            torch.cuda.empty_cache()
    except Exception as ex:
        out["torch_gpu_port_error"] = repr(ex)[:200]
    return out


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path (oracle port) on the host cores: `--steps`
    steps after `--warmup` warm-ups, each step a bounded sample (1,024 tokens) of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample_rows = 1024
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    # ~0.3-0.5 s per step on 16-32 host threads: the driver's --steps 20 --warmup 5 ends within ~15 s
    from oracle import cpu_baseline as CB
    threads = os.cpu_count() or 1
    dt, used = CB.time_steps(sample_rows, VOCAB, steps, warmup, CFG, TEMPERATURE, threads=threads)
    v = sample_rows / dt
    cb = {"value": v, "unit": "tokens/s", "cores": used, "kind": "port",
          "sample": f"each step = {sample_rows} tokens x V={VOCAB} bf16 (bounded sample of the {T_TOKENS}-token "
                    f"workload, a rate: tokens/s does not depend on the sample size); {steps} steps after {warmup} warm-up"}
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE.json configs[1] shape (V=151936 bf16), bounded sample of "
                                   f"{sample_rows} tokens per step on the host CPU; same step composition "
                                   "(packed GAE + logprob/entropy/PPO loss fwd + bwd)"},
            "cpu_baseline": cb,
            "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def emit(line: dict):
    """Print exactly one JSON line on the real stdout (fd 1 was pointed at stderr in main(), so NCCL /
    library chatter cannot pollute it)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="hook", choices=["hook", "twopass", "kernel"],
                    help="hook: engine loop + loss-scale hook (single pass); twopass: engine loop, forward then "
                         "backward; kernel: fused_actor_loss(loss_scale=s) called directly")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs: skip the host-buffer leg")
    ap.add_argument("--skip-cpu", action="store_true", help="profiling runs: skip the CPU baseline leg")
    ap.add_argument("--skip-blocks", action="store_true", help="profiling runs: skip the side blocks")
    ap.add_argument("--only", default=None, help="comma list of parts to run: headline,e2e,cpu,gae_sweep,intervals,"
                                                  "cfg2_step,kernels,multi_gpu_parity,vocab_sharded,cfg2_strong")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
