#!/usr/bin/env python
"""Headline benchmark of the RL post-processing hot path (BASELINE.json `metric`).

One STEP = what one PPO/GRPO micro-batch costs on this path, for a packed batch of T tokens:
  (1) packed GAE over the batch               (areal_b200_gae1d_nolp_misalign)
  (2) fused log-prob + entropy + decoupled PPO actor loss from [T, V] bf16 logits
  (3) (loss * loss_scale).backward() -> dlogits [T, V] bf16
called through the public Python API (`areal_b200.functional.fused_actor_loss`, autograd), i.e.
through the C ABI.  Default mode "single": the caller announces loss_scale, and (2)+(3) are ONE kernel
that reads the logits once (areal_b200_fused_actor_fwd_bwd, cluster-resident rows, 4TV bytes).
`--mode twopass` times the drop-in split instead (areal_b200_fused_actor_fwd then
areal_b200_logits_bwd, 6TV bytes); its per-kernel numbers are always reported under "two_pass".  N = 1 workload: BASELINE.json configs[1] (Qwen2.5-1.5B GRPO math shape,
T = 16,384 packed tokens, V = 151,936).  N > 1 (torchrun, one rank per GPU): the batch is sharded by
packed sequence, every rank holds its own T tokens (weak scaling), and the only exchange is one
NCCL all-reduce of (loss numerator, token count) per step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

`--impl reference` times the reference's own CPU path (oracle port, see oracle/cpu_baseline.py) on the
box's host cores on a bounded sample of the same workload.
"""
import argparse
import json
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
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

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


def make_batch(torch, dev, T, V, seed):
    """Synthetic packed batch of the cfg-2 shape: sequences of U[128,4096] tokens packed to
    T, prompt prefix (5-25 % of each sequence) masked out, logits = randn * 2 in bf16."""
    import numpy as np
    rng = np.random.RandomState(seed)
    lens, left = [], T
    while left > 0:
        n = min(left, int(rng.randint(128, 4097)))
        lens.append(n)
        left -= n
    B = len(lens)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    mask = np.ones(T, dtype=bool)
    for i in range(B):
        mask[cu[i]:cu[i] + int(lens[i] * rng.uniform(0.05, 0.25))] = False
    g = torch.Generator(device=dev).manual_seed(seed)
    logits = torch.empty(T, V, dtype=torch.bfloat16, device=dev)
    for i in range(0, T, 2048):
        n = min(2048, T - i)
        logits[i:i + n] = (torch.randn(n, V, generator=g, device=dev) * 2.0).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g, device=dev)
    batch = dict(
        logits=logits, labels=labels,
        loss_mask=torch.from_numpy(mask).to(dev),
        advantages=torch.randn(T, generator=g, device=dev),
        rewards=torch.randn(T, generator=g, device=dev),
        values=torch.randn(T + B, generator=g, device=dev),
        cu_seqlens=torch.from_numpy(cu).to(dev),
        bootstrap=torch.from_numpy(rng.rand(B) > 0.5).to(dev),
        n_seqs=B,
    )
    return batch


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from areal_b200 import functional as F
    from areal_b200 import ppo_functional as P
    from areal_b200 import _lib

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

    T, V = T_TOKENS, VOCAB
    b = make_batch(torch, dev, T, V, seed=rank)
    # proximal / behaviour log-probs around the true ones
    with torch.no_grad():
        lp_true = F.gather_logprobs(b["logits"], b["labels"], TEMPERATURE)
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        b["prox_logp"] = lp_true + 0.02 * torch.randn(T, generator=g, device=dev)
        b["old_logp"] = lp_true + 0.05 * torch.randn(T, generator=g, device=dev)
    logits = b["logits"].requires_grad_()
    loss_scale = torch.full((1,), 0.5, dtype=torch.float32, device=dev)   # the engine's loss_scale (fsdp_engine.py:578-586)
    red_pool = [torch.zeros(2, dtype=torch.float32, device=dev) for _ in range(8)]
    n_valid = float(b["loss_mask"].sum().item())
    cnt_vec = torch.tensor([n_valid, 0.0], device=dev)     # loss * count -> numerator
    cnt_vec_add = torch.tensor([0.0, n_valid], device=dev)

    def timed_loop(mode, steps, warmup, main):
        """W warm-up steps, then `steps` timed steps between barrier+synchronize, CUDA events on the launch
        stream; returns per-step time (max over ranks), per-segment spreads, host enqueue time, clocks."""
        ev = {k: [] for k in ("start", "gae", "fwd", "bwd")}
        pool = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(steps)]
        works = []
        out = {}

        def step(record=False):
            if record:
                e = pool[len(ev["start"])]
                e[0].record()
            P.cugae1d_nolp_misalign_func(b["rewards"], b["values"], b["cu_seqlens"], b["bootstrap"], GAMMA, LAM)
            if record:
                e[1].record()
            logits.grad = None
            loss, st = F.fused_actor_loss(logits, b["labels"], TEMPERATURE, b["prox_logp"], b["old_logp"],
                                          b["advantages"], b["loss_mask"],
                                          loss_scale=loss_scale if mode == "single" else None, **CFG)
            if record:
                e[2].record()
            (loss * loss_scale[0]).backward()
            if record:
                e[3].record()
                for k, x in zip(("start", "gae", "fwd", "bwd"), e):
                    ev[k].append(x)
            if world > 1:
                # the path's only exchange: (loss numerator, token count) -> global masked-mean loss.
                # Issued asynchronously on NCCL's stream so it overlaps the next step's kernels; every
                # handle is waited on before the timed region closes.
                buf = red_pool[len(works) % len(red_pool)]
                torch.addcmul(cnt_vec_add, cnt_vec, loss.detach(), out=buf)   # [loss*count, count]
                works.append(dist.all_reduce(buf, async_op=True))
            return loss, st

        for _ in range(warmup):
            step()
        for w in works:
            w.wait()
        works.clear()
        sampler = ClockSampler(local_rank, period=float(os.environ.get('BENCH_CLOCK_PERIOD', '0.02')))
        if main and os.environ.get('BENCH_NO_SAMPLER') != '1':
            sampler.start()
        sync_all()
        t0 = time.perf_counter()
        e_start = torch.cuda.Event(enable_timing=True)
        e_stop = torch.cuda.Event(enable_timing=True)
        if main:
            torch.cuda.profiler.start()      # ncu --profile-from-start off captures exactly the timed region
        e_start.record()
        h0 = time.perf_counter()
        for _ in range(steps):
            loss, st = step(record=True)
        for w in works:
            w.wait()                      # stream-level wait: the all-reduces are inside the timed region
        e_stop.record()
        host_ms = (time.perf_counter() - h0) * 1e3 / steps     # CPU enqueue time per step
        sync_all()
        if main:
            torch.cuda.profiler.stop()
        t1 = time.perf_counter()
        sampler.stop_flag = True
        ms_total = e_start.elapsed_time(e_stop)
        if world > 1:
            mt = torch.tensor([ms_total], device=dev)
            dist.all_reduce(mt, op=dist.ReduceOp.MAX)
            ms_total = float(mt.item())

        def spread(a, c):
            v = sorted(x.elapsed_time(y) for x, y in zip(ev[a], ev[c]))
            return {"min": v[0], "median": v[len(v) // 2], "mean": float(np.mean(v)), "max": v[-1]}
        out["ms_step"] = ms_total / steps
        out["spreads"] = {"gae": spread("start", "gae"), "fwd": spread("gae", "fwd"), "bwd": spread("fwd", "bwd")}
        out["host_ms"] = host_ms
        out["clocks"] = sampler.summary(t0, t1) if main else None
        out["loss"] = float(loss.detach())
        out["global_loss"] = None
        if world > 1 and works:
            last = red_pool[(len(works) - 1) % len(red_pool)]
            out["global_loss"] = float((last[0] / last[1]).item())
        return out

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def graph_replay(mode, steps):
        """The same step captured once in a CUDA graph and replayed: what a trainer that graphs its
        post-processing pays per step (no per-launch host work).  N=1 only; reported beside `value`."""
        def step():
            P.cugae1d_nolp_misalign_func(b["rewards"], b["values"], b["cu_seqlens"], b["bootstrap"], GAMMA, LAM)
            logits.grad = None
            loss, st = F.fused_actor_loss(logits, b["labels"], TEMPERATURE, b["prox_logp"], b["old_logp"],
                                          b["advantages"], b["loss_mask"],
                                          loss_scale=loss_scale if mode == "single" else None, **CFG)
            (loss * loss_scale[0]).backward()
            return loss
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3):
                step()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            loss = step()
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            g.replay()
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / steps
        return {"ms_per_step": ms, "tokens_per_s": T / (ms * 1e-3), "steps": steps, "loss": float(loss.detach())}

    W = max(args.warmup, 3)
    main = timed_loop(args.mode, args.steps, W, True)
    other_mode = "twopass" if args.mode == "single" else "single"
    other = timed_loop(other_mode, max(3, min(args.steps, 20)), 3, False)
    single, twopass = (main, other) if args.mode == "single" else (other, main)
    ms_step = main["ms_step"]
    value = world * T / (ms_step * 1e-3)
    graph = None
    if world == 1 and os.environ.get("BENCH_NO_GRAPH") != "1":
        try:
            graph = graph_replay(args.mode, max(3, min(args.steps, 30)))
        except Exception as exc:            # reported, never fatal: the eager loop above is the headline
            graph = {"error": str(exc)[:200]}
    spreads = main["spreads"]
    host_ms = main["host_ms"]

    # per-kernel launch duration = MEDIAN over the timed steps of the event-to-event gap on the launch
    # stream (the mean also counts the host's launch latency whenever the queue runs dry, e.g. on the
    # first timed step right after the barrier; both are reported in kernel_ms_spread)
    coef_rows = int(torch.count_nonzero(logits.grad[:, :8].float().abs().sum(-1)).item()) if T else 0
    active = coef_rows / T
    peak, peak_src = measured_peaks()
    ms_gae = spreads["gae"]["median"]
    gae_bytes = 16.0 * T + 9.0 * b["n_seqs"] + 4

    def roof(kernel, bytes_, ms, **kw):
        gbs = bytes_ / (ms * 1e-3) / 1e9
        # DRAM bytes per launch from the committed `ncu --set full` captures (not measured live: ncu is
        # never run inside a bench); only valid for the shape they were captured on
        tr = NCU_DRAM_BYTES.get((kernel, T, V))
        return dict({"bound": "hbm", "kernel": kernel, "achieved": gbs, "peak": peak, "unit": "GB/s",
                     "frac": gbs / peak, "traffic": tr[0] if tr else None,
                     "traffic_source": tr[1] if tr else None, "peak_source": peak_src, "ms": ms,
                     "algorithmic_bytes": bytes_}, **kw)
    tp_fwd, tp_bwd = twopass["spreads"]["fwd"]["median"], twopass["spreads"]["bwd"]["median"]
    roof_two = {
        "ms_per_step": twopass["ms_step"], "tokens_per_s": world * T / (twopass["ms_step"] * 1e-3),
        "fwd": roof("logits_fwd_tma_kernel", 2.0 * T * V + 64.0 * T, tp_fwd),
        "bwd": roof("logits_bwd_tma_kernel", 2.0 * T * V * (1.0 + active), tp_bwd, active_row_frac=active,
                    note="bytes = 2TV written + 2TV*active read (rows whose gradient is zero are not read); "
                         "the reference's split API needs 4TV"),
    }
    sp_ms = single["spreads"]["fwd"]["median"]
    roof_single = roof("fused_rowres_kernel", 4.0 * T * V + 64.0 * T, sp_ms,
                       note="one read of the logits (2TV) + dlogits written (2TV); the two-pass split moves 6TV",
                       ms_per_step=single["ms_step"], tokens_per_s=world * T / (single["ms_step"] * 1e-3))
    if args.mode == "single":
        roofline = roof_single
    else:
        roofline = roof_two["bwd"]
    roof_gae = {"bound": "launch-latency", "kernel": "gae1d_kernel", "ms": ms_gae,
                "achieved": gae_bytes / (ms_gae * 1e-3) / 1e9, "unit": "GB/s", "algorithmic_bytes": gae_bytes}
    loss_value = main["loss"]

    clocks = main["clocks"]

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timed region
    e2e = None
    try:
        if args.skip_e2e:
            raise RuntimeError("skipped (--skip-e2e)")
        e2e = run_e2e(torch, dist, F, P, dev, b, world, max(3, min(args.steps, 5)), args.mode)
    except Exception as ex:  # pinned allocation can fail on a small host: say so instead of faking
        e2e = {"value": None, "unit": "tokens/s", "error": repr(ex)[:200]}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "accumulate_dtype": "f32 (GAE scan f64)", "data": "synthetic",
        "config": {"workload": "BASELINE.json configs[1]: Qwen2.5-1.5B GRPO math shape, T=16384 packed tokens "
                               "per GPU, V=151936 bf16 logits; step = packed GAE + fused logprob+entropy+decoupled "
                               "PPO loss + dlogits (mode=" + args.mode + ")", "tokens_per_gpu": T, "vocab": V,
                   "parallelism": f"dp{world} (batch-sharded by packed sequence)",
                   "l2": "inputs (4.98 GB logits per GPU) larger than the 126 MB L2; no flush needed",
                   "loss_cfg": {k: v for k, v in CFG.items()}, "temperature": TEMPERATURE},
        "roofline": roofline, "single_pass": roof_single, "two_pass": roof_two, "roofline_gae": roof_gae,
        "mode": args.mode,
        # our kernels per step: gae1d_prep + gae1d + (fused_rowres + rescale_if_needed | logits_fwd + scale + logits_bwd)
        "e2e": e2e, "gpu_launches": (4 if args.mode == "single" else 5) * args.steps, "clocks": clocks,
        "host_enqueue_ms_per_step": host_ms, "graph_replay": graph,
        "kernel_ms_spread": spreads,
        "loss": loss_value,
    }
    if main["global_loss"] is not None:
        line["global_loss"] = main["global_loss"]
    if world == 1 and not args.skip_cpu:
        line["cpu_baseline"] = cpu_baseline()
    emit(line)
    if world > 1:
        dist.destroy_process_group()


# dram__bytes_read.sum + dram__bytes_write.sum of one launch, (kernel, T, V) -> (bytes, capture)
NCU_DRAM_BYTES = {
    ("fused_rowres_kernel", 16384, 151936): (4.980306e9 + 4.924858e9, "profiles/ncu_rowres_r1g.txt"),
    ("logits_fwd_tma_kernel", 16384, 151936): (4.979488e9 + 0.003926e9, "profiles/ncu_logits_r1.txt"),
    ("logits_bwd_tma_kernel", 16384, 151936): (4.091398e9 + 4.932738e9, "profiles/ncu_logits_r1.txt"),
}


def run_e2e(torch, dist, F, P, dev, b, world, steps, mode="single"):
    """Same step, but the inputs start in pinned HOST memory every step and the results are read back:
    H2D of logits + per-token vectors, GAE + fused fwd + bwd on the device, D2H of the loss and the
    per-token logp / entropy / advantages."""
    T, V = b["logits"].shape
    scale = torch.full((1,), 0.5, dtype=torch.float32, device=dev)
    host = {}
    for k in ("logits", "labels", "loss_mask", "advantages", "rewards", "values", "cu_seqlens", "bootstrap",
              "prox_logp", "old_logp"):
        t = b[k].detach()
        h = torch.empty(t.shape, dtype=t.dtype, device="cpu", pin_memory=True)
        h.copy_(t)
        host[k] = h
    # two device-side input sets: step i's host->device copies run on a copy stream while step i-1
    # computes (what a trainer streaming micro-batches from host memory does); every byte of every
    # step is still copied inside the timed region
    dst2 = [{k: torch.empty_like(b[k].detach()) for k in host} for _ in range(2)]
    out_host = {k: torch.empty(T, dtype=torch.float32, pin_memory=True) for k in ("logp", "entropy", "adv", "ret")}
    loss_host = torch.empty((), dtype=torch.float32, pin_memory=True)
    h2d = sum(h.numel() * h.element_size() for h in host.values())
    d2h = 4 + sum(h.numel() * 4 for h in out_host.values())
    copy_stream = torch.cuda.Stream(dev)
    main_stream = torch.cuda.current_stream(dev)
    copied = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    for e in consumed:
        e.record(main_stream)

    def upload(i):
        s = i & 1
        copy_stream.wait_event(consumed[s])           # the compute that last read this set is done
        with torch.cuda.stream(copy_stream):
            for k in host:
                dst2[s][k].copy_(host[k], non_blocking=True)
            copied[s].record(copy_stream)

    def compute(i):
        s = i & 1
        dst = dst2[s]
        main_stream.wait_event(copied[s])
        x = dst["logits"].requires_grad_()
        adv, ret = P.cugae1d_nolp_misalign_func(dst["rewards"], dst["values"], dst["cu_seqlens"], dst["bootstrap"], GAMMA, LAM)
        loss, st = F.fused_actor_loss(x, dst["labels"], TEMPERATURE, dst["prox_logp"], dst["old_logp"],
                                      dst["advantages"], dst["loss_mask"],
                                      loss_scale=scale if mode == "single" else None, **CFG)
        (loss * scale[0]).backward()
        loss_host.copy_(loss.detach(), non_blocking=True)
        out_host["logp"].copy_(st["logprobs"], non_blocking=True)
        out_host["entropy"].copy_(st["entropy"], non_blocking=True)
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

    run(2)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    run(steps)
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / steps
    if world > 1:
        mt = torch.tensor([ms], device=dev)
        dist.all_reduce(mt, op=dist.ReduceOp.MAX)
        ms = float(mt.item())
    return {"value": world * T / (ms * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "ms_per_step": ms, "steps": steps,
            "note": "dominated by the 4.98 GB host->device copy of the logits over PCIe; copies of step i+1 overlap "
                    "the kernels of step i (two device-side input sets, copy stream)"}


def cpu_baseline(sample_rows=1024, steps=2, warmup=1):
    """The oracle port of the reference's CPU path, timed on this box's host cores on a bounded sample."""
    import torch
    from oracle import cpu_baseline as CB
    threads = os.cpu_count() or 1
    dt, used = CB.time_steps(sample_rows, VOCAB, steps, warmup, CFG, TEMPERATURE, threads=threads)
    return {"value": sample_rows / dt, "unit": "tokens/s", "cores": used, "kind": "port",
            "sample": f"{sample_rows} tokens x V={VOCAB} bf16 (1/{T_TOKENS // sample_rows} of the workload), "
                      f"{steps} steps after {warmup} warm-up; torch CPU eager log_softmax/entropy/PPO loss + backward "
                      f"on {used} threads, GAE as the reference's per-token Python loop (1 thread)",
            "seconds_per_step": dt}


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path (oracle port) on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample_rows = 1024
    steps = max(1, min(args.steps, 5))
    warmup = max(1, min(args.warmup, 2))
    import torch
    from oracle import cpu_baseline as CB
    threads = os.cpu_count() or 1
    dt, used = CB.time_steps(sample_rows, VOCAB, steps, warmup, CFG, TEMPERATURE, threads=threads)
    v = sample_rows / dt
    cb = {"value": v, "unit": "tokens/s", "cores": used, "kind": "port",
          "sample": f"each step = {sample_rows} tokens x V={VOCAB} bf16 (bounded sample of the {T_TOKENS}-token "
                    f"workload); {steps} steps after {warmup} warm-up"}
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
    ap.add_argument("--mode", default="single", choices=["single", "twopass"],
                    help="single: one-pass fused forward+gradient (loss_scale announced); twopass: forward then backward")
    ap.add_argument("--skip-e2e", action="store_true", help="profiling runs: skip the host-buffer leg")
    ap.add_argument("--skip-cpu", action="store_true", help="profiling runs: skip the CPU baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
