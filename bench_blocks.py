"""Clocked side blocks of bench.py: the BASELINE.json configs beside the headline one, each measured with
CUDA events after warm-up, on inputs larger than L2 (or a ring of input sets whose total is), with the SM
clock and throttle reasons sampled while the block runs (`clocks` in every block).

N = 1:  gae_sweep   packed GAE at 32 k / 4 M / 2^27 tokens: eager time per call and GPU time per call
                    (the same calls replayed from a CUDA graph), GB/s, fraction of the measured peak
        intervals   configs[4]: slice / set on a 7.6 G-element bf16 buffer, 1 k and 100 k intervals of the
                    reference test's pattern, and the 200,704-row 7B TP 1->2 re-shard; bit-exact spot checks
        cfg2_step   configs[2] on one GPU: 32,768 packed tokens, padded masked GAE + packed GAE + actor loss
                    (engine loop: hook = single pass, and two passes) + critic loss fwd+bwd
        kernels     padded GAE, fused advantages, vocabulary-shard forward [131072, 18992] (configs[3]'s
                    per-GPU kernel), AEnt clamped entropy, LM-head forward-only log-probs
N > 1:  multi_gpu_parity   vocab-sharded == unsharded, peer-memory exchange == all-gather (small T)
        vocab_sharded      configs[3]: T = 131,072, shard [T, V/N]: fwd and fwd+bwd, both exchanges
        cfg2_strong        configs[2]: the 32,768-token batch split by sequence over the N ranks
"""
import functools
import math
import os
import types

import numpy as np

V = 151936
CFG = dict(eps_clip=0.2, eps_clip_higher=None, c_clip=None, behav_imp_weight_cap=5.0)


class Blocks:
    def __init__(self, torch, dist, dev, world, rank, local_rank, peak, Clocked, EngineLoop):
        self.torch, self.dist, self.dev = torch, dist, dev
        self.world, self.rank, self.local_rank, self.peak, self.Clocked = world, rank, local_rank, peak, Clocked
        self.EngineLoop = EngineLoop

    # ---- helpers ---------------------------------------------------------------------------------------
    def run(self, name):
        torch = self.torch
        try:
            clk = self.Clocked(self.local_rank)
            with clk:
                out = getattr(self, name)()
            if isinstance(out, dict):
                out["clocks"] = clk.clocks
            return out
        except Exception as ex:               # a side block never takes the headline down; the error is the record
            import traceback
            return {"error": repr(ex)[:300], "where": traceback.format_exc()[-600:]}
        finally:
            torch.cuda.synchronize(self.dev)
            torch.cuda.empty_cache()

    def sync(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def time_ms(self, fn, iters=10, warmup=3, max_over_ranks=True):
        """mean ms per call of `iters` back-to-back calls between two events on the launch stream"""
        torch = self.torch
        for _ in range(warmup):
            fn()
        self.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        self.sync()
        ms = e0.elapsed_time(e1) / iters
        if self.world > 1 and max_over_ranks:
            t = torch.tensor([ms], device=self.dev)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def frac(self, nbytes, ms):
        gbs = nbytes / (ms * 1e-3) / 1e9
        return {"ms": ms, "gbs": gbs, "frac_of_measured_peak": gbs / self.peak, "algorithmic_bytes": int(nbytes)}

    def logits(self, T, Vs, seed):
        torch = self.torch
        g = torch.Generator(device=self.dev).manual_seed(seed)
        x = torch.empty(T, Vs, dtype=torch.bfloat16, device=self.dev)
        rows = max(1, (1 << 28) // Vs)
        for i in range(0, T, rows):
            n = min(rows, T - i)
            x[i:i + n] = (torch.randn(n, Vs, generator=g, device=self.dev) * 2.0).bfloat16()
        return x, g

    # ---- N = 1 -----------------------------------------------------------------------------------------
    def gae_sweep(self):
        torch = self.torch
        from areal_b200 import ppo_functional as P
        rows = []
        for T in (32768, 1 << 22, 1 << 27):
            rng = np.random.RandomState(T % 1000)
            per_set = 16 * T
            n_sets = int(min(512, max(1, math.ceil((300 << 20) / per_set)))) if per_set < (300 << 20) else 1
            sets = []
            for s in range(n_sets):
                lens, left = [], T
                while left > 0:
                    n = min(left, int(rng.randint(128, 4097)))
                    lens.append(n)
                    left -= n
                B = len(lens)
                cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(self.dev)
                sets.append((torch.randn(T, device=self.dev), torch.randn(T + B, device=self.dev), cu,
                             torch.rand(B, device=self.dev) > 0.5, B))
            nbytes = 16.0 * T + 9.0 * sets[0][4] + 4

            def sweep():
                for r, v, cu, boot, _ in sets:
                    P.cugae1d_nolp_misalign_func(r, v, cu, boot, 1.0, 0.95)
            reps = 1 if n_sets > 1 else 4

            def eager():
                for _ in range(reps):
                    sweep()
            ms_eager = self.time_ms(eager, iters=5, warmup=2) / (n_sets * reps)
            # the same calls replayed from a CUDA graph: GPU time per call without the wrapper's host time
            side = torch.cuda.Stream(self.dev)
            side.wait_stream(torch.cuda.current_stream(self.dev))
            with torch.cuda.stream(side):
                sweep()
            torch.cuda.current_stream(self.dev).wait_stream(side)
            torch.cuda.synchronize(self.dev)
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                eager()
            ms_gpu = self.time_ms(gr.replay, iters=5, warmup=2) / (n_sets * reps)
            del gr
            row = {"tokens": T, "sequences": sets[0][4], "input_sets_in_ring": n_sets,
                   "l2": "ring of input sets larger than L2 in total" if n_sets > 1 else "inputs larger than L2",
                   "eager_us_per_call": ms_eager * 1e3, "gpu_us_per_call": ms_gpu * 1e3,
                   "tokens_per_s_gpu_time": T / (ms_gpu * 1e-3)}
            row.update({"eager": self.frac(nbytes, ms_eager), "gpu_time": self.frac(nbytes, ms_gpu)})
            rows.append(row)
            del sets
            torch.cuda.empty_cache()
        return {"kernel": "gae1d_prep_kernel + gae1d_kernel (areal_b200_gae1d_nolp_misalign)", "gamma_lam": [1.0, 0.95],
                "bytes_per_token": 16, "rows": rows}

    def intervals(self):
        torch = self.torch
        from areal_b200 import flatten_param as fp
        n_elem = int(os.environ.get("INTERVAL_BUFFER_ELEMS", 7_600_000_000))
        src = torch.empty(n_elem, dtype=torch.bfloat16, device=self.dev)
        src.view(torch.int16).random_(-30000, 30000)
        rng = np.random.RandomState(0)
        rows = []

        def spot_check(packed, iv, k=16):
            off = np.concatenate([[0], np.cumsum(iv[:, 1] - iv[:, 0])])
            for j in list(range(min(k, len(iv)))) + [len(iv) - 1]:
                a, e = int(iv[j, 0]), int(iv[j, 1])
                if not torch.equal(packed[int(off[j]):int(off[j + 1])].view(torch.int16), src[a:e].view(torch.int16)):
                    return False
            return True
        for n in (1000, 100000):
            s = n_elem // n
            iv = np.stack([np.arange(n, dtype=np.int64) * s, np.arange(n, dtype=np.int64) * s + s // 2], 1)
            rng.shuffle(iv)                                           # realhf/tests/cpp_extensions/test_interval_ops.py:20-31
            total = n * (s // 2)
            ivc = torch.from_numpy(iv).to(self.dev)
            nbytes = 4.0 * total + 16.0 * n
            ms = self.time_ms(lambda: fp.slice_intervals_bf16(src, ivc, total, s // 2), iters=5, warmup=2)
            packed = fp.slice_intervals_bf16(src, ivc, total, s // 2)
            ok = spot_check(packed, iv)
            rows.append(dict(op="slice_intervals", n_intervals=n, interval_elems=s // 2, bit_exact_spot_check=ok,
                             **self.frac(nbytes, ms)))
            ms = self.time_ms(lambda: fp.set_intervals_bf16(packed, src, ivc, s // 2), iters=5, warmup=2)
            rows.append(dict(op="set_intervals", n_intervals=n, interval_elems=s // 2, **self.frac(nbytes, ms)))
            del packed
        # TP 1->2 re-shard of a Qwen2.5-7B-shaped model: one interval per matrix row of the row-parallel weights
        # (realhf/impl/model/parallelism/tensor_parallel/real_llm_parallel.py:101-107, 170-176)
        iv, base = [], 0
        for layer in range(28):
            for width in (3584, 18944):
                r = np.arange(3584, dtype=np.int64)
                iv.append(np.stack([base + r * width + width // 2, base + (r + 1) * width], 1))
                base += 3584 * width
        iv = np.concatenate(iv)
        total = int((iv[:, 1] - iv[:, 0]).sum())
        ivc = torch.from_numpy(iv).to(self.dev)
        ms = self.time_ms(lambda: fp.slice_intervals_bf16(src, ivc, total, 9472), iters=5, warmup=2)
        packed = fp.slice_intervals_bf16(src, ivc, total, 9472)
        rows.append(dict(op="slice_intervals (7B TP 1->2 re-shard rows)", n_intervals=len(iv), bit_exact_spot_check=spot_check(packed, iv),
                         **self.frac(4.0 * total + 16.0 * len(iv), ms)))
        del src, packed
        return {"kernel": "interval_scan_kernel + interval_copy_kernel", "buffer_elems": n_elem, "dtype": "bf16",
                "l2": "15.2 GB buffer, every call streams GBs: larger than L2", "rows": rows}

    def _cfg2_inputs(self, T, seed):
        """32 k-token PPO-with-critic micro-batch: packed view for the losses / packed GAE, padded [B, S] view for
        the areal-lite GAE (areal/engine/ppo/actor.py:131-152)."""
        rng = np.random.RandomState(seed)
        lens, left = [], T
        while left > 0:
            n = min(left, int(rng.randint(128, 4097)))
            lens.append(n)
            left -= n
        return self._cfg2_inputs_from_lens(lens, seed)

    def _cfg2_step_fn(self, d, mode):
        from areal_b200 import actor as A, critic as C, engine_hooks as EH, ppo_functional as P, stats_tracker as ST
        torch = self.torch
        tracker = ST.DistributedStatsTracker()
        engine = self.EngineLoop(torch, self.dist, self.dev, self.world)
        if mode == "hook":
            EH.install_loss_scale_hook(engine)
        actor_fn = functools.partial(A.grpo_loss_fn, temperature=1.0, tracker=tracker, **CFG)
        T = d["logits"].shape[0]

        def step():
            d["logits"].grad = None
            d["new_values"].grad = None
            adv2, ret2 = P.gae_padded_masked(d["pad_rewards"], d["pad_values"], d["pad_mask"], d["no_eos"], 1.0, 0.95)
            adv, ret = P.cugae1d_nolp_misalign_func(d["rewards"], d["values"], d["cu"], d["boot"], 1.0, 0.95)
            mb = dict(_logits=d["logits"], input_ids=d["labels"], rolled_input_ids=d["labels"], logprobs=d["old"],
                      prox_logp=d["prox"], advantages=adv, loss_mask=d["mask"])
            engine.train_batch([mb], actor_fn, lambda x: x["loss_mask"].count_nonzero())
            cmb = dict(values=d["values"][:T], returns=ret, loss_mask=d["mask"])
            closs = C.ppo_loss_fn(d["new_values"], cmb, eps_clip=0.2, tracker=tracker)
            closs.backward()
        return step, tracker

    def cfg2_step(self):
        torch = self.torch
        T = 32768
        d = self._cfg2_inputs(T, seed=2)
        out = {"workload": "BASELINE.json configs[2] on one GPU: 32,768 packed tokens (V=151936 bf16): padded masked GAE "
                           f"[{d['B']},{d['S']}] + packed GAE + actor loss through the engine loop + critic loss, fwd+bwd",
               "tokens": T, "l2": "9.96 GB of logits per step: larger than L2"}
        for mode, nbytes in (("hook", 4.0 * T * V), ("twopass", 6.0 * T * V)):
            step, tracker = self._cfg2_step_fn(d, mode)
            ms = self.time_ms(step, iters=10, warmup=3)
            tracker.export()
            out[mode] = dict(tokens_per_s=T / (ms * 1e-3), **self.frac(nbytes, ms))
        out["hook"]["note"] = "single pass: logits read once + dlogits written (4TV)"
        out["twopass"]["note"] = "forward + backward kernels (6TV upper bound; rows with a zero gradient are not read)"
        return out

    def kernels(self):
        torch = self.torch
        from areal_b200 import aent_functional as AE, dist as D, functional as F, ppo_functional as P
        from areal_b200.actor import compute_advantages
        rows = []
        for B, S in ((16, 32768), (512, 4096), (4096, 4096)):
            nbytes = 20 * B * S
            n_sets = int(min(64, max(1, math.ceil((300 << 20) / nbytes))))
            sets = [(torch.randn(B, S, device=self.dev), torch.randn(B, S, device=self.dev),
                     (torch.rand(B, S, device=self.dev) > 0.2).float(), torch.rand(B, device=self.dev) > 0.5)
                    for _ in range(n_sets)]

            def sweep():
                for r, v, m, ne in sets:
                    P.gae_padded_masked(r, v, m, ne, 1.0, 0.95)
            ms_eager = self.time_ms(sweep, iters=5, warmup=2) / n_sets
            ms_gpu = self._graph_time(sweep) / n_sets
            rows.append(dict(kernel="gae_padded_masked", B=B, S=S, input_sets_in_ring=n_sets,
                             l2="ring of input sets larger than L2 in total",
                             eager_us_per_call=ms_eager * 1e3, gpu_us_per_call=ms_gpu * 1e3,
                             tokens_per_s=B * S / (ms_gpu * 1e-3), **self.frac(20.0 * B * S, ms_gpu)))
            del sets
            torch.cuda.empty_cache()
        B, S = 2048, 8192
        lens = torch.randint(2, S + 1, (B,), device=self.dev)
        am = torch.arange(S, device=self.dev)[None, :] < lens[:, None]
        base = dict(input_ids=torch.zeros(B, S, dtype=torch.long, device=self.dev), attention_mask=am,
                    loss_mask=am & (torch.rand(B, S, device=self.dev) > 0.2), logprobs=-torch.rand(B, S, device=self.dev),
                    ref_logp=-torch.rand(B, S, device=self.dev), values=torch.randn(B, S, device=self.dev),
                    rewards=torch.randn(B, device=self.dev))
        actor = types.SimpleNamespace(
            config=types.SimpleNamespace(use_decoupled_loss=True, recompute_logprob=True, overlong_reward_penalty=False),
            reward_bias=0.0, reward_scaling=1.0, reward_clip=20.0, kl_ctl=0.1,
            kl_estimator=types.SimpleNamespace(kl_estimator="k1", apply_clamp=True), discount=1.0, gae_lambda=0.95,
            mask_no_eos_with_zero=False, adv_norm=None, reward_norm=None)
        nbytes = (2 + 4 + 8 + 4 + 24) * B * S
        ms = self.time_ms(lambda: compute_advantages(actor, dict(base)), iters=10, warmup=3)
        rows.append(dict(kernel="advantages_padded (compute_advantages, one launch)", B=B, S=S,
                         tokens_per_s=B * S / (ms * 1e-3), **self.frac(nbytes, ms)))
        del base, am
        # few long rows (long chains of thought): a cluster of 8 CTAs per row; GPU time from a graph replay over a
        # ring of input sets that is larger than L2 in total
        B, S = 16, 32768
        n_sets = 16
        ring = []
        for _ in range(n_sets):
            lens = torch.randint(S // 2, S + 1, (B,), device=self.dev)
            am = torch.arange(S, device=self.dev)[None, :] < lens[:, None]
            ring.append(dict(input_ids=torch.zeros(B, S, dtype=torch.long, device=self.dev), attention_mask=am,
                             loss_mask=am & (torch.rand(B, S, device=self.dev) > 0.2),
                             logprobs=-torch.rand(B, S, device=self.dev), ref_logp=-torch.rand(B, S, device=self.dev),
                             values=torch.randn(B, S, device=self.dev), rewards=torch.randn(B, device=self.dev)))

        def adv_sweep():
            for d in ring:
                compute_advantages(actor, dict(d))
        ms_eager = self.time_ms(adv_sweep, iters=5, warmup=2) / n_sets
        ms_gpu = self._graph_time(adv_sweep) / n_sets
        rows.append(dict(kernel="advantages_padded (compute_advantages), few long rows", B=B, S=S,
                         input_sets_in_ring=n_sets, eager_us_per_call=ms_eager * 1e3, gpu_us_per_call=ms_gpu * 1e3,
                         note="gpu time includes the wrapper's small kernels (mask casts, clones) of the same call",
                         **self.frac((2 + 4 + 8 + 4 + 24) * B * S, ms_gpu)))
        del ring
        # configs[3]'s per-GPU kernel on one GPU: a [131072, V/8] vocabulary shard
        T, Vs = 131072, V // 8
        x, g = self.logits(T, Vs, 5)
        lab = torch.randint(0, V, (T,), generator=g, device=self.dev)
        ms = self.time_ms(lambda: D._cuda_partials(x, lab, 0, 1.0), iters=10, warmup=3)
        rows.append(dict(kernel="logits_partial_fwd (vocabulary shard [131072, 18992])", **self.frac(2.0 * T * Vs, ms)))
        ms = self.time_ms(lambda: F.gather_logprobs_entropy(x, lab % Vs, 1.0), iters=10, warmup=3)
        rows.append(dict(kernel="logits_fwd (rows of 18992 bf16)", **self.frac(2.0 * T * Vs, ms)))
        del x
        torch.cuda.empty_cache()
        # AEnt clamped entropy (recipe/AEnt/functional.py:15-80)
        T = 8192
        x, g = self.logits(T, V, 6)
        lab = torch.randint(0, V, (T,), generator=g, device=self.dev)
        xr = x.requires_grad_()
        w = torch.randn(T, generator=g, device=self.dev)
        with torch.no_grad():
            ms = self.time_ms(lambda: AE.gather_logprobs_clamped_entropy(x.detach(), lab, 0.3, 1.0), iters=5, warmup=2)
        rows.append(dict(kernel="clamped_entropy_fwd (AEnt, clamp 0.3)", T=T, tokens_per_s=T / (ms * 1e-3), **self.frac(2.0 * T * V, ms)))

        def fb():
            xr.grad = None
            lp, h = AE.gather_logprobs_clamped_entropy(xr, lab, 0.3, 1.0)
            ((lp * w).sum() + (h * w).sum()).backward()
        ms = self.time_ms(fb, iters=5, warmup=2)
        rows.append(dict(kernel="clamped_entropy fwd+bwd", T=T, tokens_per_s=T / (ms * 1e-3), **self.frac(6.0 * T * V, ms)))
        del x, xr
        torch.cuda.empty_cache()
        rows += self._lm_head_rows()
        return {"rows": rows}

    def e2e_from_hidden(self):
        """LABELLED EXTRA, not the headline `e2e`: the same step at a different boundary -- the engine hands over the
        last hidden state [T, H] and the LM-head weight instead of [T, V] logits (areal_b200.actor.
        grpo_loss_fn_from_hidden) -- with host buffers: every step copies the hidden state (50 MB instead
        of 4.98 GB) and the per-token vectors from pinned host memory, runs packed GAE + loss forward + backward down
        to d_hidden / d_weight, and reads back the loss and the advantages / returns.  The weight stays on the device
        (it is a parameter)."""
        torch = self.torch
        import functools as ft
        from areal_b200 import actor as A, ppo_functional as P, stats_tracker as ST
        T, Hd = 16384, 1536
        g = torch.Generator(device=self.dev).manual_seed(3)
        weight = (torch.randn(V, Hd, generator=g, device=self.dev) * 0.03).bfloat16().requires_grad_()
        rng = np.random.RandomState(3)
        lens, left = [], T
        while left > 0:
            n = min(left, int(rng.randint(128, 4097)))
            lens.append(n)
            left -= n
        B = len(lens)
        dev_in = dict(hidden=torch.randn(T, Hd, generator=g, device=self.dev).bfloat16(),
                      labels=torch.randint(0, V, (T,), generator=g, device=self.dev),
                      loss_mask=torch.rand(T, generator=g, device=self.dev) > 0.15,
                      advantages=torch.randn(T, generator=g, device=self.dev),
                      old_logp=-torch.rand(T, generator=g, device=self.dev) * 5,
                      prox_logp=-torch.rand(T, generator=g, device=self.dev) * 5,
                      rewards=torch.randn(T, generator=g, device=self.dev),
                      values=torch.randn(T + B, generator=g, device=self.dev),
                      cu_seqlens=torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(self.dev),
                      bootstrap=torch.rand(B, generator=g, device=self.dev) > 0.5)
        host = {}
        for k, t in dev_in.items():
            host[k] = torch.empty(t.shape, dtype=t.dtype, device="cpu", pin_memory=True)
            host[k].copy_(t)
        dst2 = [{k: torch.empty_like(t) for k, t in dev_in.items()} for _ in range(2)]
        out_host = {k: torch.empty(T, dtype=torch.float32, pin_memory=True) for k in ("adv", "ret")}
        loss_host = torch.empty((), dtype=torch.float32, pin_memory=True)
        h2d = sum(h.numel() * h.element_size() for h in host.values())
        copy_stream, main = torch.cuda.Stream(self.dev), torch.cuda.current_stream(self.dev)
        copied = [torch.cuda.Event() for _ in range(2)]
        consumed = [torch.cuda.Event() for _ in range(2)]
        for e in consumed:
            e.record(main)
        tracker = ST.DistributedStatsTracker()
        loss_fn = ft.partial(A.grpo_loss_fn_from_hidden, temperature=1.0, tracker=tracker, **CFG)

        def upload(i):
            s = i & 1
            copy_stream.wait_event(consumed[s])
            with torch.cuda.stream(copy_stream):
                for k in host:
                    dst2[s][k].copy_(host[k], non_blocking=True)
                copied[s].record(copy_stream)

        def compute(i):
            s = i & 1
            d = dst2[s]
            main.wait_event(copied[s])
            h = d["hidden"].requires_grad_()
            adv, ret = P.cugae1d_nolp_misalign_func(d["rewards"], d["values"], d["cu_seqlens"], d["bootstrap"], 1.0, 0.95)
            mb = dict(input_ids=d["labels"], rolled_input_ids=d["labels"], logprobs=d["old_logp"],
                      prox_logp=d["prox_logp"], advantages=d["advantages"], loss_mask=d["loss_mask"])
            loss = loss_fn(h, weight, mb)
            loss.backward()
            loss_host.copy_(loss.detach(), non_blocking=True)
            out_host["adv"].copy_(adv, non_blocking=True)
            out_host["ret"].copy_(ret, non_blocking=True)
            h.grad = None
            weight.grad = None
            d["hidden"] = h.detach()
            consumed[s].record(main)

        def run(n):
            upload(0)
            for i in range(n):
                if i + 1 < n:
                    upload(i + 1)
                compute(i)
        run(2)
        torch.cuda.synchronize(self.dev)
        steps = 5
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(steps)
        e1.record()
        torch.cuda.synchronize(self.dev)
        ms = e0.elapsed_time(e1) / steps
        tracker.export()
        return {"label": "EXTRA, another boundary than `e2e`: hidden state [T, H] in, no [T, V] tensor anywhere",
                "api": "areal_b200.actor.grpo_loss_fn_from_hidden (row blocks: library GEMM for the block's logits, "
                       "our single-pass kernel, two library GEMMs for d_hidden / d_weight)",
                "tokens": T, "hidden_size": Hd, "vocab": V, "value": T / (ms * 1e-3), "unit": "tokens/s",
                "ms_per_step": ms, "steps": steps, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4 + 8 * T,
                "loss": float(loss_host)}

    def _graph_time(self, fn, iters=5):
        """ms per replay of `fn` captured once in a CUDA graph: GPU time without the wrappers' host time"""
        torch = self.torch
        side = torch.cuda.Stream(self.dev)
        side.wait_stream(torch.cuda.current_stream(self.dev))
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream(self.dev).wait_stream(side)
        torch.cuda.synchronize(self.dev)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            fn()
        ms = self.time_ms(gr.replay, iters=iters, warmup=2)
        del gr
        return ms

    def _ring_time(self, fn, nbytes):
        """time one small call with the L2 flushed (a 256 MB write) before each timed call"""
        torch = self.torch
        if nbytes >= (256 << 20):
            return self.time_ms(fn, iters=10, warmup=3)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)
        for _ in range(3):
            fn()
        ts = []
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize(self.dev)
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        return ts[len(ts) // 2]

    def _lm_head_rows(self):
        """Forward-only log-probs from the hidden state (PPOActor.compute_logp's shape, areal/engine/ppo/actor.py:52-70)
        at T=16384, V=151936 and H=1536 (Qwen2.5-1.5B) / 3584 (7B): the tcgen05 kernel (no [T,V] tensor) against
        library GEMM + streaming pass.  `frac_of_measured_bf16_peak`: against MEASURED_PEAKS.json bf16_tflops (burst:
        these kernels are timed alone)."""
        import json
        torch = self.torch
        from areal_b200 import lm_head as LM
        try:
            tpeak = float(json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")))["bf16_tflops"])
            tsrc = "measured (MEASURED_PEAKS.json bf16_tflops, burst)"
        except Exception:
            tpeak, tsrc = 1590.0, "fallback (B200_PROFILING.md)"
        rows = []
        for Hd in (1536, 3584):
            T = 16384
            g = torch.Generator(device=self.dev).manual_seed(3)
            h = torch.randn(T, Hd, generator=g, device=self.dev).bfloat16()
            w = (torch.randn(V, Hd, generator=g, device=self.dev) * (2.0 / Hd ** 0.5)).bfloat16()
            lab = torch.randint(0, V, (T,), generator=g, device=self.dev)
            flops = 2.0 * T * Hd * V
            ms = self.time_ms(lambda: LM.lm_head_logprobs(h, w, lab, 1.0, chunk_rows=4096, with_entropy=True, impl="library"), iters=5, warmup=2)
            rows.append(dict(kernel="lm_head_logprobs: cuBLAS GEMM per 4096-row block + logits_fwd (library + streaming)", H=Hd,
                             ms=ms, tflops=flops / (ms * 1e-3) / 1e12, tokens_per_s=T / (ms * 1e-3)))
            ms2 = self.time_ms(lambda: LM.lm_head_logprobs_tc(h, w, lab, 1.0, with_entropy=True), iters=5, warmup=2)
            a = LM.lm_head_logprobs_tc(h, w, lab, 1.0, with_entropy=True)
            b = LM.lm_head_logprobs(h, w, lab, 1.0, chunk_rows=4096, with_entropy=True, impl="library")
            tf = flops / (ms2 * 1e-3) / 1e12
            rows.append(dict(kernel="lm_head_logprobs_tc: tcgen05 GEMM with the online-softmax epilogue (no [T,V] tensor)", H=Hd,
                             ms=ms2, tflops=tf, tokens_per_s=T / (ms2 * 1e-3), bound="tensor", peak_tflops=tpeak,
                             peak_source=tsrc, frac_of_measured_bf16_peak=tf / tpeak, speedup_vs_library_path=ms / ms2,
                             max_abs_diff_logp_vs_library_path=float((a[0] - b[0]).abs().max()),
                             max_abs_diff_entropy_vs_library_path=float((a[1] - b[1]).abs().max())))
            del h, w
            torch.cuda.empty_cache()
        return rows

    # ---- N > 1 -----------------------------------------------------------------------------------------
    def multi_gpu_parity(self):
        torch, dist = self.torch, self.dist
        from areal_b200 import dist as D, functional as F
        g = torch.Generator(device=self.dev).manual_seed(7)          # same stream on every rank
        T = 96
        logits = (torch.randn(T, V, generator=g, device=self.dev) * 2).bfloat16()
        labels = torch.randint(0, V, (T,), generator=g, device=self.dev)
        w_lp = torch.randn(T, generator=g, device=self.dev)
        w_h = torch.randn(T, generator=g, device=self.dev) * 0.1
        Vs = V // self.world
        lo = self.rank * Vs
        shard = logits[:, lo:lo + Vs].detach().clone().requires_grad_()
        lp, ent = D.vocab_parallel_logprobs_entropy(shard, labels, 0.9)
        ((lp * w_lp).sum() + (ent * w_h).sum()).backward()
        assert Vs * self.world == V, "the Qwen vocabulary divides by 2, 4 and 8"
        full = logits.detach().clone().requires_grad_()
        lab_ok = labels
        lp1, ent1 = F.gather_logprobs_entropy(full, lab_ok, 0.9)
        ((lp1 * w_lp).sum() + (ent1 * w_h).sum()).backward()
        checks = {"vocab_sharded_logp_eq_unsharded": bool(torch.allclose(lp, lp1, rtol=1e-5, atol=2e-5)),
                  "vocab_sharded_entropy_eq_unsharded": bool(torch.allclose(ent, ent1, rtol=1e-5, atol=2e-5)),
                  "vocab_sharded_grad_eq_unsharded": bool(torch.allclose(shard.grad.float(), full.grad[:, lo:lo + Vs].float(),
                                                                         rtol=2 ** -7, atol=1e-9))}
        try:
            ex = D.PeerExchange(max_tokens=4096)
            sh = logits[:, lo:lo + Vs].contiguous()
            same = True
            for Tp in (96, 1, 77):
                a = D.vocab_parallel_forward(sh[:Tp], lab_ok[:Tp], lo, 0.9)
                b = ex.forward(sh[:Tp], lab_ok[:Tp], lo, 0.9)
                same = same and all(torch.equal(u, v) for u, v in zip(a, b))
            ex.check()
            checks["peer_exchange_bit_identical_to_all_gather"] = bool(same)
            fx = D.FusedPeerExchange(max_tokens=4096)
            same = True
            for Tp in (96, 1, 77, 96):
                a = D.vocab_parallel_forward(sh[:Tp], lab_ok[:Tp], lo, 0.9)
                b = fx.forward(sh[:Tp], lab_ok[:Tp], lo, 0.9)
                same = same and all(torch.equal(u, v) for u, v in zip(a, b))
            fx.check()
            checks["fused_push_exchange_bit_identical_to_all_gather"] = bool(same)
        except Exception as exc:
            checks["peer_exchange"] = "unavailable: " + repr(exc)[:120]
        flags = torch.tensor([int(all(v for v in checks.values() if isinstance(v, bool)))], device=self.dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        return {"multi_gpu_parity": "ok" if int(flags.item()) == 1 else "FAILED", "checks_rank0": checks, "world": self.world}

    def vocab_sharded(self):
        torch, dist = self.torch, self.dist
        from areal_b200 import dist as D
        T = int(os.environ.get("VP_TOKENS", 131072))
        Vs = V // self.world
        shard, g = self.logits(T, Vs, 11 + self.rank)
        gl = torch.Generator(device=self.dev).manual_seed(1234)     # same labels / weights on every rank
        labels = torch.randint(0, V, (T,), generator=gl, device=self.dev)
        w = torch.randn(T, generator=gl, device=self.dev)
        w[torch.rand(T, generator=gl, device=self.dev) < 0.2] = 0.0
        active = float((w != 0).float().mean().item())
        out = {"workload": f"BASELINE.json configs[3]: T={T} tokens, V={V} bf16 vocab-sharded over {self.world} GPUs "
                           f"(shard [T, {Vs}]): partial statistics, exchange, combine, local backward",
               "tokens": T, "shard_width": Vs, "l2": "shard larger than L2"}
        try:
            ex = D.PeerExchange(max_tokens=T)
            fx = D.FusedPeerExchange(max_tokens=T)
        except Exception as exc:
            ex, fx, out["peer_memory"] = None, None, {"unavailable": repr(exc)[:200]}
        x = shard.requires_grad_()
        fwd_bytes = 2.0 * T * Vs
        bwd_bytes = 2.0 * T * Vs * (1.0 + active)
        out["variants"] = {"allgather": "partial kernel, NCCL all-gather of the partials, combine kernel",
                           "peer_memory": "partial kernel, combine kernel that reads the peers' partials over NVLink",
                           "fused_push": "ONE launch: the streaming kernel's row epilogue stores the partials into every "
                                         "rank's buffer over NVLink, the same CTAs combine once all rows are in"}
        for name, exch in (("allgather", None), ("peer_memory", ex), ("fused_push", fx)):
            if exch is None and name != "allgather":
                continue

            def fwd():
                with torch.no_grad():
                    return D.vocab_parallel_logprobs_entropy(x.detach(), labels, 1.0, exchange=exch)

            def fwd_bwd():
                x.grad = None
                lp, ent = D.vocab_parallel_logprobs_entropy(x, labels, 1.0, exchange=exch)
                (lp * w).sum().backward()
            ms_f = self.time_ms(fwd, iters=10, warmup=3)
            ms_fb = self.time_ms(fwd_bwd, iters=10, warmup=3)
            out[name] = {"fwd_ms": ms_f, "fwd_bwd_ms": ms_fb, "tokens_per_s": T / (ms_fb * 1e-3),
                         "fwd_hbm_frac": fwd_bytes / (ms_f * 1e-3) / 1e9 / self.peak,
                         "fwd_bwd_hbm_frac": (fwd_bytes + bwd_bytes) / (ms_fb * 1e-3) / 1e9 / self.peak}
            if exch is not None:
                exch.check()
        # the per-shard kernel and the exchange on their own
        ms_k = self.time_ms(lambda: D._cuda_partials(shard.detach(), labels, self.rank * Vs, 1.0), iters=10, warmup=3)
        parts = D._cuda_partials(shard.detach(), labels, self.rank * Vs, 1.0)
        ms_x = self.time_ms(lambda: D.gather_partials(parts), iters=20, warmup=5)
        out["partial_kernel"] = self.frac(fwd_bytes, ms_k)
        # Why the synchronised forward costs more than the slowest rank's kernel ALONE: every step waits for the rank
        # that happens to be slowest in THAT step.  Per-step kernel times of every rank (one event pair per step, the
        # ranks released together by a barrier), then mean over steps of the max over ranks against the max over
        # ranks of the per-rank means.
        n_it = 20
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_it)]
        for a, b in evs:
            dist.barrier()
            a.record()
            D._cuda_partials(shard.detach(), labels, self.rank * Vs, 1.0)
            b.record()
        torch.cuda.synchronize(self.dev)
        mine = torch.tensor([a.elapsed_time(b) for a, b in evs], device=self.dev)
        allt = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(allt, mine)
        allt = torch.stack(allt)                                     # [ranks, steps]
        out["partial_kernel_per_step"] = {
            "steps": n_it, "max_over_ranks_of_mean_ms": float(allt.mean(1).max()),
            "mean_over_steps_of_max_over_ranks_ms": float(allt.max(0).values.mean()),
            "fastest_rank_mean_ms": float(allt.mean(1).min()), "slowest_single_step_ms": float(allt.max()),
            "note": "a synchronised step (any exchange) costs the second figure plus the exchange, not the first"}
        out["all_gather_us"] = ms_x * 1e3
        out["nccl_payload_bytes_per_rank"] = 16 * T
        # full-size properties (size-independent): log-prob <= 0, 0 <= entropy <= ln V, gradient rows sum to 0
        xx = shard.detach().requires_grad_()
        lp_a, ent_a = D.vocab_parallel_logprobs_entropy(xx, labels, 1.0)
        checks = {"logp_le_0": bool((lp_a <= 0).all()),
                  "entropy_in_range": bool(((ent_a >= 0) & (ent_a <= math.log(V) + 1e-3)).all())}
        lp_a.sum().backward()
        rs = xx.grad[:4096].float().sum(-1)
        dist.all_reduce(rs)
        checks["grad_rows_sum_to_zero"] = bool((rs.abs() <= 3e-2).all())
        out["checks"] = checks
        return out

    def cfg2_strong(self):
        torch, dist = self.torch, self.dist
        from areal_b200 import dist as D
        T = 32768
        rng = np.random.RandomState(2)
        lens, left = [], T
        while left > 0:
            n = min(left, int(rng.randint(128, 4097)))
            lens.append(n)
            left -= n
        mine = D.partition_sequences(lens, self.world)[self.rank]
        my_lens = [lens[i] for i in mine]
        Tm = int(sum(my_lens))
        d = self._cfg2_inputs_from_lens(my_lens, seed=20 + self.rank)
        out = {"workload": f"BASELINE.json configs[2]: the 32,768-token batch ({len(lens)} sequences) split by sequence over "
                           f"{self.world} GPUs (areal_b200.dist.partition_sequences), per rank: padded GAE + packed GAE + "
                           "actor loss through the engine loop + critic loss, fwd+bwd; strong scaling",
               "tokens_total": T, "tokens_this_rank": Tm, "scaling": "strong"}
        for mode, per_tok in (("hook", 4.0 * V), ("twopass", 6.0 * V)):
            step, tracker = self._cfg2_step_fn(d, mode)
            ms = self.time_ms(step, iters=10, warmup=3)
            tracker.export(reduce_group=dist.group.WORLD)
            out[mode] = {"ms_per_step_max_over_ranks": ms, "tokens_per_s": T / (ms * 1e-3)}
        toks = torch.tensor([float(Tm)], device=self.dev)
        mx = toks.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        out["max_tokens_on_a_rank"] = int(mx.item())
        out["hook"]["hbm_frac_busiest_rank"] = 4.0 * V * float(mx.item()) / (out["hook"]["ms_per_step_max_over_ranks"] * 1e-3) / 1e9 / self.peak
        return out

    def _cfg2_inputs_from_lens(self, lens, seed):
        """as _cfg2_inputs, for a given list of sequence lengths"""
        torch = self.torch
        from areal_b200 import functional as F
        T, B, S = int(sum(lens)), len(lens), 4096
        cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(self.dev)
        x, g = self.logits(T, V, seed)
        lab = torch.randint(0, V, (T,), generator=g, device=self.dev)
        with torch.no_grad():
            lp = F.gather_logprobs(x, lab, 1.0)
        return dict(lens=lens, B=B, S=S, cu=cu, logits=x.requires_grad_(), labels=lab,
                    prox=lp + 0.02 * torch.randn(T, generator=g, device=self.dev),
                    old=lp + 0.05 * torch.randn(T, generator=g, device=self.dev),
                    mask=torch.rand(T, generator=g, device=self.dev) > 0.2,
                    rewards=torch.randn(T, generator=g, device=self.dev),
                    values=torch.randn(T + B, generator=g, device=self.dev),
                    boot=torch.rand(B, generator=g, device=self.dev) > 0.5,
                    new_values=torch.randn(T, generator=g, device=self.dev, requires_grad=True),
                    pad_rewards=torch.randn(B, S, generator=g, device=self.dev),
                    pad_values=torch.randn(B, S, generator=g, device=self.dev),
                    pad_mask=(torch.arange(S, device=self.dev)[None, :] < torch.tensor(lens, device=self.dev)[:, None]).float(),
                    no_eos=torch.rand(B, generator=g, device=self.dev) > 0.5)
