"""BASELINE.json configs[3] shape: T = 131,072 packed tokens, V = 151,936 bf16 logits vocab-sharded over the
ranks of one node (Megatron-TP layout).  Per step: partial statistics of the local [T, V/R] shard, exchange,
combine, local backward.  Launch: python -m torch.distributed.run --nproc-per-node R bench_vocab_parallel.py
Prints one JSON line (rank 0); time = max over ranks of the CUDA-event time.  (Older stand-alone form of the
`vocab_sharded` block of `bench.py --gpus N`, which also times the one-launch exchange and the per-step jitter.)"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from areal_b200 import dist as D
    T, V = int(os.environ.get("VP_TOKENS", 131072)), 151936
    Vs = V // world
    steps = int(os.environ.get("VP_STEPS", 20))
    g = torch.Generator(device=dev).manual_seed(rank)
    shard = torch.empty(T, Vs, dtype=torch.bfloat16, device=dev)
    for i in range(0, T, 8192):
        shard[i:i + 8192] = (torch.randn(min(8192, T - i), Vs, generator=g, device=dev) * 2).bfloat16()
    gl = torch.Generator(device=dev).manual_seed(1234)          # same labels / weights on every rank
    labels = torch.randint(0, V, (T,), generator=gl, device=dev)
    w = torch.randn(T, generator=gl, device=dev)
    w[torch.rand(T, generator=gl, device=dev) < 0.2] = 0.0
    peak = 6539.9
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    try:
        ex = D.PeerExchange(max_tokens=T)
    except Exception as exc:
        ex = None
        why = repr(exc)[:200]
    out = {"workload": f"BASELINE.json configs[3]: T={T} tokens, V={V} bf16 vocab-sharded over {world} GPUs "
                       f"(shard [T, {Vs}]), logprob+entropy forward and dlogits backward per step",
           "n_gpus": world, "steps": steps, "unit": "tokens/s", "peak_gbs": peak}
    from bench import Clocked                   # SM clock + throttle reasons while the timed loops run (rank's own GPU)
    for name, exch in (("allgather", None), ("peer_memory", ex)):
        if name == "peer_memory" and ex is None:
            out[name] = {"unavailable": why}
            continue
        x = shard.requires_grad_()
        clk = Clocked(local)
        clk.__enter__()

        def step():
            x.grad = None
            lp, ent = D.vocab_parallel_logprobs_entropy(x, labels, 1.0, exchange=exch)
            return lp, ent

        def step_bwd():
            lp, ent = step()
            (lp * w).sum().backward()
        res = {}
        for label, fn in (("fwd", step), ("fwd_bwd", step_bwd)):
            for _ in range(3):
                fn()
            dist.barrier()
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            dist.barrier()
            torch.cuda.synchronize(dev)
            ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            res[label + "_ms"] = float(ms.item())
        active = float((w != 0).float().mean().item())
        fwd_bytes = 2.0 * T * Vs
        bwd_bytes = 2.0 * T * Vs * (1.0 + active)
        res["tokens_per_s"] = T / (res["fwd_bwd_ms"] * 1e-3)
        res["fwd_hbm_frac"] = fwd_bytes / (res["fwd_ms"] * 1e-3) / 1e9 / peak
        res["fwd_bwd_hbm_frac"] = (fwd_bytes + bwd_bytes) / (res["fwd_bwd_ms"] * 1e-3) / 1e9 / peak
        clk.__exit__(None, None, None)
        res["clocks"] = clk.clocks
        out[name] = res
        if exch is not None:
            exch.check()
    # full-size properties: log-prob <= 0, 0 <= entropy <= ln V, both exchanges agree bit for bit, the
    # gradient of sum(logp) sums to ~0 over the WHOLE vocabulary row (all-reduced over the shards)
    import math
    x = shard.detach().requires_grad_()
    lp_a, ent_a = D.vocab_parallel_logprobs_entropy(x, labels, 1.0)
    checks = {"logp_le_0": bool((lp_a <= 0).all()), "entropy_in_range": bool(((ent_a >= 0) & (ent_a <= math.log(V) + 1e-3)).all())}
    if ex is not None:
        lp_b, ent_b = D.vocab_parallel_logprobs_entropy(shard.detach(), labels, 1.0, exchange=ex)
        checks["exchanges_bit_identical"] = bool(torch.equal(lp_a.detach(), lp_b) and torch.equal(ent_a.detach(), ent_b))
        ex.check()
    lp_a.sum().backward()
    rs = x.grad[:4096].float().sum(-1)
    dist.all_reduce(rs)
    checks["grad_rows_sum_to_zero"] = bool((rs.abs() <= 3e-2).all())
    out["checks"] = checks
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
