"""torchrun worker for tests/test_gpu_multi.py (pattern of the reference's
areal/tests/torchrun/run_*.py): vocab-sharded and batch-sharded paths over NCCL vs the single-GPU
kernels on the full problem.  Exits non-zero on any mismatch."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from areal_b200 import dist as D
    from areal_b200 import functional as F
    from areal_b200 import ppo_functional as P

    # ---------------- vocab-sharded (cfg 4 in miniature): V = 151,936 split over `world` ranks ---------
    g = torch.Generator(device=dev).manual_seed(7)          # same stream on every rank
    T, V = 96, 151936
    logits = (torch.randn(T, V, generator=g, device=dev) * 2).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g, device=dev)
    w_lp = torch.randn(T, generator=g, device=dev)
    w_h = torch.randn(T, generator=g, device=dev) * 0.1
    Vs = V // world
    shard = logits[:, rank * Vs:(rank + 1) * Vs].detach().clone().requires_grad_()
    lp, ent = D.vocab_parallel_logprobs_entropy(shard, labels, 0.9)
    ((lp * w_lp).sum() + (ent * w_h).sum()).backward()
    full = logits.detach().clone().requires_grad_()
    lp1, ent1 = F.gather_logprobs_entropy(full, labels, 0.9)
    ((lp1 * w_lp).sum() + (ent1 * w_h).sum()).backward()
    assert torch.allclose(lp, lp1, rtol=1e-5, atol=2e-5), (lp - lp1).abs().max()
    assert torch.allclose(ent, ent1, rtol=1e-5, atol=2e-5), (ent - ent1).abs().max()
    gs = full.grad[:, rank * Vs:(rank + 1) * Vs].float()
    assert torch.allclose(shard.grad.float(), gs, rtol=2 ** -7, atol=1e-9), (shard.grad.float() - gs).abs().max()

    # ---------------- the two distributed branches inside gather_logprobs(_entropy) ---------------------
    # (areal/utils/functional.py:50-58, 91-98: vocabulary shards; :76-79, 120-125: Ulysses token slices)
    F.set_parallel_groups(tp_group=dist.group.WORLD)
    shard_tp = logits[:, rank * Vs:(rank + 1) * Vs].detach().clone().requires_grad_()
    lp_tp, ent_tp = F.gather_logprobs_entropy(shard_tp, labels, 0.9)
    ((lp_tp * w_lp).sum() + (ent_tp * w_h).sum()).backward()
    assert torch.equal(lp_tp, lp) and torch.equal(ent_tp, ent) and torch.equal(shard_tp.grad, shard.grad)
    assert torch.equal(F.gather_logprobs(shard_tp.detach(), labels, 0.9), lp.detach())
    F.set_parallel_groups(sp_group=dist.group.WORLD)
    Tr = T // world
    rows = logits[rank * Tr:(rank + 1) * Tr].detach().clone().requires_grad_()
    lp_sp, ent_sp = F.gather_logprobs_entropy(rows, labels[rank * Tr:(rank + 1) * Tr], 0.9)
    assert lp_sp.shape == (Tr * world,)
    assert torch.equal(lp_sp.detach(), lp1.detach()[:Tr * world]) and torch.equal(ent_sp.detach(), ent1.detach()[:Tr * world])
    ((lp_sp * w_lp[:Tr * world]).sum() + (ent_sp * w_h[:Tr * world]).sum()).backward()
    # every rank back-propagates the same (replicated) loss, so the gathered gradient is world x the local one
    want_rows = full.grad[rank * Tr:(rank + 1) * Tr].float() * world
    assert torch.allclose(rows.grad.float(), want_rows, rtol=2 ** -6, atol=1e-9), (rows.grad.float() - want_rows).abs().max()
    F.reset_parallel_groups()

    # ---------------- the same exchange over NVLink peer memory (no all-gather) -------------------------
    try:
        ex = D.PeerExchange(max_tokens=4096)
    except Exception as exc:                                  # no peer mapping on this box: reported, not fatal
        ex = None
        if rank == 0:
            print("P2P_UNAVAILABLE", repr(exc)[:300])
    if ex is not None:
        for step, Tp in enumerate((96, 1, 4096, 777, 0, 96)):   # both slots several times, ragged and empty T
            gp = torch.Generator(device=dev).manual_seed(100 + step)
            lg_p = (torch.randn(max(Tp, 1), V, generator=gp, device=dev) * 2).bfloat16()[:Tp]
            lab_p = torch.randint(0, V, (Tp,), generator=gp, device=dev)
            sh = lg_p[:, rank * Vs:(rank + 1) * Vs].contiguous()
            a = D.vocab_parallel_forward(sh, lab_p, rank * Vs, 0.9) if Tp else None
            b = ex.forward(sh, lab_p, rank * Vs, 0.9)
            if Tp:
                for u, v in zip(a, b):
                    assert torch.equal(u, v), (step, Tp, (u - v).abs().max())
        ex.check()
        # autograd path through the exchange
        shard2 = logits[:, rank * Vs:(rank + 1) * Vs].detach().clone().requires_grad_()
        lp2, ent2 = D.vocab_parallel_logprobs_entropy(shard2, labels, 0.9, exchange=ex)
        ((lp2 * w_lp).sum() + (ent2 * w_h).sum()).backward()
        assert torch.equal(lp2, lp) and torch.equal(ent2, ent) and torch.equal(shard2.grad, shard.grad)
        # device time of one vocab-sharded forward (T = 4096, V/world columns), both ways
        Tp = 4096
        sh = (torch.randn(Tp, Vs, generator=g, device=dev) * 2).bfloat16()
        lab_p = torch.randint(0, V, (Tp,), generator=g, device=dev)

        def t_nccl():
            D.vocab_parallel_forward(sh, lab_p, rank * Vs, 0.9)   # partial kernel, all-gather, combine

        def t_p2p():
            ex.forward(sh, lab_p, rank * Vs, 0.9)                 # partial kernel, peer-memory combine
        res = {}
        for name, fn in (("partial+allgather+combine", t_nccl), ("partial+peer_combine", t_p2p)):
            for _ in range(5):
                fn()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(50):
                fn()
            e1.record()
            torch.cuda.synchronize()
            res[name] = e0.elapsed_time(e1) / 50 * 1e3
        ex.check()
        if rank == 0:
            print("P2P_OK", world, {k: round(v, 1) for k, v in res.items()}, "us per call")

    # ---------------- ONE launch: partials pushed into every rank's buffer by the streaming kernel, combined there ----
    try:
        fx = D.FusedPeerExchange(max_tokens=4096)
    except Exception as exc:                                  # no peer mapping on this box: reported, not fatal
        fx = None
        if rank == 0:
            print("P2P_PUSH_UNAVAILABLE", repr(exc)[:300])
    if fx is not None:
        for step, Tp in enumerate((96, 1, 4096, 777, 0, 96, 3000, 5)):   # both slots several times, ragged and empty T
            gp = torch.Generator(device=dev).manual_seed(200 + step)
            lg_p = (torch.randn(max(Tp, 1), V, generator=gp, device=dev) * 2).bfloat16()[:Tp]
            lab_p = torch.randint(0, V, (Tp,), generator=gp, device=dev)
            sh = lg_p[:, rank * Vs:(rank + 1) * Vs].contiguous()
            a = D.vocab_parallel_forward(sh, lab_p, rank * Vs, 0.9) if Tp else None
            mn = torch.empty(Tp, dtype=torch.float32, device=dev)
            b = fx.forward(sh, lab_p, rank * Vs, 0.9, raw_min=mn)
            if Tp:
                for u, v in zip(a, b):
                    assert torch.equal(u, v), (step, Tp, (u - v).abs().max())
                assert torch.equal(mn, sh.float().min(dim=-1).values), (step, Tp)
        fx.check()
        # autograd path through the fused exchange
        shard3 = logits[:, rank * Vs:(rank + 1) * Vs].detach().clone().requires_grad_()
        lp3, ent3 = D.vocab_parallel_logprobs_entropy(shard3, labels, 0.9, exchange=fx)
        ((lp3 * w_lp).sum() + (ent3 * w_h).sum()).backward()
        assert torch.equal(lp3, lp) and torch.equal(ent3, ent) and torch.equal(shard3.grad, shard.grad)
        Tp = 4096
        sh = (torch.randn(Tp, Vs, generator=g, device=dev) * 2).bfloat16()
        lab_p = torch.randint(0, V, (Tp,), generator=g, device=dev)
        for _ in range(5):
            fx.forward(sh, lab_p, rank * Vs, 0.9)
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            fx.forward(sh, lab_p, rank * Vs, 0.9)
        e1.record()
        torch.cuda.synchronize()
        fx.check()
        if rank == 0:
            print("P2P_PUSH_OK", world, round(e0.elapsed_time(e1) / 50 * 1e3, 1), "us per call (one launch)")

    # ---------------- parameter reallocation: slice straight into the neighbour's staging buffer -----------
    from areal_b200 import flatten_param as FP
    try:
        pb = D.PeerPackedBuffer(numel=200_000, dtype=torch.bfloat16)
    except Exception as exc:                                  # no peer mapping on this box: reported, not fatal
        pb = None
        if rank == 0:
            print("PEERBUF_UNAVAILABLE", repr(exc)[:300])
    if pb is not None:
        gsrc = torch.Generator(device=dev).manual_seed(55)    # same stream on every rank: row r = rank r's parameters
        flat_all = torch.randn(world, 400_000, generator=gsrc, device=dev).bfloat16()
        rs = np.random.RandomState(9)
        sizes = rs.randint(1, 300, 900)
        starts = np.cumsum(sizes + rs.randint(0, 100, 900)) - sizes
        iv_np = np.stack([starts, starts + sizes], 1).astype(np.int64)
        rs.shuffle(iv_np)
        total = int(sizes.sum())
        ivc = torch.from_numpy(iv_np).to(dev)
        nxt, prv = (rank + 1) % world, (rank - 1) % world
        for offset in (16, 3):                                # 16-byte aligned, then a 6-byte shift (funnel path)
            pb.slice_into(nxt, flat_all[rank].contiguous(), ivc, total, offset=offset)
            pb.fence()
            want = FP.slice_intervals_any(flat_all[prv].contiguous(), ivc, total)   # what my left neighbour sent
            got = pb.local()[offset: offset + total]
            assert torch.equal(got.view(torch.int16), want.view(torch.int16)), offset
            mine, ref = torch.zeros(400_000, dtype=torch.bfloat16, device=dev), torch.zeros(400_000, dtype=torch.bfloat16, device=dev)
            pb.set_from_local(mine, ivc, total, offset=offset)
            FP.set_intervals_any(want, ref, ivc)
            assert torch.equal(mine.view(torch.int16), ref.view(torch.int16))
            pb.fence(host=(offset == 3))                      # receivers are done before the next round overwrites
        pb.check()                                            # (device-side fences: nobody timed out)
        # device time of a 256 MB re-shard write into the neighbour (128 Mi bf16 in 4,096 pieces), against the same
        # slice into local memory
        big = D.PeerPackedBuffer(numel=1 << 27, dtype=torch.bfloat16)
        flat_big = torch.empty(1 << 28, dtype=torch.bfloat16, device=dev)
        flat_big.view(torch.int16).random_(-30000, 30000)
        piece = (1 << 28) // 4096
        iv_big = torch.stack([torch.arange(4096, device=dev) * piece, torch.arange(4096, device=dev) * piece + piece // 2], 1)
        res = {}
        for name, target in (("into the neighbour over NVLink", nxt), ("into local memory", rank)):
            for _ in range(3):
                big.slice_into(target, flat_big, iv_big, 1 << 27)
            big.fence()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                big.slice_into(target, flat_big, iv_big, 1 << 27)
            e1.record()
            big.fence()
            torch.cuda.synchronize()                          # (the fence is stream-ordered, not a host sync)
            res[name] = round(e0.elapsed_time(e1) / 10, 3)
        big.check()
        if rank == 0:
            print("PEERBUF_OK", world, {k: f"{v} ms ({0.268435456 / v * 1e3:.0f} GB/s written)" for k, v in res.items()})

    # ---------------- batch-sharded (cfg 3 in miniature) ------------------------------------------------
    rng = np.random.RandomState(3)
    lens = rng.randint(16, 200, 24)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
    Tn, Vn = int(cu[-1]), 4096
    lg = (torch.randn(Tn, Vn, generator=g, device=dev) * 2).bfloat16()
    lab = torch.randint(0, Vn, (Tn,), generator=g, device=dev)
    prox = -torch.rand(Tn, generator=g, device=dev) * 9
    old = prox + 0.05 * torch.randn(Tn, generator=g, device=dev)
    adv = torch.randn(Tn, generator=g, device=dev)
    mask = torch.rand(Tn, generator=g, device=dev) > 0.3
    rewards = torch.randn(Tn, generator=g, device=dev)
    values = torch.randn(Tn + 24, generator=g, device=dev)
    boot = torch.rand(24, generator=g, device=dev) > 0.5
    cfg = dict(eps_clip=0.4, behav_imp_weight_cap=5.0)
    # single GPU reference on the whole batch
    x_full = lg.clone().requires_grad_()
    loss_full, st_full = F.fused_actor_loss(x_full, lab, 1.0, prox, old, adv, mask, **cfg)
    loss_full.backward()
    a_full, r_full = P.cugae1d_nolp_misalign_func(rewards, values, cu, boot, 0.99, 0.95)
    # my shard
    mine = D.partition_sequences(lens, world)[rank]
    take = lambda t, extra=0: D.take_sequences(t, cu, mine, extra)[0]
    my_cu = D.take_sequences(rewards, cu, mine)[1]
    x = take(lg).clone().requires_grad_()
    my_mask = take(mask)
    loss, st = F.fused_actor_loss(x, take(lab), 1.0, take(prox), take(old), take(adv), my_mask, **cfg)
    scale, total = D.global_loss_scale(my_mask.sum())
    (loss * scale).backward()
    glob = D.all_reduce_loss(loss, my_mask.sum())
    assert abs(float(glob) - float(loss_full)) <= 1e-5 * max(1.0, abs(float(loss_full))), (float(glob), float(loss_full))
    assert torch.allclose(x.grad.float(), take(x_full.grad).float(), rtol=2 ** -7, atol=1e-9)
    assert torch.equal(st["logprobs"], take(st_full["logprobs"]))
    a, r = P.cugae1d_nolp_misalign_func(take(rewards), take(values, 1), my_cu, boot[torch.tensor(mine, device=dev)], 0.99, 0.95)
    assert torch.equal(a, take(a_full)) and torch.equal(r, take(r_full))
    # ---------------- stats export over the NCCL group with scalar keys only (after a reset) -------------
    from tests.run_tracker_nccl import check_scalar_only_export
    check_scalar_only_export(dist.group.WORLD)
    dist.barrier()
    torch.cuda.synchronize()
    if rank == 0:
        print("MULTI_GPU_OK", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
