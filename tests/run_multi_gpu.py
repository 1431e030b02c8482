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
    dist.barrier()
    torch.cuda.synchronize()
    if rank == 0:
        print("MULTI_GPU_OK", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
