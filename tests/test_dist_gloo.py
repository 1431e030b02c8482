"""CPU, world_size 2, gloo: the N>1 host logic -- sequence partitioning, the loss-normalisation
all-reduce, and the vocab-sharded all-gather + combine protocol (CUDA kernels replaced by the oracle's
CPU stand-ins, injected through `partials_fn` / `combine_fn`)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from areal_b200 import dist as D
        from oracle import cpu_oracle as O
        rng = np.random.RandomState(0)            # same data on every rank
        # ---- batch sharding: partition, local masked means, one all-reduce -> global masked mean
        lens = rng.randint(5, 400, 23)
        parts = D.partition_sequences(lens, world)
        assert sorted(sum(parts, [])) == list(range(23))
        loads = [int(lens[p].sum()) for p in parts]
        assert max(loads) - min(loads) <= int(lens.max())
        cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32))
        T = int(cu[-1])
        per_tok = torch.from_numpy(rng.randn(T).astype(np.float32))
        mask = torch.from_numpy(rng.rand(T) > 0.4)
        values = torch.from_numpy(rng.randn(T + 23).astype(np.float32))
        mine, my_cu = D.take_sequences(per_tok, cu, parts[rank])
        my_mask, _ = D.take_sequences(mask, cu, parts[rank])
        my_vals, _ = D.take_sequences(values, cu, parts[rank], extra_per_seq=1)
        assert my_vals.numel() == mine.numel() + len(parts[rank]) and int(my_cu[-1]) == mine.numel()
        i0 = parts[rank][0]
        assert torch.equal(my_vals[: lens[i0] + 1], values[int(cu[i0]) + i0: int(cu[i0 + 1]) + i0 + 1])
        cnt = my_mask.sum().float()
        local = torch.where(my_mask, mine, torch.zeros(())).sum() / torch.clamp(cnt, min=1)
        want = float(torch.where(mask, per_tok, torch.zeros(())).sum() / mask.sum())
        got = float(D.all_reduce_loss(local, cnt))
        assert abs(got - want) < 1e-5, (got, want)
        scale, total = D.global_loss_scale(cnt)
        assert float(total) == float(mask.sum())
        contrib = torch.tensor([float(local * scale)])
        dist.all_reduce(contrib)
        assert abs(float(contrib) - want) < 1e-5
        # ---- vocab sharding: partials -> all_gather -> combine == unsharded oracle
        Tn, V = 37, 1000
        logits = (rng.randn(Tn, V) * 3).astype(np.float32)
        labels = rng.randint(0, V, Tn)
        Vs = V // world
        shard = torch.from_numpy(logits[:, rank * Vs:(rank + 1) * Vs])

        def partials_fn(sh, lab, vocab_start, temp):
            return torch.from_numpy(O.shard_partials(sh.numpy(), lab.numpy(), vocab_start, temp).astype(np.float32))

        def combine_fn(allp, temp):
            assert allp.shape == (world, 4, Tn)
            return tuple(torch.from_numpy(a) for a in O.combine_partials(allp.numpy(), temp))

        logp, ent, row_max, ln_s = D.vocab_parallel_forward(shard, torch.from_numpy(labels), rank * Vs, 0.7,
                                                            partials_fn=partials_fn, combine_fn=combine_fn)
        ref_lp, ref_ent, ref_lse, _, ref_max = O.gather_logprobs_entropy(logits, labels, 0.7)
        assert np.allclose(logp.numpy(), ref_lp, atol=1e-4) and np.allclose(ent.numpy(), ref_ent, atol=1e-4)
        assert np.array_equal(row_max.numpy(), ref_max.astype(np.float64))
        assert np.allclose(row_max.numpy() / 0.7 + ln_s.numpy(), ref_lse, atol=1e-4)
        # ---- Ulysses branch of gather_logprobs (areal/utils/functional.py:76-79): differentiable
        # all-gather of the per-rank token slices, concatenated in rank order; group look-up rules
        from areal_b200 import functional as F
        F.reset_parallel_groups()
        assert F._tp_group() is None and F._sp_group() is None        # nothing imported, nothing set
        F.set_parallel_groups(sp_group=dist.group.WORLD)
        assert F._tp_group() is None and F._sp_group() is dist.group.WORLD
        mine_lp = torch.arange(5, dtype=torch.float32) + 10.0 * rank
        mine_lp.requires_grad_()
        both = F._sp_all_gather(mine_lp, F._sp_group())
        assert torch.equal(both.detach(), torch.cat([torch.arange(5.0), torch.arange(5.0) + 10.0]))
        wts = torch.arange(10, dtype=torch.float32) + 1.0
        (both * wts).sum().backward()                                  # every rank back-propagates the same loss
        assert torch.equal(mine_lp.grad, world * wts[rank * 5:(rank + 1) * 5])
        F.set_parallel_groups()                                        # pinned off
        assert F._sp_group() is None
        F.reset_parallel_groups()
        out[rank] = 1
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo():
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(out) == {0: 1, 1: 1}


def test_partition_sequences_edge_cases():
    from areal_b200 import dist as D
    assert D.partition_sequences([], 4) == [[], [], [], []]
    assert D.partition_sequences([7], 2) == [[0], []]
    p = D.partition_sequences([5, 5, 5, 5, 100], 2)
    assert sorted(sum(p, [])) == [0, 1, 2, 3, 4] and [4] in p
    with pytest.raises(ValueError):
        D.partition_sequences([1], 0)
