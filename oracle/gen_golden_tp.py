"""TEST INFRASTRUCTURE ONLY (build container).  Golden vectors for the vocabulary-sharded log-prob / entropy path
from the reference's own sharded code, run on two gloo ranks on the CPU: what `gather_logprobs_entropy` executes
under Megatron tensor parallelism (areal/utils/functional.py:91-98) --
    _logits = logits.float() / temperature
    entropy = _VocabParallelEntropy.apply(_logits)                       (areal/utils/mcore/functional.py:8-52)
    logprobs = -vocab_parallel_cross_entropy(_logits, labels)           (the tree's copy of Megatron's function,
                                                                         realhf/.../tensor_parallel/modules.py:1060-1186)
-- and its backward to each rank's shard.  Writes tests/golden/vocab_parallel.npz.
usage: python oracle/gen_golden_tp.py"""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")
WORLD = 2
CASES = ((40, 4096, 1.0, 2.0), (6, 37984, 0.7, 2.0), (16, 2048, 0.9, 8.0))     # T, V (all ranks), temperature, logit scale


def case_inputs(k):
    T, V, temp, scale = CASES[k]
    g = torch.Generator().manual_seed(7100 + k)
    logits = (torch.randn(T, V, generator=g) * scale).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g)
    labels[0], labels[1] = 0, V - 1
    logits[2, labels[2]] += 25.0                         # a peaked row
    return logits, labels, temp, torch.randn(T, generator=g), torch.randn(T, generator=g) * 0.1


def worker(rank, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    torch.set_num_threads(4)
    from oracle import ref_loader as R
    Entropy, cross_entropy, restore = R.load_vocab_parallel(dist.group.WORLD, rank, WORLD)
    try:
        res = {}
        for k in range(len(CASES)):
            logits, labels, temp, g_lp, g_h = case_inputs(k)
            Vs = logits.shape[1] // WORLD
            shard = logits[:, rank * Vs:(rank + 1) * Vs].clone().requires_grad_()
            scaled = shard.float() / temp                        # functional.py:93
            entropy = Entropy.apply(scaled)                      # :94
            logprobs = -cross_entropy(scaled, labels)            # :95-97
            ((logprobs * g_lp).sum() + (entropy * g_h).sum()).backward()
            res[k] = dict(logp=logprobs.detach().numpy(), entropy=entropy.detach().numpy(),
                          grad=shard.grad.view(torch.int16).numpy().view(np.uint16))
        out[rank] = res
    finally:
        restore()
        dist.destroy_process_group()


def main():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    procs = [ctx.Process(target=worker, args=(r, port, out)) for r in range(WORLD)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=600)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    z = {"n_cases": np.array(len(CASES)), "world": np.array(WORLD)}
    for k in range(len(CASES)):
        logits, labels, temp, g_lp, g_h = case_inputs(k)
        a, b = out[0][k], out[1][k]
        assert np.array_equal(a["logp"], b["logp"]) and np.array_equal(a["entropy"], b["entropy"])   # replicated results
        p = f"c{k}_"
        z.update({p + "logits_bits": logits.view(torch.int16).numpy().view(np.uint16), p + "labels": labels.numpy(),
                  p + "temp": np.array(temp), p + "g_lp": g_lp.numpy(), p + "g_h": g_h.numpy(),
                  p + "logp": a["logp"], p + "entropy": a["entropy"],
                  p + "grad0_bits": a["grad"], p + "grad1_bits": b["grad"]})
    np.savez_compressed(os.path.join(OUT, "vocab_parallel.npz"), **z)
    print("vocab_parallel golden cases:", len(CASES))


if __name__ == "__main__":
    main()
