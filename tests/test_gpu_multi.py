"""GPU, N > 1: NCCL runs of the batch-sharded and vocab-sharded paths (spawned through
torch.distributed.run on 127.0.0.1, one rank per GPU).  Skipped on a single-GPU box; the world_size-2
gloo test (tests/test_dist_gloo.py) covers the host logic everywhere."""
import os
import subprocess
import sys

import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu


def test_single_rank_vocab_parallel_equals_unsharded():
    """dist.vocab_parallel_logprobs_entropy with a world of 1: partial kernel + combine on one GPU."""
    from areal_b200 import dist as D, functional as F
    g = torch.Generator().manual_seed(1)
    logits = (torch.randn(50, 18992, generator=g) * 2).bfloat16().to("cuda:0")
    labels = torch.randint(0, 18992, (50,), generator=g).to("cuda:0")
    a = logits.clone().requires_grad_()
    b = logits.clone().requires_grad_()
    lp, ent = D.vocab_parallel_logprobs_entropy(a, labels, 0.8)
    lp1, ent1 = F.gather_logprobs_entropy(b, labels, 0.8)
    assert torch.allclose(lp, lp1, rtol=1e-5, atol=1e-5) and torch.allclose(ent, ent1, rtol=1e-5, atol=1e-5)
    w = torch.randn(50, device="cuda:0")
    (lp * w).sum().backward()
    (lp1 * w).sum().backward()
    assert torch.allclose(a.grad.float(), b.grad.float(), rtol=2 ** -7, atol=1e-9)


def test_tracker_scalar_only_export_over_nccl_world1():
    """export_all(reduce_group=NCCL group) with only SCALAR keys recorded after a reset (the reference's
    example loop): the scalar table must be reduced on the GPU.  One rank, one GPU; the same check runs
    inside tests/run_multi_gpu.py at world 2/4/8."""
    r = subprocess.run([sys.executable, os.path.join(H.ROOT, "tests", "run_tracker_nccl.py")], capture_output=True,
                       text=True, timeout=300, env=dict(os.environ, MASTER_PORT="29541", RANK="0", WORLD_SIZE="1",
                                                        LOCAL_RANK="0"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "TRACKER_NCCL_OK 1" in r.stdout


@pytest.mark.parametrize("world", [2, 4, 8])
def test_nccl_sharded_paths(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    port = 29500 + world
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(H.ROOT, "tests", "run_multi_gpu.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert f"MULTI_GPU_OK {world}" in r.stdout
    # the NVLink peer-memory exchange ran bit-identically to the all-gather path, or the box cannot map peers
    assert f"P2P_OK {world}" in r.stdout or "P2P_UNAVAILABLE" in r.stdout, r.stdout[-2000:]
    print(next(line for line in r.stdout.splitlines() if line.startswith("P2P_")))
    # ... and so did the one-launch variant (partials pushed from the streaming kernel's row epilogue)
    assert f"P2P_PUSH_OK {world}" in r.stdout or "P2P_PUSH_UNAVAILABLE" in r.stdout, r.stdout[-2000:]
    print(next(line for line in r.stdout.splitlines() if line.startswith("P2P_PUSH_")))
    # slice_intervals wrote into the neighbour's staging buffer over NVLink, bit-exact, or no peer mapping here
    assert f"PEERBUF_OK {world}" in r.stdout or "PEERBUF_UNAVAILABLE" in r.stdout, r.stdout[-2000:]
    print(next(line for line in r.stdout.splitlines() if line.startswith("PEERBUF_")))


def test_vocab_sharded_kernels_vs_reference_golden():
    """The per-shard partial kernel, the combine kernel and the sharded backward, run shard by shard on ONE GPU,
    against golden vectors of the reference's own vocabulary-sharded code on two gloo ranks
    (tests/golden/vocab_parallel.npz, oracle/gen_golden_tp.py: `_VocabParallelEntropy`,
    areal/utils/mcore/functional.py:8-52, and the tree's copy of Megatron's vocab_parallel_cross_entropy,
    realhf/impl/model/parallelism/tensor_parallel/modules.py:1060-1186)."""
    import numpy as np
    from areal_b200 import dist as D
    from areal_b200.functional import _logits_backward
    z = H.load("vocab_parallel.npz")
    world = int(z["world"])
    dev = "cuda:0"
    for k, c in H.cases(z):
        logits = H.bf16_from_bits(c["logits_bits"]).to(dev)
        labels = torch.from_numpy(c["labels"]).to(dev)
        temp = float(c["temp"])
        Vs = logits.shape[1] // world
        shards = [logits[:, r * Vs:(r + 1) * Vs].contiguous() for r in range(world)]
        parts = torch.stack([D._cuda_partials(shards[r], labels, r * Vs, temp) for r in range(world)])
        logp, ent, row_max, ln_s = D._cuda_combine(parts, temp)
        assert np.allclose(logp.cpu().numpy(), c["logp"], rtol=H.TOL_LOGITS, atol=H.TOL_LOGITS), k
        assert np.allclose(ent.cpu().numpy(), c["entropy"], rtol=H.TOL_LOGITS, atol=H.TOL_LOGITS), k
        g_lp, g_h = torch.from_numpy(c["g_lp"]).to(dev), torch.from_numpy(c["g_h"]).to(dev)
        slack = 1e-6 * (np.abs(c["g_lp"]) + 12 * np.abs(c["g_h"]))[:, None] / temp
        for r in range(world):
            d = _logits_backward(shards[r], labels, temp, row_max, ln_s, ent, g_lp, g_h, vocab_start=r * Vs)
            got = d.float().cpu().numpy().astype(np.float64)
            want = H.bf16_from_bits(c[f"grad{r}_bits"]).double().numpy()
            # both are bf16 roundings of the same gradient (the reference's with its fp32 cancellation at the label)
            assert np.all(np.abs(got - want) <= 2 ** -6 * np.abs(want) + slack + 1e-30), (k, r)


def test_vocab_parallel_backward_keeps_the_input_shape():
    """A [B, S, V/R] shard (more than two dimensions) gets a gradient of its own shape."""
    from areal_b200 import dist as D
    g = torch.Generator().manual_seed(2)
    x = (torch.randn(3, 7, 4096, generator=g) * 2).bfloat16().to("cuda:0").requires_grad_()
    labels = torch.randint(0, 4096, (3, 7), generator=g).to("cuda:0")
    lp, ent = D.vocab_parallel_logprobs_entropy(x, labels, 1.0)
    (lp.sum() + ent.sum()).backward()
    assert x.grad.shape == x.shape
    y = x.detach().reshape(21, 4096).clone().requires_grad_()
    lp2, ent2 = D.vocab_parallel_logprobs_entropy(y, labels.reshape(-1), 1.0)
    (lp2.sum() + ent2.sum()).backward()
    assert torch.equal(x.grad.reshape(21, 4096), y.grad)
