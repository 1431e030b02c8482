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
    # slice_intervals wrote into the neighbour's staging buffer over NVLink, bit-exact, or no peer mapping here
    assert f"PEERBUF_OK {world}" in r.stdout or "PEERBUF_UNAVAILABLE" in r.stdout, r.stdout[-2000:]
    print(next(line for line in r.stdout.splitlines() if line.startswith("PEERBUF_")))
