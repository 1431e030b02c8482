"""GPU: the allocation-free GAE entry point (areal_b200_gae1d_nolp_misalign_static).

Some users of this library cannot hand a kernel device scratch: a host written against the C ABI
without PyTorch has no caching allocator, a trainer may already have reserved all of HBM, and a
latency-critical thread cannot afford the device-wide synchronisation hidden in cudaMalloc / cudaFree.
For them this entry keeps the look-back descriptors in a fixed 65,600-byte array inside the library
(limit: 4,194,304 tokens per call).  Here it is called straight through the C ABI, twice in a row on
different batches -- the second, smaller call runs over whatever the first left in the shared array --
and each result must equal the workspace-based entry's on the same inputs.

This file is named so that pytest collects it last: a run that stops at the first failure has
already exercised every other test.  It passes alone in a fresh process (first use of the library and
of the device) and after the rest of the suite.
"""
import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
D = "cuda:0"
MAX_TOKENS = 4194304   # AREAL_B200_GAE1D_STATIC_MAX_TOKENS in include/areal_b200.h


def _batch(seed, lens):
    rng = np.random.RandomState(seed)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    T, B = int(cu[-1]), len(lens)
    dev = torch.device(D)
    return dict(
        rewards=torch.from_numpy(rng.randn(T).astype(np.float32)).to(dev),
        values=torch.from_numpy(rng.randn(T + B).astype(np.float32)).to(dev),
        cu=torch.from_numpy(cu).to(dev),
        boot=torch.from_numpy(rng.rand(B) > 0.5).to(dev),
        T=T, B=B, cu_np=cu)


def _static_call(b, gamma, lam):
    """Straight through the C ABI: the caller provides inputs and outputs, nothing else."""
    from areal_b200 import _lib
    from areal_b200._lib import call, ptr, stream_ptr
    dev = torch.device(D)
    adv = torch.full((b["T"],), float("nan"), device=dev)
    ret = torch.full((b["T"],), float("nan"), device=dev)
    with torch.cuda.device(dev):
        call("areal_b200_gae1d_nolp_misalign_static", ptr(b["rewards"]), ptr(b["values"]), ptr(b["cu"]),
             ptr(b["boot"].view(torch.uint8)), b["T"], b["B"], gamma, lam, ptr(adv), ptr(ret), stream_ptr(dev))
    return adv, ret


def _workspace_call(b, gamma, lam):
    from areal_b200 import ppo_functional as P
    return P.cugae1d_nolp_misalign_func(b["rewards"], b["values"], b["cu"], b["boot"], gamma, lam)


def _close(got, ref, cu):
    got = got.cpu().numpy().astype(np.float64)
    ref = ref.cpu().numpy().astype(np.float64)
    assert np.isfinite(got).all()
    for i in range(len(cu) - 1):
        s, e = int(cu[i]), int(cu[i + 1])
        if e > s:
            scale = max(1.0, float(np.abs(ref[s:e]).max()))
            assert float(np.abs(got[s:e] - ref[s:e]).max()) <= H.TOL_GAE * scale, i


def test_static_scratch_entry_matches_workspace_entry_twice_in_a_row():
    rng = np.random.RandomState(0)
    # call 1: many tiles (one 150k-token sequence keeps the look-back chain long) ...
    big = _batch(1, np.concatenate([[150000], rng.randint(1, 4097, 120)]))
    # ... call 2, right after it: fewer tiles, so it starts on the first call's leftovers
    small = _batch(2, rng.randint(128, 4097, 16))
    for b, (gamma, lam) in ((big, (0.999, 0.97)), (small, (1.0, 0.95))):
        adv, ret = _static_call(b, gamma, lam)
        adv_ref, ret_ref = _workspace_call(b, gamma, lam)
        _close(adv, adv_ref, b["cu_np"])
        _close(ret, ret_ref, b["cu_np"])
    # a single-tile batch and a repeat of the big one, still agreeing
    tiny = _batch(3, [5, 1, 40])
    adv, ret = _static_call(tiny, 0.9, 0.5)
    adv_ref, ret_ref = _workspace_call(tiny, 0.9, 0.5)
    _close(adv, adv_ref, tiny["cu_np"])
    adv, ret = _static_call(big, 0.999, 0.97)
    adv_ref, ret_ref = _workspace_call(big, 0.999, 0.97)
    _close(adv, adv_ref, big["cu_np"])
    _close(ret, ret_ref, big["cu_np"])


def test_static_scratch_entry_rejects_oversized_batches():
    from areal_b200 import _lib
    from areal_b200._lib import ptr, stream_ptr
    dev = torch.device(D)
    one = torch.zeros(8, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().areal_b200_gae1d_nolp_misalign_static(
            ptr(one), ptr(one), ptr(one), ptr(one), MAX_TOKENS + 1, 1, 1.0, 1.0, ptr(one), ptr(one), stream_ptr(dev))
    assert rc == 1
    msg = _lib.lib().areal_b200_last_error().decode()
    assert "limit" in msg and str(MAX_TOKENS) in msg
