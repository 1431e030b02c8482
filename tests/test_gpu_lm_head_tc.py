"""GPU parity: the tcgen05 LM-head kernel (areal_b200_lm_head_logprobs: projection on the tensor cores,
online softmax as the GEMM's epilogue, no [T,V] tensor) against

  * golden vectors of the unmodified reference run from the hidden state (tests/golden/lm_head.npz:
    F.linear -> gather_logprobs_entropy), tolerance 1e-3 (north star, bf16 in / fp32 accumulate);
  * the oracle on inputs that sit on binary grids, so that every dot product is exact in fp32 and the
    bf16 logits are the same bits whatever the summation order: ragged T, V that is not a multiple of
    the 256-column tile nor of the 32-column epilogue chunk, the Qwen vocabulary, several hidden sizes,
    temperature != 1, labels in the first / last column;
  * the library path (cuBLAS GEMM per row block + streaming kernel) on random bf16 inputs, where the two
    fp32 summation orders may round a logit to neighbouring bf16 values: tolerance one bf16 ulp of the
    label's logit on the log-prob.
Both CTA-pair (cta_group::2) and single-CTA (cta_group::1) instances are run (the second in a subprocess:
the choice is read once per process from AREAL_B200_LMHEAD_CTA_GROUP)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import cpu_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu
D = "cuda:0"


def grid_inputs(T, Hd, V, seed):
    g = torch.Generator().manual_seed(seed)
    h = (torch.randint(-16, 17, (T, Hd), generator=g).float() / 8).bfloat16()
    w = (torch.randint(-16, 17, (V, Hd), generator=g).float() / 32).bfloat16()
    labels = torch.randint(0, V, (T,), generator=g)
    labels[0], labels[-1] = 0, V - 1
    return h, w, labels


def check_vs_oracle(T, Hd, V, temp, seed):
    from areal_b200.lm_head import lm_head_logprobs_tc
    h, w, labels = grid_inputs(T, Hd, V, seed)
    logits = (h.double() @ w.double().t()).float().bfloat16()          # exact products, rounded once
    lp_ref, ent_ref, _, mn_ref, mx_ref = O.gather_logprobs_entropy(logits.float().numpy(), labels.numpy(), temp)
    st = lm_head_logprobs_tc(h.to(D), w.to(D), labels.to(D), temp, with_stats=True)
    assert H.rel_err(st["logp"].cpu().numpy(), lp_ref) <= H.TOL_LOGITS, (T, Hd, V)
    assert H.rel_err(st["entropy"].cpu().numpy(), ent_ref) <= H.TOL_LOGITS, (T, Hd, V)
    assert np.array_equal(st["vocab_max_logits"].cpu().numpy(), mx_ref.astype(np.float32))
    assert np.array_equal(st["vocab_min_logits"].cpu().numpy(), mn_ref.astype(np.float32))
    # measured accuracy is far inside the bar
    assert H.rel_err(st["logp"].cpu().numpy(), lp_ref) <= 2e-5
    lp_only = lm_head_logprobs_tc(h.to(D), w.to(D), labels.to(D), temp)
    assert torch.equal(lp_only, st["logp"])


def test_golden_from_the_reference():
    from areal_b200.lm_head import lm_head_logprobs_tc
    z = H.load("lm_head.npz")
    for k, c in H.cases(z):
        h = H.bf16_from_bits(c["h_bits"]).to(D)
        w = H.bf16_from_bits(c["w_bits"]).to(D)
        lab = torch.from_numpy(c["labels"]).to(D)
        lp, ent = lm_head_logprobs_tc(h, w, lab, float(c["temp"]), with_entropy=True)
        assert H.rel_err(lp.cpu().numpy(), c["logp"]) <= H.TOL_LOGITS, k
        assert H.rel_err(ent.cpu().numpy(), c["entropy"]) <= H.TOL_LOGITS, k


@pytest.mark.parametrize("T,Hd,V,temp", [(256, 64, 512, 1.0), (128, 128, 256, 1.0), (300, 192, 5000, 0.9),
                                        (1, 64, 4104, 1.0), (517, 1536, 8200, 0.7), (70, 64, 151936, 1.0),
                                        (2500, 256, 33000, 1.0), (257, 72, 300, 1.0)])
def test_exact_inputs_vs_oracle(T, Hd, V, temp):
    check_vs_oracle(T, Hd, V, temp, seed=T + V)


def test_random_inputs_vs_library_path():
    from areal_b200.lm_head import lm_head_logprobs, lm_head_logprobs_tc
    g = torch.Generator().manual_seed(9)
    T, Hd, V = 1024, 1536, 20000
    h = torch.randn(T, Hd, generator=g).bfloat16().to(D)
    w = (torch.randn(V, Hd, generator=g) * (2.0 / Hd ** 0.5)).bfloat16().to(D)
    lab = torch.randint(0, V, (T,), generator=g).to(D)
    a_lp, a_h = lm_head_logprobs_tc(h, w, lab, 1.0, with_entropy=True)
    b_lp, b_h = lm_head_logprobs(h, w, lab, 1.0, chunk_rows=512, with_entropy=True, impl="library")
    c_lp, c_h = lm_head_logprobs(h, w, lab, 1.0, with_entropy=True)           # "auto" takes the tensor-core kernel
    assert torch.equal(c_lp, a_lp) and torch.equal(c_h, a_h)
    x_lab = (h.float() @ w.float().t()).gather(1, lab[:, None])[:, 0]
    tol = 2.0 ** -7 * x_lab.abs() + 2e-3            # the two GEMMs may round a logit to neighbouring bf16 values
    assert bool(((a_lp - b_lp).abs() <= tol).all()), float((a_lp - b_lp).abs().max())
    assert float(((a_lp - b_lp).abs() <= 1e-4).float().mean()) > 0.95      # ... and almost never do
    assert torch.allclose(a_h, b_h, rtol=2e-3, atol=2e-3)


def test_auto_dispatch_by_hidden_size():
    """`impl="auto"`: the tensor-core kernel up to a hidden size of 2,048, the row-blocked library path above."""
    from areal_b200.lm_head import lm_head_logprobs, lm_head_logprobs_tc
    h, w, labels = grid_inputs(96, 2304, 4096, 11)
    a = lm_head_logprobs(h.to(D), w.to(D), labels.to(D), 1.0)
    b = lm_head_logprobs(h.to(D), w.to(D), labels.to(D), 1.0, impl="library")
    assert torch.equal(a, b)
    c = lm_head_logprobs_tc(h.to(D), w.to(D), labels.to(D), 1.0)              # exact inputs: same logits either way
    assert H.rel_err(c.cpu().numpy(), a.cpu().numpy()) <= 2e-5
    h2, w2, l2 = grid_inputs(96, 2048, 4096, 12)
    assert torch.equal(lm_head_logprobs(h2.to(D), w2.to(D), l2.to(D), 1.0), lm_head_logprobs_tc(h2.to(D), w2.to(D), l2.to(D), 1.0))


def test_strided_rows_and_bad_arguments():
    from areal_b200 import _lib
    from areal_b200.lm_head import lm_head_logprobs_tc
    h, w, labels = grid_inputs(200, 64, 1024, 5)
    big_h = torch.zeros(200, 128, dtype=torch.bfloat16, device=D)
    big_h[:, :64] = h.to(D)
    a = lm_head_logprobs_tc(big_h[:, :64], w.to(D), labels.to(D), 1.0)        # row stride 128, still one TMA map
    b = lm_head_logprobs_tc(h.to(D), w.to(D), labels.to(D), 1.0)
    assert torch.equal(a, b)
    bad = labels.clone()
    bad[3] = 5000
    c = lm_head_logprobs_tc(h.to(D), w.to(D), bad.to(D), 1.0)
    assert torch.isnan(c[3]) and torch.equal(c[4:], b[4:])
    with pytest.raises(TypeError):
        lm_head_logprobs_tc(h.float().to(D), w.float().to(D), labels.to(D), 1.0)
    with pytest.raises(_lib.ArealB200Error):
        lm_head_logprobs_tc(h, w, labels, 1.0)                                # CPU tensors: no fallback


def test_single_cta_instance_in_a_subprocess():
    """cta_group::1 (128 x 256 tiles, one CTA) -- the pair instance is the default everywhere else."""
    code = ("import sys; sys.path.insert(0, %r); from tests.test_gpu_lm_head_tc import check_vs_oracle; "
            "check_vs_oracle(300, 192, 5000, 0.9, 1); check_vs_oracle(517, 1536, 8200, 0.7, 2); print('CG1_OK')" % H.ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, AREAL_B200_LMHEAD_CTA_GROUP="1"))
    assert r.returncode == 0 and "CG1_OK" in r.stdout, r.stdout[-1500:] + r.stderr[-3000:]
