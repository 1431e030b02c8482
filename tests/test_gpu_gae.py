"""GPU parity: packed / padded / 2D GAE through the C ABI vs the pinned oracle and golden vectors."""
import numpy as np
import pytest
import torch

from oracle import c_oracle as C
from oracle import cpu_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda:0")


def run1d(rewards, values, cu, boot, gamma, lam):
    from areal_b200 import ppo_functional as P
    d = dev()
    adv, ret = P.cugae1d_nolp_misalign_func(
        torch.from_numpy(np.ascontiguousarray(rewards)).to(d), torch.from_numpy(np.ascontiguousarray(values)).to(d),
        torch.from_numpy(np.ascontiguousarray(cu, dtype=np.int32)).to(d),
        torch.from_numpy(np.ascontiguousarray(boot).astype(bool)).to(d), gamma, lam)
    return adv.cpu().numpy(), ret.cpu().numpy()


def per_seq_rel(got, ref, cu):
    worst = 0.0
    for i in range(len(cu) - 1):
        s, e = int(cu[i]), int(cu[i + 1])
        if e > s:
            scale = max(1.0, float(np.abs(ref[s:e]).max()))
            worst = max(worst, float(np.abs(got[s:e].astype(np.float64) - ref[s:e]).max()) / scale)
    return worst


def test_gae1d_reference_test_grid_golden():
    """Cases of realhf/tests/cpp_extensions/test_cugae.py (+ mixed bootstrap), outputs of the unmodified
    pygae1d_nolp_misalign.  Bars: (1) north_star: <= 1e-5 relative (to max|ref| of the sequence);
    (2) the reference test's atol=1e-5, up to the reference's OWN rounding: the kernel returns the
    correctly rounded recurrence (fp64 scan), the sequential fp32 reference drifts from it by up to
    ~2e-5 at max_seqlen=512, gamma=lam=1 -- so |ours - ref| <= |ref - exact| + 1 ulp, and ours is
    within 1e-5 absolute of the exact value everywhere."""
    z = H.load("gae1d.npz")
    n_within_ref_atol = 0
    for k, c in H.cases(z):
        g, l = (float(x) for x in c["gl"])
        adv, ret = run1d(c["rewards"], c["values"], c["cu"], c["boot"], g, l)
        assert per_seq_rel(adv, c["adv"], c["cu"]) <= H.TOL_GAE, k
        assert per_seq_rel(ret, c["ret"], c["cu"]) <= H.TOL_GAE, k
        a64, _ = O.gae1d_nolp_misalign_f64(c["rewards"], c["values"], c["cu"], c["boot"], g, l)
        ulp = np.spacing(np.abs(c["adv"]).astype(np.float32)).astype(np.float64)
        assert np.all(np.abs(adv - a64) <= 0.5 * ulp + 1e-12), k                 # correctly rounded
        assert np.allclose(adv, a64, atol=1e-5), k
        assert np.all(np.abs(adv.astype(np.float64) - c["adv"]) <= np.abs(c["adv"] - a64) + ulp), k
        n_within_ref_atol += bool(np.allclose(adv, c["adv"], atol=1e-5) and np.allclose(ret, c["ret"], atol=1e-5))
    assert n_within_ref_atol >= 40, n_within_ref_atol


@pytest.mark.parametrize("gamma,lam", [(1.0, 0.95), (1.0, 1.0), (0.99, 0.95)])
def test_gae1d_config0_vs_oracle(gamma, lam):
    """BASELINE.json configs[0]: 64 sequences, lengths 128-4096; oracle = C restatement (bit-equal to
    the reference on the golden set).  Both are also measured against the fp64 recurrence."""
    rng = np.random.RandomState(0)
    lens = rng.randint(128, 4097, 64)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    T = int(cu[-1])
    r = rng.randn(T).astype(np.float32)
    v = rng.randn(T + 64).astype(np.float32)
    boot = rng.rand(64) > 0.5
    adv, ret = run1d(r, v, cu, boot, gamma, lam)
    a_ref, r_ref = C.gae1d_nolp_misalign(r, v, cu, boot, gamma, lam)
    a64, _ = O.gae1d_nolp_misalign_f64(r, v, cu, boot, gamma, lam)
    assert per_seq_rel(adv, a_ref, cu) <= H.TOL_GAE
    assert per_seq_rel(ret, r_ref, cu) <= H.TOL_GAE
    ours64 = per_seq_rel(adv, a64, cu)
    ref64 = per_seq_rel(a_ref, a64, cu)
    assert ours64 <= 1e-7 + 2 ** -23, ours64      # correctly rounded recurrence
    assert ours64 <= ref64 + 1e-12


def test_gae1d_edge_cases():
    rng = np.random.RandomState(5)
    # empty sequences, length-1 sequences, a sequence straddling many tiles, T not a multiple of 8
    lens = np.array([0, 1, 1, 0, 5000, 1, 0, 0, 2047, 2049, 3, 0, 7001, 1], dtype=np.int64)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    B, T = len(lens), int(cu[-1])
    r = rng.randn(T).astype(np.float32)
    v = rng.randn(T + B).astype(np.float32)
    boot = rng.rand(B) > 0.5
    for gamma, lam in ((0.9, 0.5), (1.0, 1.0)):
        adv, ret = run1d(r, v, cu, boot, gamma, lam)
        a_ref, r_ref = C.gae1d_nolp_misalign(r, v, cu, boot, gamma, lam)
        assert per_seq_rel(adv, a_ref, cu) <= H.TOL_GAE and per_seq_rel(ret, r_ref, cu) <= H.TOL_GAE
    # tiny inputs
    adv, ret = run1d(np.array([2.0], np.float32), np.array([0.5, 3.0], np.float32), [0, 1], [True], 0.9, 0.5)
    assert np.allclose(adv, [2.0 + 0.9 * 3.0 - 0.5]) and np.allclose(ret, adv + 0.5)
    adv, ret = run1d(np.array([2.0], np.float32), np.array([0.5, 3.0], np.float32), [0, 1], [False], 0.9, 0.5)
    assert np.allclose(adv, [1.5])
    adv, ret = run1d(np.zeros(0, np.float32), np.zeros(2, np.float32), [0, 0, 0], [True, False], 0.9, 0.5)
    assert adv.shape == (0,)


def test_gae1d_long_chain_and_workspace_reuse():
    """One 300k-token sequence (look-back across ~150 tiles, no early exit) and many short ones in the
    same launch; repeated calls must agree bit for bit (the workspace cleans itself)."""
    rng = np.random.RandomState(9)
    lens = np.concatenate([[300000], rng.randint(1, 300, 2000)])
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    B, T = len(lens), int(cu[-1])
    r = rng.randn(T).astype(np.float32)
    v = rng.randn(T + B).astype(np.float32)
    boot = rng.rand(B) > 0.5
    a_ref, r_ref = C.gae1d_nolp_misalign(r, v, cu, boot, 0.999, 0.99)
    a64, _ = O.gae1d_nolp_misalign_f64(r, v, cu, boot, 0.999, 0.99)
    outs = [run1d(r, v, cu, boot, 0.999, 0.99) for _ in range(3)]
    for adv, ret in outs:
        assert per_seq_rel(adv, a_ref, cu) <= H.TOL_GAE and per_seq_rel(ret, r_ref, cu) <= H.TOL_GAE
        assert per_seq_rel(adv, a64, cu) <= 2 ** -23
        assert np.array_equal(adv, outs[0][0]) and np.array_equal(ret, outs[0][1])


@pytest.mark.parametrize("odd", [0, 3])
def test_gae1d_large_batch(odd):
    """~3.7M tokens (1800 tiles, more than one wave of resident CTAs): a 1M-token sequence, runs of
    length-1 sequences (2048 sequence ends in one tile), empty sequences, and T not a multiple of 4
    (scalar load/store path).  The static-scratch entry must give the same bits."""
    rng = np.random.RandomState(31 + odd)
    lens = np.concatenate([rng.randint(1, 3000, 600), [1 << 20], np.ones(5000, np.int64), [0, 0, 7],
                           rng.randint(1, 40, 20000), rng.randint(1, 9000, 300)])
    lens = np.concatenate([lens, [4 + (odd - int(lens.sum())) % 4]])          # T % 4 == odd
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    B, T = len(lens), int(cu[-1])
    assert 2_600_000 < T <= 4194304 and T % 4 == odd
    r = rng.randn(T).astype(np.float32)
    v = rng.randn(T + B).astype(np.float32)
    boot = rng.rand(B) > 0.5
    a_ref, r_ref = C.gae1d_nolp_misalign(r, v, cu, boot, 0.9995, 0.98)
    a64, _ = O.gae1d_nolp_misalign_f64(r, v, cu, boot, 0.9995, 0.98)
    outs = [run1d(r, v, cu, boot, 0.9995, 0.98) for _ in range(2)]
    for adv, ret in outs:
        assert per_seq_rel(adv, a_ref, cu) <= H.TOL_GAE and per_seq_rel(ret, r_ref, cu) <= H.TOL_GAE
        assert per_seq_rel(adv, a64, cu) <= 2 ** -23
        assert np.array_equal(adv, outs[0][0]) and np.array_equal(ret, outs[0][1])
    if T <= 4194304:
        # the static-scratch entry runs the one-tile-per-CTA kernel on the same batch: same bits
        from areal_b200 import _lib
        from areal_b200._lib import call, ptr, stream_ptr
        d = dev()
        tr, tv = torch.from_numpy(r).to(d), torch.from_numpy(v).to(d)
        tcu, tb = torch.from_numpy(cu).to(d), torch.from_numpy(boot.astype(np.uint8)).to(d)
        adv2, ret2 = torch.empty(T, device=d), torch.empty(T, device=d)
        call("areal_b200_gae1d_nolp_misalign_static", ptr(tr), ptr(tv), ptr(tcu), ptr(tb), T, B, 0.9995, 0.98,
             ptr(adv2), ptr(ret2), stream_ptr(d))
        assert np.array_equal(adv2.cpu().numpy(), outs[0][0]) and np.array_equal(ret2.cpu().numpy(), outs[0][1])


def test_gae1d_alternating_sizes_share_a_workspace():
    """Small and large batches alternate on the same cached workspace (2 to 300 tiles, ragged last
    tiles, one long sequence spanning every tile); each is checked against the oracle, twice in a row."""
    rng = np.random.RandomState(77)
    for ntok in (2 * 2048 + 1, 300 * 2048 - 5, 64 * 2048, 2049, 65 * 2048, 64 * 2048 - 7, 5000, 63 * 2048 + 9):
        if ntok % 3 == 0:
            lens = np.array([ntok])                                  # one sequence across all tiles
        else:
            lens = []
            left = ntok
            while left > 0:
                n = min(left, int(rng.randint(1, 5000)))
                lens.append(n)
                left -= n
            lens = np.array(lens)
        cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        B, T = len(lens), int(cu[-1])
        r = rng.randn(T).astype(np.float32)
        v = rng.randn(T + B).astype(np.float32)
        boot = rng.rand(B) > 0.5
        a_ref, r_ref = C.gae1d_nolp_misalign(r, v, cu, boot, 0.9999, 0.97)
        a64, _ = O.gae1d_nolp_misalign_f64(r, v, cu, boot, 0.9999, 0.97)
        first = None
        for _ in range(2):
            adv, ret = run1d(r, v, cu, boot, 0.9999, 0.97)
            assert per_seq_rel(adv, a64, cu) <= 2 ** -23, ntok
            assert per_seq_rel(adv, a_ref, cu) <= H.TOL_GAE * 5 and per_seq_rel(ret, r_ref, cu) <= H.TOL_GAE * 5, ntok
            if first is None:
                first = (adv, ret)
            else:
                assert np.array_equal(adv, first[0]) and np.array_equal(ret, first[1]), ntok


def test_gae1d_inputs_untouched_and_checks():
    from areal_b200 import ppo_functional as P
    d = dev()
    r = torch.randn(10, device=d)
    v = torch.randn(12, device=d)
    v0 = v.clone()
    cu = torch.tensor([0, 4, 10], dtype=torch.int32, device=d)
    boot = torch.tensor([False, True], device=d)
    P.cugae1d_nolp_misalign_func(r, v, cu, boot, 1.0, 1.0)
    assert torch.equal(v, v0)          # pygae1d zeroes values[v_end] for bootstrap=False; CUDA must not
    with pytest.raises(RuntimeError):
        P.cugae1d_nolp_misalign_func(r, v, cu.long(), boot, 1.0, 1.0)
    with pytest.raises(RuntimeError):
        P.cugae1d_nolp_misalign_func(r, v[:-1], cu, boot, 1.0, 1.0)
    with pytest.raises(RuntimeError):
        P.cugae1d_nolp_misalign_func(r, v, cu, boot.float(), 1.0, 1.0)
    adv, ret = P.get_packed_advantages_and_returns(1.0, 0.95, v, r, cu.long(), boot.float())
    a_ref, r_ref = C.gae1d_nolp_misalign(r.cpu().numpy(), v.cpu().numpy(), cu.cpu().numpy(), boot.cpu().numpy(), 1.0, 0.95)
    assert np.allclose(adv.cpu().numpy(), a_ref, atol=1e-5)


def run_padded(r, v, m, ne, gamma, lam):
    from areal_b200 import ppo_functional as P
    d = dev()
    adv, ret = P.gae_padded_masked(torch.from_numpy(r).to(d), torch.from_numpy(v).to(d), torch.from_numpy(m).to(d),
                                   torch.from_numpy(np.asarray(ne).astype(bool)).to(d), gamma, lam)
    return adv.cpu().numpy(), ret.cpu().numpy()


def row_rel(got, ref):
    scale = np.maximum(1.0, np.abs(ref).max(axis=1, keepdims=True))
    return float((np.abs(got.astype(np.float64) - ref) / scale).max())


def test_gae_padded_golden():
    z = H.load("gae_padded.npz")
    for k, c in H.cases(z):
        g, l = (float(x) for x in c["gl"])
        adv, ret = run_padded(c["rewards"], c["values"], c["mask"], c["no_eos"], g, l)
        assert row_rel(adv, c["adv"]) <= H.TOL_GAE and row_rel(ret, c["ret"]) <= H.TOL_GAE, k
        assert np.array_equal(adv[:, -1], np.zeros(adv.shape[0], np.float32))


@pytest.mark.parametrize("B,S", [(1, 1), (2, 2), (37, 5000), (300, 2049),
                                 # few long rows: one cluster of up to 8 CTAs per row (4096 tokens each); unaligned S;
                                 # the longest row a cluster takes; beyond it (one CTA per row again)
                                 (16, 32768), (5, 20001), (3, 32769), (2, 40000), (9, 4098)])
def test_gae_padded_vs_oracle(B, S):
    rng = np.random.RandomState(B * 7 + S)
    r = rng.randn(B, S).astype(np.float32)
    v = rng.randn(B, S).astype(np.float32)
    lens = rng.randint(1, S + 1, B)
    prompt = (lens * rng.rand(B) * 0.5).astype(np.int64)
    pos = np.arange(S)[None, :]
    m = ((pos >= prompt[:, None]) & (pos < lens[:, None]) & (rng.rand(B, S) > 0.05)).astype(np.float32)
    m = np.roll(m, -1, axis=1)
    if B > 2:
        m[0] = 0.0        # nothing valid
        m[1] = 1.0        # everything valid
    ne = lens == S
    for gamma, lam in ((1.0, 1.0), (0.99, 0.95)):
        adv, ret = run_padded(r, v, m, ne, gamma, lam)
        a_ref, r_ref = C.gae_padded_masked(r, v, m, ne, gamma, lam)
        assert row_rel(adv, a_ref) <= H.TOL_GAE and row_rel(ret, r_ref) <= H.TOL_GAE


def test_gae2d_olp_vs_oracle():
    """No reference caller / test exists; the golden of the kernel itself is checked below.  Here: the restatement of
    gae_kernel_2d_olp (gae.cu:62-89), incl. truncated segments and the nolp alias."""
    from areal_b200 import ppo_functional as P
    d = dev()
    rng = np.random.RandomState(4)
    for B, Hh in ((3, 40), (5, 4100), (16, 513)):
        r = rng.randn(B, Hh).astype(np.float32)
        v = rng.randn(B, Hh + 1).astype(np.float32)
        done = rng.rand(B, Hh + 1) < 0.01
        done[:, 0] = False
        idx, cu, trunc = [], [0], []
        for b in range(B):
            ys = np.nonzero(done[b])[0]
            ys = ys[ys <= Hh]
            idx.extend(ys.tolist())
            cu.append(len(idx))
            trunc.extend((rng.rand(len(ys)) < 0.3).tolist())
        idx = np.array(idx, np.int32)
        cu = np.array(cu, np.int32)
        trunc = np.array(trunc, bool)
        a_ref, r_ref = O.gae_2d_olp(r, v, done, idx, cu, trunc, 0.99, 0.95)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(d)
        for fn in (P.gae_2d_olp, P.gae_2d_nolp):
            adv, ret = fn(t(r), t(v), t(done), t(idx), t(cu), int(np.diff(cu).max(initial=0)), t(trunc), 0.99, 0.95)
            assert row_rel(adv.cpu().numpy(), a_ref) <= H.TOL_GAE
            assert row_rel(ret.cpu().numpy(), r_ref) <= H.TOL_GAE


def test_gae2d_vs_reference_kernel_golden():
    """Our segment-table 2D GAE against outputs of the reference's own CUDA kernels (tests/golden/gae2d.npz, made on a
    B200 by oracle/gen_golden_gpu.py from csrc/cugae/gae.cu compiled where it lies): 1e-5 relative, both entry points."""
    from areal_b200 import ppo_functional as P
    d = dev()
    z = H.load("gae2d.npz")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(d)
    for k, c in H.cases(z):
        mx = int(np.diff(c["cu"]).max(initial=0))
        for fn, tag in ((P.gae_2d_olp, "olp"), (P.gae_2d_nolp, "nolp")):
            adv, ret = fn(t(c["rewards"]), t(c["values"]), t(c["done"]), t(c["idx"]), t(c["cu"]), mx, t(c["trunc"]),
                          float(c["gamma"]), float(c["lam"]))
            assert row_rel(adv.cpu().numpy(), c["adv_" + tag]) <= H.TOL_GAE, (k, tag)
            assert row_rel(ret.cpu().numpy(), c["ret_" + tag]) <= H.TOL_GAE, (k, tag)
