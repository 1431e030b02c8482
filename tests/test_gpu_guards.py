"""GPU: out-of-bounds WRITE detection with guard bands (compute-sanitizer is closed on this pool).
Every output is a window of a larger sentinel-filled buffer (and dlogits rows have padding columns); the
sentinels must be intact after the kernels ran.  Called straight through the C ABI."""
import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
D = torch.device("cuda:0")
SENT = -7777.0


def guarded(n, dtype=torch.float32, guard=256):
    buf = torch.full((n + 2 * guard,), 0 if dtype == torch.uint8 else SENT, dtype=dtype, device=D)
    return buf, buf[guard:guard + n]


def intact(buf, n, guard=256):
    return bool((buf[:guard] == SENT).all()) and bool((buf[guard + n:] == SENT).all())


@pytest.mark.parametrize("V", [151936, 4099, 18992])
def test_logits_kernels_write_only_their_windows(V):
    from areal_b200 import _lib
    from areal_b200._lib import call, ptr, stream_ptr
    g = torch.Generator().manual_seed(V)
    T, pad = 37, 24
    logits = (torch.randn(T, V, generator=g) * 2).bfloat16().to(D)
    labels = torch.randint(0, V, (T,), generator=g).to(D)
    labels[0], labels[1] = 0, V - 1
    outs = {k: guarded(T) for k in ("logp", "ent", "rmax", "lns", "rmin", "coef", "sl", "sr", "sk", "bw", "bk")}
    masks = {k: guarded(T, torch.uint8) for k in ("cm", "dm", "bm")}
    for b, _ in masks.values():
        b.fill_(201)
    loss, inv = torch.zeros(1, device=D), torch.zeros(1, device=D)
    prox = torch.randn(T, device=D) - 9
    old = prox + 0.05 * torch.randn(T, device=D)
    adv = torch.randn(T, device=D)
    mask = (torch.rand(T, device=D) > 0.3)
    ws = torch.zeros(_lib.lib().areal_b200_loss_workspace_bytes(), dtype=torch.uint8, device=D)
    dl_buf = torch.full((T + 2, V + pad), SENT, dtype=torch.bfloat16, device=D)
    dl = dl_buf[1:T + 1]
    scale = torch.full((1,), 0.5, device=D)
    for single in (False, True):
        dl_buf.fill_(SENT)
        args = [ptr(logits), 0, T, V, V, ptr(labels), 1.0, ptr(prox), ptr(old), ptr(adv), ptr(mask.view(torch.uint8)),
                0.4, float("nan"), 3.0, 5.0, ptr(outs["logp"][1]), ptr(outs["ent"][1]), ptr(outs["rmax"][1]),
                ptr(outs["lns"][1]), ptr(outs["rmin"][1]), ptr(loss), ptr(inv), ptr(outs["coef"][1]), ptr(outs["sl"][1]),
                ptr(outs["sr"][1]), ptr(outs["sk"][1]), ptr(masks["cm"][1]), ptr(masks["dm"][1]), ptr(outs["bw"][1]),
                ptr(outs["bk"][1]), ptr(masks["bm"][1])]
        if single:
            if V % 8:
                continue
            call("areal_b200_fused_actor_fwd_bwd", *args, ptr(scale), ptr(dl), V + pad, 0, 0, ptr(ws), ws.numel(), stream_ptr(D))
        else:
            call("areal_b200_fused_actor_fwd", *args, ptr(ws), ws.numel(), stream_ptr(D))
            call("areal_b200_logits_bwd", ptr(logits), 0, T, V, V, ptr(labels), 0, 1.0, ptr(outs["rmax"][1]),
                 ptr(outs["lns"][1]), None, ptr(outs["coef"][1]), None, ptr(inv), ptr(scale), ptr(dl), V + pad, stream_ptr(D))
        torch.cuda.synchronize()
        for k, (buf, _) in outs.items():
            assert intact(buf, T), (k, single)
        for k, (buf, _) in masks.items():
            assert bool((buf[:256] == 201).all()) and bool((buf[256 + T:] == 201).all()), (k, single)
        assert bool((dl_buf[0] == SENT).all()) and bool((dl_buf[-1] == SENT).all()), single
        assert bool((dl_buf[1:T + 1, V:] == SENT).all()), single
        assert not bool((dl[:, :V] == SENT).any())


def test_gae_and_intervals_write_only_their_windows():
    from areal_b200 import _lib
    from areal_b200._lib import call, ptr, stream_ptr
    rng = np.random.RandomState(2)
    lens = rng.randint(1, 3000, 11)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    T, B = int(cu[-1]), len(lens)
    r = torch.randn(T, device=D)
    v = torch.randn(T + B, device=D)
    boot = torch.rand(B, device=D) > 0.5
    cu_t = torch.from_numpy(cu).to(D)
    abuf, a = guarded(T)
    rbuf, rr = guarded(T)
    wsbytes = _lib.lib().areal_b200_gae1d_workspace_bytes(T)
    wbuf, w = guarded(wsbytes, torch.uint8)
    wbuf.fill_(77)
    call("areal_b200_gae1d_nolp_misalign", ptr(r), ptr(v), ptr(cu_t), ptr(boot.view(torch.uint8)), T, B, 0.99, 0.95,
         ptr(a), ptr(rr), ptr(w), wsbytes, stream_ptr(D))
    call("areal_b200_gae1d_nolp_misalign_static", ptr(r), ptr(v), ptr(cu_t), ptr(boot.view(torch.uint8)), T, B, 0.99,
         0.95, ptr(a), ptr(rr), stream_ptr(D))
    torch.cuda.synchronize()
    assert intact(abuf, T) and intact(rbuf, T)
    assert bool((wbuf[:256] == 77).all()) and bool((wbuf[256 + wsbytes:] == 77).all())
    # padded GAE
    Bp, S = 5, 2100
    obuf, o = guarded(Bp * S)
    o2buf, o2 = guarded(Bp * S)
    x = torch.randn(Bp, S, device=D)
    call("areal_b200_gae_padded_masked", ptr(x), ptr(x), ptr((x > 0).float()), ptr(torch.ones(Bp, dtype=torch.uint8, device=D)),
         Bp, S, 1.0, 0.95, ptr(o), ptr(o2), stream_ptr(D))
    torch.cuda.synchronize()
    assert intact(obuf, Bp * S) and intact(o2buf, Bp * S)
    # interval ops: odd sizes / offsets
    n = 50000
    starts = np.sort(rng.choice(n - 200, 300, replace=False)).astype(np.int64)
    ends = np.minimum(starts + rng.randint(1, 150, 300), np.append(starts[1:], n))
    iv = np.stack([starts, ends], 1)
    rng.shuffle(iv)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    ivc = torch.from_numpy(iv).to(D)
    flat_buf = torch.full((n + 512,), SENT, dtype=torch.bfloat16, device=D)
    flat = flat_buf[256:256 + n]
    flat.copy_(torch.randn(n, device=D).bfloat16())
    pbuf = torch.full((total + 512,), SENT, dtype=torch.bfloat16, device=D)
    packed = pbuf[256:256 + total]
    wsb = _lib.lib().areal_b200_interval_workspace_bytes(300)
    ws = torch.empty(wsb, dtype=torch.uint8, device=D)
    call("areal_b200_slice_intervals", ptr(flat), n, ptr(packed), total, ptr(ivc), 300, 2, ptr(ws), wsb, stream_ptr(D))
    torch.cuda.synchronize()
    assert bool((pbuf[:256] == SENT).all()) and bool((pbuf[256 + total:] == SENT).all())
    before = flat.clone()
    call("areal_b200_set_intervals", ptr(packed), total, ptr(flat), n, ptr(ivc), 300, 2, ptr(ws), wsb, stream_ptr(D))
    torch.cuda.synchronize()
    assert bool((flat_buf[:256] == SENT).all()) and bool((flat_buf[256 + n:] == SENT).all())
    assert torch.equal(flat, before)        # set(slice(x)) is the identity
