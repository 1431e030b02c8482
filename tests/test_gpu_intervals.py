"""GPU parity: slice_intervals / set_intervals through the C ABI, bit-exact vs golden vectors of the
reference's `_slice_intervals_py` / `_set_intervals_py` and vs the C oracle at the reference test's
sizes (realhf/tests/cpp_extensions/test_interval_ops.py: 1e8 elements, up to 100k intervals)."""
import numpy as np
import pytest
import torch

from oracle import c_oracle as C
from oracle import cpu_oracle as O
from tests import helpers as H

pytestmark = pytest.mark.gpu
D = "cuda:0"
NP_OF = {"bf16": np.uint16, "fp16": np.uint16, "fp32": np.uint32}
TORCH_OF = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}


def test_golden_bit_exact():
    from areal_b200 import flatten_param as fp
    z = H.load("intervals.npz")
    for k, c in H.cases(z):
        dt = str(c["dtype"])
        iv = c["intervals"]
        src = H.bits_to_torch(c["src"], dt).to(D)
        iv_cuda = torch.from_numpy(iv).to(D)
        out = fp.slice_intervals(src, [tuple(x) for x in iv.tolist()], intervals_cuda=iv_cuda,
                                 output_size=int(c["packed"].shape[0]), max_interval_size=int((iv[:, 1] - iv[:, 0]).max()))
        assert out.dtype == src.dtype
        assert np.array_equal(H.torch_bits(out), c["packed"]), k
        out2 = fp.slice_intervals(src, [tuple(x) for x in iv.tolist()])       # cpu list only
        assert np.array_equal(H.torch_bits(out2), c["packed"]), k
        dst = H.bits_to_torch(c["dst"], dt).to(D)
        assert fp.set_intervals(H.bits_to_torch(c["newvals"], dt).to(D), dst, None, intervals_cuda=iv_cuda) is None
        assert np.array_equal(H.torch_bits(dst), c["dst_after"]), k


@pytest.mark.parametrize("n_intervals", [100000, 10000, 4000, 2000, 1000, 500, 100, 1])
@pytest.mark.parametrize("dt", ["bf16", "fp32", "fp16"])
def test_reference_test_pattern_1e8(n_intervals, dt):
    """Same generator and sizes as the reference's (skipped) test_get / test_set."""
    if dt != "bf16" and n_intervals not in (100000, 1000, 1):
        pytest.skip("bf16 covers the full grid; other dtypes the extremes")
    from areal_b200 import flatten_param as fp
    rng = np.random.RandomState(0)
    n = int(1e8)
    iv, total, max_sz = O.make_intervals(n, n_intervals, rng)
    npdt = NP_OF[dt]
    src_bits = rng.randint(0, np.iinfo(npdt).max, n, dtype=npdt)
    want = C.slice_intervals(src_bits, iv)
    src = torch.from_numpy(src_bits.view(np.int16 if npdt == np.uint16 else np.int32)).to(D).view(TORCH_OF[dt])
    iv_cuda = torch.from_numpy(iv).to(D)
    fn = getattr(fp, f"slice_intervals_{dt}")
    out = fn(src, iv_cuda, total, max_sz)
    assert np.array_equal(H.torch_bits(out), want)
    # set: scatter new values, everything outside the intervals untouched
    new_bits = rng.randint(0, np.iinfo(npdt).max, total, dtype=npdt)
    dst_bits = src_bits.copy()
    C.set_intervals(new_bits, dst_bits, iv)
    newv = torch.from_numpy(new_bits.view(np.int16 if npdt == np.uint16 else np.int32)).to(D).view(TORCH_OF[dt])
    getattr(fp, f"set_intervals_{dt}")(newv, src, iv_cuda, max_sz)
    assert np.array_equal(H.torch_bits(src), dst_bits)
    # round trip
    back = fn(src, iv_cuda, total, max_sz)
    assert np.array_equal(H.torch_bits(back), new_bits)


def test_ragged_misaligned_and_row_split_intervals():
    """Odd starts / sizes (every relative alignment 2..16 B), empty intervals, and the realistic shape:
    a [3584, 18944] weight split on dim 1 for TP=2 -> one 9472-element interval per row
    (realhf/impl/model/nn/real_llm_parallel.py:101-107)."""
    from areal_b200 import flatten_param as fp
    rng = np.random.RandomState(3)
    n = 3584 * 18944 + 12345
    src_bits = rng.randint(0, 65535, n, dtype=np.uint16)
    rows = np.arange(3584, dtype=np.int64)
    row_split = np.stack([rows * 18944 + 9472, rows * 18944 + 18944], 1) + 12345
    starts = np.sort(rng.choice(12000, 500, replace=False)).astype(np.int64)
    ends = np.minimum(starts + rng.randint(0, 40, 500), np.append(starts[1:], 12345))
    ragged = np.stack([starts, ends], 1)
    iv = np.concatenate([ragged, row_split])
    rng.shuffle(iv)
    want = C.slice_intervals(src_bits, iv)
    src = torch.from_numpy(src_bits.view(np.int16)).to(D).view(torch.bfloat16)
    out, status = fp.slice_intervals_any(src, torch.from_numpy(iv).to(D), want.shape[0], return_status=True)
    assert int(status[0]) == 0
    assert np.array_equal(H.torch_bits(out), want)
    for es, tdt, npdt in ((1, torch.uint8, np.uint8), (8, torch.float64, np.uint64)):
        sb = rng.randint(0, 255, 50000 * es, dtype=np.uint8).view(npdt)
        ivs = np.stack([np.arange(0, 49000, 70), np.arange(0, 49000, 70) + rng.randint(0, 70, 700)], 1).astype(np.int64)
        w = C.slice_intervals(sb, ivs)
        o = fp.slice_intervals_any(torch.from_numpy(sb.view(np.uint8)).to(D).view(tdt), torch.from_numpy(ivs).to(D), w.shape[0])
        assert np.array_equal(o.cpu().numpy().view(npdt) if es == 1 else o.view(torch.int64).cpu().numpy().view(npdt), w)


def test_invalid_tables_write_nothing():
    from areal_b200 import flatten_param as fp
    src = torch.arange(1000, dtype=torch.float32, device=D)
    dst = torch.full((1000,), -1.0, device=D)
    new = torch.ones(20, device=D)
    for bad, code in (([[0, 10], [995, 1005]], 1), ([[10, 0], [0, 30]], 1), ([[-5, 5], [20, 30]], 1),
                      ([[0, 10], [20, 25]], 2)):
        st = fp.set_intervals_any(new, dst, torch.tensor(bad, dtype=torch.long, device=D), return_status=True)
        assert int(st[0]) == code
        assert bool((dst == -1.0).all())
    out, st = fp.slice_intervals_any(src, torch.zeros((0, 2), dtype=torch.long, device=D), 0, return_status=True)
    assert out.numel() == 0 and int(st[0]) == 0
    with pytest.raises(RuntimeError):
        fp.set_intervals_any(new.half(), dst, torch.tensor([[0, 20]], dtype=torch.long, device=D))
    with pytest.raises(RuntimeError):
        fp.slice_intervals_any(src, torch.tensor([[0, 20]], dtype=torch.int32, device=D), 20)


def test_every_byte_shift_long_pieces():
    """1-byte elements, pieces of 256..6000 bytes whose source and destination offsets take every
    relative shift 0..15 (the funnel-shift path and its head / tail handling), gather and scatter; pieces
    end flush against the end of the buffers (the aligned over-read stays inside the last 16-byte block)."""
    from areal_b200 import flatten_param as fp
    rng = np.random.RandomState(17)
    n = 1 << 22
    src = rng.randint(0, 256, n, dtype=np.uint8)
    starts, ends, pos = [], [], 3
    while True:
        size = int(rng.randint(256, 6000))
        if pos + size > n:
            break
        starts.append(pos)
        ends.append(pos + size)
        pos += size + int(rng.randint(0, 37))
    starts.append(n - 1000 - 5)             # last piece ends exactly at the end of the flat buffer
    ends.append(n)
    iv = np.stack([np.array(starts), np.array(ends)], 1).astype(np.int64)
    rng.shuffle(iv)
    shifts = set()
    off = 0
    for s, e in iv:
        shifts.add((off - s) & 15)
        off += e - s
    assert shifts == set(range(16))
    want = C.slice_intervals(src, iv)
    tsrc = torch.from_numpy(src).to(D)
    tiv = torch.from_numpy(iv).to(D)
    out = fp.slice_intervals_any(tsrc, tiv, want.shape[0])
    assert np.array_equal(out.cpu().numpy(), want)
    new = rng.randint(0, 256, want.shape[0], dtype=np.uint8)
    dst = torch.from_numpy(src.copy()).to(D)
    fp.set_intervals_any(torch.from_numpy(new).to(D), dst, tiv)
    ref = src.copy()
    C.set_intervals(new, ref, iv)
    assert np.array_equal(dst.cpu().numpy(), ref)


def test_many_intervals_span_several_look_back_rounds():
    """300,000 intervals = 74 scan tiles: the offset scan's warp-wide look-back needs three rounds for the last
    tiles.  Ragged sizes incl. empty intervals and odd offsets, shuffled order; bit-exact vs numpy, round trip
    through set_intervals, and an invalid interval in a late tile still stops every write."""
    from areal_b200 import flatten_param as fp
    rng = np.random.RandomState(11)
    n = 300_000
    sizes = rng.randint(0, 90, n).astype(np.int64)
    sizes[rng.rand(n) < 0.05] = 0
    gaps = rng.randint(0, 7, n).astype(np.int64)
    starts = np.cumsum(sizes + gaps) - sizes
    numel = int(starts[-1] + sizes[-1] + 5)
    iv = np.stack([starts, starts + sizes], 1)
    rng.shuffle(iv)
    total = int(sizes.sum())
    src_np = rng.randint(0, 65536, numel).astype(np.uint16)
    src = torch.from_numpy(src_np.view(np.int16)).view(torch.bfloat16).to(D)
    ivc = torch.from_numpy(iv).to(D)
    out, st = fp.slice_intervals_any(src, ivc, total, return_status=True)
    assert int(st[0]) == 0
    want = np.concatenate([src_np[a:b] for a, b in iv])
    assert np.array_equal(H.torch_bits(out), want)
    dst = torch.zeros_like(src)
    fp.set_intervals_any(out, dst, ivc)
    covered = np.zeros(numel, bool)
    for a, b in iv:
        covered[a:b] = True
    assert np.array_equal(H.torch_bits(dst), np.where(covered, src_np, 0))
    bad = iv.copy()
    bad[n - 1000] = (numel - 3, numel + 40)          # beyond the flat buffer, in tile 73
    sentinel = torch.full((total,), 7.0, dtype=torch.bfloat16, device=D)
    st = fp.set_intervals_any(sentinel, dst, torch.from_numpy(bad).to(D), return_status=True)
    assert int(st[0]) == 1
    assert np.array_equal(H.torch_bits(dst), np.where(covered, src_np, 0))
