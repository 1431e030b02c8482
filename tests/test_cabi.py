"""CPU: the C-ABI shared library loads and exports every symbol include/areal_b200.h declares; the
host-only entry points work; argument validation rejects bad calls before touching the GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

from tests import helpers as H

HEADER = os.path.join(H.ROOT, "include", "areal_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"AREAL_B200_API\s+[\w\s\*]+?\b(areal_b200_\w+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from areal_b200 import build
    build.build()
    from areal_b200 import _lib
    return _lib.lib()


def test_header_symbols_exported(lib):
    from areal_b200 import _lib
    syms = declared_symbols()
    assert len(syms) >= 19
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/areal_b200.h but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes signature"
    assert sorted(_lib.SIGNATURES) == syms


def test_host_only_entry_points(lib):
    assert lib.areal_b200_version() >= 100
    assert lib.areal_b200_gae1d_workspace_bytes(100) == 64
    assert lib.areal_b200_gae1d_workspace_bytes(2049) == 64 + 5 * 32      # sized for the smallest (512-token) tile
    assert lib.areal_b200_interval_workspace_bytes(10) == (14 + 4 + 1) * 8    # status[2] | off[n+1] | any_shift | scan scratch[4] | one word per 4096 intervals
    assert lib.areal_b200_interval_workspace_bytes(200704) == (200704 + 4 + 4 + 49) * 8
    assert lib.areal_b200_loss_workspace_bytes() >= 16 + 8 * 8


def test_merge_intervals_host():
    from areal_b200 import flatten_param as fp
    from oracle import cpu_oracle as O
    rng = np.random.RandomState(1)
    pts = np.sort(rng.choice(10000, 400, replace=False))
    iv = [(int(a), int(b)) for a, b in zip(pts[:-1], pts[1:]) if rng.rand() > 0.2]
    for case in ([], [(0, 4)], [(0, 4), (4, 9), (9, 10)], [(5, 7), (0, 5), (7, 9), (9, 9), (9, 12)], iv):
        assert fp.merge_intervals(case) == O.merge_intervals(case)


def test_argument_validation_without_gpu(lib):
    one = ctypes.c_void_p(16)   # never dereferenced: validation fails first
    rc = lib.areal_b200_logits_fwd(one, 0, 4, 0, 0, one, 1.0, one, None, one, None, None, None)
    assert rc == 1 and b"bad shape" in lib.areal_b200_last_error()
    rc = lib.areal_b200_logits_fwd(one, 7, 4, 8, 8, one, 1.0, one, None, one, None, None, None)
    assert rc == 1 and b"dtype" in lib.areal_b200_last_error()
    rc = lib.areal_b200_logits_fwd(one, 0, 4, 8, 8, one, 0.0, one, None, one, None, None, None)
    assert rc == 1 and b"temperature" in lib.areal_b200_last_error()
    rc = lib.areal_b200_slice_intervals(one, 10, one, 4, one, 2, 3, one, 1 << 20, None)
    assert rc == 1 and b"elem_size" in lib.areal_b200_last_error()
    rc = lib.areal_b200_slice_intervals(one, 10, one, 4, one, 2, 2, one, 8, None)
    assert rc == 3 and b"workspace" in lib.areal_b200_last_error()
    rc = lib.areal_b200_gae1d_nolp_misalign(one, one, one, one, 5000, 3, 1.0, 1.0, one, one, None, 0, None)
    assert rc == 3
    rc = lib.areal_b200_ppo_actor_loss(one, one, one, one, one, 8, 0.2, float("nan"), 0.5, float("nan"),
                                       one, one, one, *([None] * 8), one, 1 << 20, None)
    assert rc == 1 and b"c_clip" in lib.areal_b200_last_error()


def test_product_does_not_import_oracle():
    """The product path must never route through oracle/ (checked textually)."""
    pkg = os.path.join(H.ROOT, "areal_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("oracle/", "").lower() or f == "build.py", (dirpath, f)


def test_cpu_tensors_fail_loudly():
    import torch
    from areal_b200 import _lib, functional as F
    with pytest.raises(_lib.ArealB200Error):
        F.gather_logprobs(torch.randn(2, 8), torch.zeros(2, dtype=torch.long))
