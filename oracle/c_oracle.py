"""TEST INFRASTRUCTURE ONLY -- ctypes front end of oracle/liboracle.so (see oracle/oracle.c)."""
import ctypes
import os

import numpy as np

from . import build_oracle

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "liboracle.so")
        if not os.path.exists(path):
            build_oracle.build_c_oracle()
        _lib = ctypes.CDLL(path)
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def gae1d_nolp_misalign(rewards, values, cu_seqlens, bootstrap, gamma, lam):
    r = np.ascontiguousarray(rewards, dtype=np.float32)
    v = np.ascontiguousarray(values, dtype=np.float32)
    cu = np.ascontiguousarray(cu_seqlens, dtype=np.int32)
    b = np.ascontiguousarray(np.asarray(bootstrap).astype(np.uint8))
    adv = np.zeros_like(r)
    ret = np.zeros_like(r)
    lib().oracle_gae1d_nolp_misalign(_p(r), _p(v), _p(cu), _p(b), ctypes.c_int64(cu.shape[0] - 1),
                                     ctypes.c_float(np.float32(gamma)), ctypes.c_float(np.float32(gamma * lam)),
                                     _p(adv), _p(ret))
    return adv, ret


def gae_padded_masked(rewards, values, mask, no_eos, gamma, lam):
    r = np.ascontiguousarray(rewards, dtype=np.float32)
    v = np.ascontiguousarray(values, dtype=np.float32)
    m = np.ascontiguousarray(mask, dtype=np.float32)
    ne = np.ascontiguousarray(np.asarray(no_eos).astype(np.uint8))
    adv = np.zeros_like(r)
    ret = np.zeros_like(r)
    lib().oracle_gae_padded_masked(_p(r), _p(v), _p(m), _p(ne), ctypes.c_int64(r.shape[0]),
                                   ctypes.c_int64(r.shape[1]), ctypes.c_float(np.float32(gamma)),
                                   ctypes.c_float(np.float32(gamma * lam)), _p(adv), _p(ret))
    return adv, ret


def slice_intervals(src: np.ndarray, intervals):
    iv = np.ascontiguousarray(intervals, dtype=np.int64).reshape(-1, 2)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    out = np.empty(total, dtype=src.dtype)
    lib().oracle_slice_intervals(_p(np.ascontiguousarray(src)), _p(out), _p(iv), ctypes.c_int64(iv.shape[0]),
                                 ctypes.c_int64(src.dtype.itemsize))
    return out


def set_intervals(packed: np.ndarray, dst: np.ndarray, intervals):
    iv = np.ascontiguousarray(intervals, dtype=np.int64).reshape(-1, 2)
    assert dst.flags["C_CONTIGUOUS"]
    lib().oracle_set_intervals(_p(np.ascontiguousarray(packed)), _p(dst), _p(iv), ctypes.c_int64(iv.shape[0]),
                               ctypes.c_int64(dst.dtype.itemsize))


def merge_intervals(intervals):
    iv = np.ascontiguousarray(intervals, dtype=np.uint64).reshape(-1, 2)
    out = np.empty_like(iv)
    f = lib().oracle_merge_intervals
    f.restype = ctypes.c_int64
    k = f(_p(iv), ctypes.c_int64(iv.shape[0]), _p(out))
    return [(int(a), int(b)) for a, b in out[:k]]
