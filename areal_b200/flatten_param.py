"""Host-side mirror of the interval ops of `realhf/impl/model/nn/flatten_param.py:65-186` and of the
native modules `realhf._C.interval_op_cuda` (csrc/interval_op/interval_op.cu:106-113) and
`realhf._C.interval_op` (csrc/interval_op/interval_op.cpp:27-29).  CUDA tensors only."""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import call, ptr, stream_ptr


def _check(cond, msg):
    if not cond:
        raise RuntimeError(msg)


def _intervals_cuda(intervals_cpu, intervals_cuda, device):
    if intervals_cuda is None:
        arr = np.ascontiguousarray(np.asarray(intervals_cpu, dtype=np.int64).reshape(-1, 2))
        intervals_cuda = torch.from_numpy(arr).to(device, non_blocking=True)
    _check(intervals_cuda.is_cuda, "intervals must be on CUDA")
    _check(intervals_cuda.is_contiguous(), "intervals must be contiguous")
    _check(intervals_cuda.dtype == torch.long, "intervals must be of type long")
    _check(intervals_cuda.dim() == 2 and intervals_cuda.size(1) == 2, "intervals must have shape (N, 2)")
    return intervals_cuda


def _run(slice_: bool, flat: torch.Tensor, packed: torch.Tensor, iv: torch.Tensor):
    dev = flat.device
    n = iv.size(0)
    with torch.cuda.device(dev):
        nbytes = _lib.lib().areal_b200_interval_workspace_bytes(n)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        if slice_:
            call("areal_b200_slice_intervals", ptr(flat), flat.numel(), ptr(packed), packed.numel(), ptr(iv), n,
                 flat.element_size(), ptr(ws), nbytes, stream_ptr(dev))
        else:
            call("areal_b200_set_intervals", ptr(packed), packed.numel(), ptr(flat), flat.numel(), ptr(iv), n,
                 flat.element_size(), ptr(ws), nbytes, stream_ptr(dev))
    return ws


def slice_intervals_any(src: torch.Tensor, intervals_cuda: torch.Tensor, total_size: int,
                        max_interval_size: int = 0, return_status: bool = False):
    """`slice_intervals_{fp32,fp16,bf16}(src, intervals, total_size, max_interval_size)`
    (interval_op.cu:68-104): returns a new packed tensor.  `max_interval_size` only sized the
    reference's grid; ignored."""
    _check(src.is_cuda, "src must be on CUDA")
    _check(src.is_contiguous(), "src must be contiguous")
    _check(src.dim() == 1, "src must be 1-D")
    iv = _intervals_cuda(None, intervals_cuda, src.device)
    dst = torch.empty(int(total_size), dtype=src.dtype, device=src.device)
    ws = _run(True, src, dst, iv)
    return (dst, ws[:8].view(torch.int64)) if return_status else dst


def set_intervals_any(src: torch.Tensor, dst: torch.Tensor, intervals_cuda: torch.Tensor,
                      max_interval_size: int = 0, return_status: bool = False):
    """`set_intervals_{fp32,fp16,bf16}(src, dst, intervals, max_interval_size)` (interval_op.cu:28-66):
    writes `dst` in place, returns None."""
    _check(src.is_cuda and dst.is_cuda, "src and dst must be on CUDA")
    _check(src.is_contiguous() and dst.is_contiguous(), "src and dst must be contiguous")
    _check(src.dtype == dst.dtype, "Source and destination tensors must have the same dtype")
    iv = _intervals_cuda(None, intervals_cuda, src.device)
    ws = _run(False, dst, src, iv)
    return ws[:8].view(torch.int64) if return_status else None


def _typed(fn, dtype):
    def f(*args, **kw):
        t = args[0]
        _check(t.dtype == dtype, f"expected {dtype}, got {t.dtype}")
        return fn(*args, **kw)
    return f


slice_intervals_fp32 = _typed(slice_intervals_any, torch.float32)
slice_intervals_fp16 = _typed(slice_intervals_any, torch.float16)
slice_intervals_bf16 = _typed(slice_intervals_any, torch.bfloat16)
set_intervals_fp32 = _typed(set_intervals_any, torch.float32)
set_intervals_fp16 = _typed(set_intervals_any, torch.float16)
set_intervals_bf16 = _typed(set_intervals_any, torch.bfloat16)


def slice_intervals(src: torch.Tensor, intervals_cpu: Optional[Sequence[Tuple[int, int]]] = None,
                    intervals_cuda: Optional[torch.Tensor] = None, output_size: Optional[int] = None,
                    max_interval_size: Optional[int] = None) -> torch.Tensor:
    """Drop-in for `slice_intervals` (flatten_param.py:97-143).  Unlike the reference there is no
    N < 1024 Python branch: the kernel is load-balanced over bytes, not over intervals, so it is
    used for every N."""
    iv = _intervals_cuda(intervals_cpu, intervals_cuda, src.device)
    if output_size is None:
        _check(intervals_cpu is not None, "output_size is required when only intervals_cuda is given")
        arr = np.asarray(intervals_cpu, dtype=np.int64).reshape(-1, 2)
        output_size = int((arr[:, 1] - arr[:, 0]).sum())
    return slice_intervals_any(src, iv, output_size, max_interval_size or 0)


def set_intervals(src: torch.Tensor, dst: torch.Tensor,
                  intervals_cpu: Optional[Sequence[Tuple[int, int]]] = None,
                  intervals_cuda: Optional[torch.Tensor] = None, max_interval_size: Optional[int] = None):
    """Drop-in for `set_intervals` (flatten_param.py:146-186)."""
    iv = _intervals_cuda(intervals_cpu, intervals_cuda, dst.device)
    return set_intervals_any(src, dst, iv, max_interval_size or 0)


def merge_intervals(intervals: List[Tuple[int, int]]) -> List[Tuple[int, int]]:
    """`realhf._C.interval_op.merge_intervals` (interval_op.cpp:4-25): merge touching neighbours in
    input order (host-side, plan time)."""
    arr = np.ascontiguousarray(np.asarray(intervals, dtype=np.uint64).reshape(-1, 2))
    out = np.empty_like(arr)
    k = _lib.lib().areal_b200_merge_intervals(arr.ctypes.data_as(ctypes.c_void_p), arr.shape[0],
                                              out.ctypes.data_as(ctypes.c_void_p))
    if k < 0:
        raise _lib.ArealB200Error("merge_intervals: bad arguments")
    return [(int(a), int(b)) for a, b in out[:k]]
