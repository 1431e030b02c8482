"""ctypes binding of libareal_b200.so (the C ABI declared in include/areal_b200.h).

There is NO fallback: if the shared library is missing, was not built for this machine, or a call
returns a non-zero status, this raises.  Build with `python -m areal_b200.build`.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_double, c_int, c_int64, c_void_p
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libareal_b200.so")

P = c_void_p
I64 = c_int64
F64 = c_double

# name -> (restype, argtypes); must list every symbol include/areal_b200.h declares
SIGNATURES = {
    "areal_b200_last_error": (c_char_p, []),
    "areal_b200_version": (c_int, []),
    "areal_b200_gae1d_workspace_bytes": (I64, [I64]),
    "areal_b200_gae1d_nolp_misalign": (c_int, [P, P, P, P, I64, I64, F64, F64, P, P, P, I64, P]),
    "areal_b200_gae1d_nolp_misalign_static": (c_int, [P, P, P, P, I64, I64, F64, F64, P, P, P]),
    "areal_b200_gae_padded_masked": (c_int, [P, P, P, P, I64, I64, F64, F64, P, P, P]),
    "areal_b200_advantages_padded": (c_int, [P, c_int, I64, I64, F64, P, c_int, P, c_int, P, P, c_int, P, I64, I64, c_int, F64, F64, F64,
                                             F64, c_int, c_int, c_int, F64, F64, P, P, P, P, P, P, P]),
    "areal_b200_vocab_p2p_buffer_bytes": (I64, [I64]),
    "areal_b200_vocab_combine_p2p": (c_int, [P, c_int, c_int, I64, I64, I64, F64, I64, P, P, P, P, P, P]),
    "areal_b200_vocab_push_buffer_bytes": (I64, [I64, c_int]),
    "areal_b200_logits_partial_fwd_push": (c_int, [P, c_int, I64, I64, I64, P, I64, F64, P, c_int, c_int, I64, I64, I64, I64, P, P,
                                                   P, P, P, P, P]),
    "areal_b200_clamped_entropy_fwd": (c_int, [P, c_int, I64, I64, I64, P, F64, I64, P, P, P, P]),
    "areal_b200_clamped_entropy_bwd": (c_int, [P, c_int, I64, I64, I64, P, F64, P, P, P, P, I64, P]),
    "areal_b200_lm_head_workspace_bytes": (I64, [I64, I64]),
    "areal_b200_lm_head_logprobs": (c_int, [P, P, P, I64, I64, I64, I64, I64, F64, P, P, P, P, P, P, I64, P]),
    "areal_b200_rms_update": (c_int, [P, I64, P, c_int, c_int, F64, P]),
    "areal_b200_rms_apply": (c_int, [P, I64, P, c_int, c_int, F64, c_int, P, P]),
    "areal_b200_packed_rewards": (c_int, [P, P, P, P, P, I64, I64, F64, F64, c_int, c_int, P, P, P]),
    "areal_b200_segment_masked_mean": (c_int, [P, P, P, I64, P, P, P]),
    "areal_b200_masked_stats_workspace_bytes": (I64, [I64, I64]),
    "areal_b200_masked_stats": (c_int, [P, P, P, ctypes.c_uint32, I64, I64, c_int, P, I64, P]),
    "areal_b200_gae2d_olp": (c_int, [P, P, P, P, P, P, I64, I64, I64, F64, F64, P, P, P]),
    "areal_b200_logits_fwd": (c_int, [P, c_int, I64, I64, I64, P, F64, P, P, P, P, P, P]),
    "areal_b200_logits_bwd": (c_int, [P, c_int, I64, I64, I64, P, I64, F64, P, P, P, P, P, P, P, P, I64, P]),
    "areal_b200_logits_partial_fwd": (c_int, [P, c_int, I64, I64, I64, P, I64, F64, P, P, P]),
    "areal_b200_vocab_combine": (c_int, [P, I64, I64, F64, P, P, P, P, P]),
    "areal_b200_legacy_entropy": (c_int, [P, c_int, I64, I64, I64, P, P, F64, P, P]),
    "areal_b200_loss_workspace_bytes": (I64, []),
    "areal_b200_ppo_actor_loss": (c_int, [P, P, P, P, P, I64, F64, F64, F64, F64,
                                          P, P, P, P, P, P, P, P, P, P, P, P, I64, P]),
    "areal_b200_ppo_critic_loss": (c_int, [P, P, P, P, I64, F64, c_int, P, P, P, P, P, P, I64, P]),
    "areal_b200_scale_by_scalars": (c_int, [P, P, P, I64, P, P]),
    "areal_b200_fused_actor_fwd": (c_int, [P, c_int, I64, I64, I64, P, F64, P, P, P, P, F64, F64, F64, F64,
                                           P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, I64, P]),
    "areal_b200_fused_actor_fwd_bwd": (c_int, [P, c_int, I64, I64, I64, P, F64, P, P, P, P, F64, F64, F64, F64,
                                               P, P, P, P, P, P, P, P, P, P, P, P, P, P, P, P,
                                               P, P, I64, c_int, c_int, P, I64, P]),
    "areal_b200_rescale_grad_if_needed": (c_int, [P, I64, I64, I64, P, P, P]),
    "areal_b200_norm_num_partials": (I64, [I64]),
    "areal_b200_masked_moments": (c_int, [P, P, I64, I64, I64, P, P]),
    "areal_b200_norm_centered_sumsq": (c_int, [P, P, c_int, c_int, c_int, I64, I64, P, I64, I64, I64, P, P]),
    "areal_b200_norm_apply": (c_int, [P, P, I64, c_int, c_int, c_int, I64, I64, P, c_int, c_int, c_int, I64, I64, P,
                                      F64, P, P, P]),
    "areal_b200_masked_normalization_apply": (c_int, [P, P, I64, P, c_int, F64, P, P]),
    "areal_b200_interval_workspace_bytes": (I64, [I64]),
    "areal_b200_slice_intervals": (c_int, [P, I64, P, I64, P, I64, I64, P, I64, P]),
    "areal_b200_set_intervals": (c_int, [P, I64, P, I64, P, I64, I64, P, I64, P]),
    "areal_b200_merge_intervals": (I64, [P, I64, P]),
}

_lib: Optional[ctypes.CDLL] = None


class ArealB200Error(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ArealB200Error(
                f"{LIB_PATH} not found: the CUDA extension is not built (run `python -m areal_b200.build`). "
                "areal_b200 has no CPU or PyTorch fallback.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)   # AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def call(name: str, *args) -> None:
    """Call an int-status entry point; raise with the library's message on failure."""
    rc = getattr(lib(), name)(*args)
    if rc != 0:
        msg = lib().areal_b200_last_error().decode("utf-8", "replace")
        raise ArealB200Error(f"{name} failed (status {rc}): {msg}")


def ptr(t: Optional[torch.Tensor]):
    return None if t is None else c_void_p(t.data_ptr())


def stream_ptr(device: torch.device):
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


DTYPE_CODE = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}


def require_cuda(*tensors: torch.Tensor) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise ArealB200Error("areal_b200 ops run on CUDA tensors only (no CPU fallback)")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise ArealB200Error(f"tensors on different devices: {dev} vs {t.device}")
    return dev


# ---- zero-initialised, self-cleaning workspaces (one per device and stream) -------------------------
_WS = {}


def workspace(kind: str, device: torch.device, nbytes: int) -> torch.Tensor:
    """Persistent device scratch that the kernels leave zeroed after use.  Keyed by stream so that
    calls issued on different streams never share a ticket counter."""
    key = (kind, device.index, torch.cuda.current_stream(device).cuda_stream)
    buf = _WS.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.zeros(max(int(nbytes), 4096), dtype=torch.uint8, device=device)
        _WS[key] = buf
    return buf
