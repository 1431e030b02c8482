"""Shared helpers for the parity tests (golden loading, bf16 bit views, tolerances)."""
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

# north_star tolerances
TOL_LOGITS = 1e-3     # log-prob / entropy / loss / dlogits: relative, bf16 in / fp32 accumulate
TOL_GAE = 1e-5        # GAE: relative to max|ref| per sequence, fp32


def load(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def cases(z, prefix="c", count_key="n_cases"):
    n = int(z[count_key])
    for k in range(n):
        p = f"{prefix}{k}_"
        yield k, {key[len(p):]: z[key] for key in z.files if key.startswith(p)}


def bf16_from_bits(bits: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(bits.view(np.int16).copy()).view(torch.bfloat16)


def bits_to_torch(bits: np.ndarray, dt_name: str) -> torch.Tensor:
    if dt_name == "fp32":
        return torch.from_numpy(bits.view(np.float32).copy())
    t = torch.from_numpy(bits.view(np.int16).copy())
    return t.view(torch.bfloat16 if dt_name == "bf16" else torch.float16)


def torch_bits(t: torch.Tensor) -> np.ndarray:
    t = t.detach().cpu().contiguous()
    if t.dtype == torch.float32:
        return t.numpy().view(np.uint32)
    return t.view(torch.int16).numpy().view(np.uint16)


def opt(x):
    return None if np.isnan(x) else float(x)


def rel_err(a, b, floor=1.0):
    """max |a-b| / max(|b|, floor): relative where the reference is large, absolute near zero."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if a.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor)))
