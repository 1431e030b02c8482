"""Host-side mirror of the legacy running value statistics,
`realhf/impl/model/modules/rms.py` (`ExponentialRunningMeanStd` :19-109, `MovingAverageRunningMeanStd`
:112-183): same class names (so the name-mangled parameter names in a state_dict match), same methods
(`update`, `mean_std`, `normalize`, `denormalize`, `reset_parameters`), same dtype path
(`high_precision` -> float64 state and arithmetic).

`update` is two launches (masked moments in fp64, scalar state update) instead of three reductions and
about ten scalar kernels, with ONE all-reduce of three numbers under data parallelism instead of three;
`normalize` / `denormalize` are one elementwise launch each.  CUDA tensors only, no fallback.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _lib
from ._lib import call, ptr, stream_ptr


class _RunningMeanStd(nn.Module):
    _MODE = 0

    def __init__(self, high_precision: bool = True, device=None, group=None):
        super().__init__()
        self._dtype = torch.float64 if high_precision else torch.float32
        self._group = group                     # the reference uses constants.data_parallel_group()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self._dev = dev
        # [mean, mean_sq, debiasing term | accumulated denominator]: one buffer the kernels update in place;
        # the reference's parameters are views of it
        self._state = torch.zeros(3, dtype=self._dtype, device=dev)

    def forward(self, *args, **kwargs):
        raise NotImplementedError()             # as upstream: the meaning would be ambiguous

    def _view(self, i) -> torch.Tensor:
        return self._state[i:i + 1]

    @torch.no_grad()
    def update(self, x: torch.Tensor, mask: Optional[torch.Tensor] = None):
        dev = _lib.require_cuda(x, mask)
        xf = x.float().contiguous().reshape(-1)
        mf = None if mask is None else mask.float().contiguous().reshape(-1)
        if mf is not None and mf.numel() != xf.numel():
            mf = mask.float().expand_as(x).contiguous().reshape(-1)
        N = xf.numel()
        P = _lib.lib().areal_b200_norm_num_partials(N)
        partials = torch.empty(P, 3, dtype=torch.float64, device=dev)
        with torch.cuda.device(dev):
            call("areal_b200_masked_moments", ptr(xf), ptr(mf), 1, N, P, ptr(partials), stream_ptr(dev))
            if dist.is_initialized():
                partials = partials.sum(0, keepdim=True)         # (factor, x_sum, x_sum_sq)
                dist.all_reduce(partials, op=dist.ReduceOp.SUM, group=self._group)
                P = 1
            call("areal_b200_rms_update", ptr(partials), P, ptr(self._state), int(self._dtype == torch.float64),
                 self._MODE, float(getattr(self, "_beta", 0.0)), stream_ptr(dev))

    def _apply_kernel(self, x: torch.Tensor, denorm: bool) -> torch.Tensor:
        dev = _lib.require_cuda(x)
        xf = x.float().contiguous()
        out = torch.empty_like(xf)
        with torch.cuda.device(dev):
            call("areal_b200_rms_apply", ptr(xf), xf.numel(), ptr(self._state), int(self._dtype == torch.float64),
                 self._MODE, float(getattr(self, "_eps", 0.0)), int(denorm), ptr(out), stream_ptr(dev))
        return out

    @torch.no_grad()
    def normalize(self, x):
        return self._apply_kernel(x, False)

    @torch.no_grad()
    def denormalize(self, x):
        return self._apply_kernel(x, True)


class ExponentialRunningMeanStd(_RunningMeanStd):
    _MODE = 0

    def __init__(self, beta=0.999, epsilon=1e-5, high_precision=True, device=None, group=None):
        super().__init__(high_precision, device, group)
        self._beta, self._eps = float(beta), float(epsilon)
        self.__mean = nn.Parameter(self._view(0), requires_grad=False)
        self.__mean_sq = nn.Parameter(self._view(1), requires_grad=False)
        self.__debiasing_term = nn.Parameter(self._view(2), requires_grad=False)
        self.reset_parameters()

    def reset_parameters(self):
        self._state.zero_()

    def _load_from_state_dict(self, *args, **kwargs):
        super()._load_from_state_dict(*args, **kwargs)
        self._rebind()

    def _rebind(self):
        """Parameters replaced by a load: copy their values into the kernels' buffer and view it again."""
        vals = torch.cat([self.__mean.data.reshape(1), self.__mean_sq.data.reshape(1),
                          self.__debiasing_term.data.reshape(1)]).to(self._state)
        self._state.copy_(vals)
        self.__mean.data, self.__mean_sq.data, self.__debiasing_term.data = self._view(0), self._view(1), self._view(2)

    @torch.no_grad()
    def mean_std(self):
        d = self.__debiasing_term.clamp(min=self._eps)
        mean = self.__mean / d
        var = (self.__mean_sq / d - mean ** 2).clamp(min=1e-2)
        return mean, var.sqrt()


class MovingAverageRunningMeanStd(_RunningMeanStd):
    _MODE = 1

    def __init__(self, high_precision=True, device=None, group=None):
        super().__init__(high_precision, device, group)
        self.__mean = nn.Parameter(self._view(0), requires_grad=False)
        self.__mean_sq = nn.Parameter(self._view(1), requires_grad=False)
        self.reset_parameters()

    def reset_parameters(self):
        self._state.zero_()

    def _load_from_state_dict(self, *args, **kwargs):
        super()._load_from_state_dict(*args, **kwargs)
        vals = torch.cat([self.__mean.data.reshape(1), self.__mean_sq.data.reshape(1)]).to(self._state)
        self._state[:2].copy_(vals)
        self.__mean.data, self.__mean_sq.data = self._view(0), self._view(1)

    @torch.no_grad()
    def mean_std(self):
        return self.__mean.clone(), (self.__mean_sq - self.__mean ** 2).clamp(min=1e-2).sqrt()
