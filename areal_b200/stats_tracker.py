"""Host-side mirror of `areal/utils/stats_tracker.py` (DistributedStatsTracker, :30-303).

Same API and error behaviour (`scope`, `denominator`, `stat`, `scalar`, `record_timing`, `export`),
different mechanics: the reference clones every recorded tensor (:92, :124) and, at export time, runs
`torch.where` + a reduction per key, per micro-batch and per statistic (:210-289) and one to four
all-reduces per key.  Here a record call reduces its tensors right away -- all keys of the call in
ONE launch of `areal_b200_masked_stats`, folded into fp64 running totals that stay on the device
(no clone, no host sync) -- and `export` does one device->host copy, or three all-reduces
(SUM, MIN, MAX) over all keys together.  CUDA tensors only, no fallback.
"""
from __future__ import annotations

import ctypes
import time
from collections import defaultdict
from contextlib import contextmanager
from enum import Enum, auto
from threading import Lock
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from . import _lib
from ._lib import call, stream_ptr

MAX_KEYS_PER_LAUNCH = 32


class ReduceType(Enum):
    AVG_MIN_MAX = auto()
    AVG = auto()
    SUM = auto()
    MIN = auto()
    MAX = auto()
    SCALAR = auto()


_IDENTITY = (0.0, 0.0, float("inf"), -float("inf"))


class DistributedStatsTracker:
    def __init__(self, name: str = ""):
        self.lock = Lock()
        self.scope_stack: List[str] = []
        if name:
            self.scope_stack.append(name.strip("/"))
        self.denominators: Dict[str, str] = {}      # key -> denominator key
        self.reduce_types: Dict[str, ReduceType] = {}
        self.scalars = defaultdict(list)            # SCALAR keys: python floats
        self._reset_device_state()

    # ---- device-side running totals ---------------------------------------------------------------
    def _reset_device_state(self):
        self._acc: Optional[torch.Tensor] = None    # f64 [capacity, 4] = (sum, count, min, max) per key
        self._slot: Dict[str, int] = {}
        self._next_slot = 0
        self._last_mask: Dict[str, torch.Tensor] = {}   # denominator key -> mask of the current micro-batch
        self._shapes: Dict[str, List[Tuple[int, ...]]] = defaultdict(list)

    def _slot_of(self, full_key: str, device) -> int:
        if full_key in self._slot:
            return self._slot[full_key]
        n = self._next_slot
        self._next_slot += 1
        if self._acc is None or n >= self._acc.shape[0]:
            cap = 64 if self._acc is None else 2 * self._acc.shape[0]
            new = torch.tensor(_IDENTITY, dtype=torch.float64, device=device).repeat(cap, 1)
            if self._acc is not None:
                new[: self._acc.shape[0]] = self._acc
            self._acc = new
        self._slot[full_key] = n
        return n

    def _reduce(self, entries):
        """entries: (full_key, value f32 tensor or None, mask bool tensor or None); one launch per
        group of <= 32 entries with the same element count."""
        if not entries:
            return
        dev = _lib.require_cuda(*[t for e in entries for t in e[1:] if t is not None])
        by_n = defaultdict(list)
        for key, v, m in entries:
            n = (v if v is not None else m).numel()
            slot = self._slot_of(key, dev)        # may grow self._acc: resolve every slot first
            by_n[n].append((slot, v, m))
        base = self._acc.data_ptr()
        with torch.cuda.device(dev):
            for n, group in by_n.items():
                for i in range(0, len(group), MAX_KEYS_PER_LAUNCH):
                    part = group[i:i + MAX_KEYS_PER_LAUNCH]
                    keep = []
                    vals = (ctypes.c_void_p * len(part))()
                    masks = (ctypes.c_void_p * len(part))()
                    outs = (ctypes.c_void_p * len(part))()
                    u8_bits = 0
                    for j, (slot, v, m) in enumerate(part):
                        if v is not None:
                            v = v.detach().reshape(-1)
                            v = v if v.is_contiguous() else v.contiguous()
                            if v.dtype in (torch.bool, torch.uint8):
                                v, u8_bits = v.view(torch.uint8), u8_bits | (1 << j)
                            keep.append(v)
                            vals[j] = v.data_ptr()
                        if m is not None:
                            m = m.detach().reshape(-1)
                            m = (m if m.is_contiguous() else m.contiguous()).view(torch.uint8)
                            keep.append(m)
                            masks[j] = m.data_ptr()
                        outs[j] = base + 32 * slot
                    nbytes = _lib.lib().areal_b200_masked_stats_workspace_bytes(len(part), n)
                    ws = _lib.workspace("masked_stats", dev, nbytes)
                    call("areal_b200_masked_stats", vals, masks, outs, u8_bits, len(part), n, 1, _lib.ptr(ws), ws.numel(),
                         stream_ptr(dev))

    # ---- the reference's API ----------------------------------------------------------------------
    def scope(self, name):
        with self.lock:
            return self.Scope(self, name)

    class Scope:
        def __init__(self, tracker, name):
            self.tracker = tracker
            self.name = name.strip("/")

        def __enter__(self):
            self.tracker.scope_stack.append(self.name)
            return self

        def __exit__(self, exc_type, exc_val, exc_tb):
            self.tracker.scope_stack.pop()

    def _get_full_key(self, key):
        if not self.scope_stack:
            return key
        return "/".join(self.scope_stack + [key])

    @contextmanager
    def disable_scope(self):
        tmp = self.scope_stack
        self.scope_stack = []
        yield
        self.scope_stack = tmp

    @contextmanager
    def record_timing(self, key):
        start_time = time.perf_counter()
        try:
            yield
        finally:
            with self.lock:
                full_key = f"timeperf/{key}"      # fixed scope, stats_tracker.py:79
                self._set_reduce_type(full_key, ReduceType.SCALAR)
                self.scalars[full_key].append(time.perf_counter() - start_time)

    def _check_denominators(self, kwargs):
        for key, value in kwargs.items():
            if not isinstance(value, torch.Tensor) or value.dtype != torch.bool:
                raise ValueError(f"`{key}` must be a pytorch bool tensor: {value.dtype}")
            if value.numel() == 0:
                raise ValueError(f"`{key}` must be non-empty")

    def _denominator_entries(self, kwargs):
        entries = []
        for key, value in kwargs.items():
            full_key = self._get_full_key(key)
            self._set_reduce_type(full_key, ReduceType.SUM)
            self._last_mask[full_key] = value
            self._shapes[full_key].append(tuple(value.shape))
            entries.append((full_key, None, value))
        return entries

    def denominator(self, **kwargs):
        with self.lock:
            self._check_denominators(kwargs)
            self._reduce(self._denominator_entries(kwargs))

    def scalar(self, **kwargs):
        with self.lock:
            for key, value in kwargs.items():
                full_key = self._get_full_key(key)
                self._set_reduce_type(full_key, ReduceType.SCALAR)
                self.scalars[full_key].append(float(value))

    def _stat_entries(self, denominator, reduce_type, kwargs, allow_bool=False):
        entries = []
        for key, value in kwargs.items():
            if allow_bool and isinstance(value, torch.Tensor) and value.dtype == torch.bool:
                pass                              # record(): `mask.float()` happens inside the kernel
            elif not isinstance(value, torch.Tensor) or value.dtype != torch.float:
                raise ValueError(f"`{key}` should be a pytorch float tensor: {value.dtype}")
            if value.numel() == 0:
                raise ValueError(f"`{key}` should be non-empty")
            if reduce_type == ReduceType.SCALAR:
                raise ValueError("Cannot use the scalar reduce type for a tensor")
            full_key = self._get_full_key(key)
            denorm = self._get_full_key(denominator)
            if denorm not in self._last_mask:
                raise ValueError(f"Denominator `{denorm}` does not exist")
            for x, y in zip(self._shapes[denorm], self._shapes[full_key] + [tuple(value.shape)]):
                assert x == y, (x, y)
            self.denominators[full_key] = denorm
            if reduce_type is None:
                reduce_type = ReduceType.AVG_MIN_MAX
            self._set_reduce_type(full_key, reduce_type)
            self._shapes[full_key].append(tuple(value.shape))
            entries.append((full_key, value, self._last_mask[denorm]))
        return entries

    def stat(self, denominator: str, reduce_type: ReduceType | None = None, **kwargs):
        with self.lock:
            self._reduce(self._stat_entries(denominator, reduce_type, kwargs))

    def record(self, denominators: Dict[str, torch.Tensor], stats: List[Tuple[str, Dict[str, torch.Tensor]]]):
        """Several `denominator(...)` / `stat(denominator=..., ...)` calls of one micro-batch as ONE
        launch (no counterpart in the reference; what `areal_b200.actor.grpo_loss_fn` uses).  Values may
        be bool tensors where the reference records `mask.float()`."""
        with self.lock:
            self._check_denominators(denominators)
            entries = self._denominator_entries(denominators)
            for denom, kw in stats:
                entries += self._stat_entries(denom, None, kw, allow_bool=True)
            self._reduce(entries)

    def _set_reduce_type(self, key, reduce_type):
        if not isinstance(reduce_type, ReduceType):
            raise ValueError("reduce_type must be a ReduceType enum")
        self.reduce_types[key] = reduce_type

    # ---- export -----------------------------------------------------------------------------------
    def _totals(self, keys, reduce_group):
        """{key: (sum, count, min, max)} for tensor keys after the cross-rank reduction."""
        tensor_keys = [k for k in keys if self.reduce_types.get(k, ReduceType.SCALAR) != ReduceType.SCALAR]
        if not tensor_keys:
            return {}
        if self._acc is not None:
            dev, have = self._acc.device, self._acc
        else:                                     # this rank recorded nothing: identities only
            dev, have = self._reduce_device(), None
        idx = [self._slot.get(k, -1) for k in tensor_keys]
        table = torch.tensor(_IDENTITY, dtype=torch.float64, device=dev).repeat(len(tensor_keys), 1)
        sel = [i for i, s in enumerate(idx) if s >= 0]
        if sel and have is not None:
            table[torch.tensor(sel, device=dev)] = have[torch.tensor([idx[i] for i in sel], device=dev)]
        if reduce_group is not None:
            sums, mins, maxs = table[:, :2].contiguous(), table[:, 2].contiguous(), table[:, 3].contiguous()
            dist.all_reduce(sums, group=reduce_group)
            dist.all_reduce(mins, op=dist.ReduceOp.MIN, group=reduce_group)
            dist.all_reduce(maxs, op=dist.ReduceOp.MAX, group=reduce_group)
            table = torch.cat([sums, mins[:, None], maxs[:, None]], dim=1)
        host = table.cpu().tolist()                # the one device->host copy of export
        return dict(zip(tensor_keys, host))

    def export(self, key=None, reduce_group=None, reset=True) -> Dict[str, float]:
        with self.lock:
            if key is not None:
                full_key = self._get_full_key(key)
                result = self._aggregate(full_key, self._totals([full_key], reduce_group),
                                         self._scalar_totals([full_key], reduce_group))
                if reset:
                    self.denominators.pop(full_key, None)
                    self.reduce_types.pop(full_key, None)
                    self.scalars.pop(full_key, None)
                    self._shapes.pop(full_key, None)
                    self._last_mask.pop(full_key, None)
                    self._slot.pop(full_key, None)    # its slot is retired, a new record starts a fresh one
                return result
            keys = list(dict.fromkeys(list(self._slot.keys()) + list(self.scalars.keys())))
            if reduce_group is not None:
                all_keys = [None for _ in range(dist.get_world_size(reduce_group))]
                dist.all_gather_object(all_keys, (keys, {k: self.reduce_types[k].name for k in keys},
                                                  dict(self.denominators)), group=reduce_group)
                for ks, rts, dens in all_keys:    # ranks that recorded nothing still take part
                    for k in ks:
                        self.reduce_types.setdefault(k, ReduceType[rts[k]])
                    for k, d in dens.items():
                        self.denominators.setdefault(k, d)
                keys = sorted(set(k for ks, _, _ in all_keys for k in ks))
            totals = self._totals(keys, reduce_group)
            scalar_totals = self._scalar_totals(keys, reduce_group)
            results = {}
            for k in keys:
                results.update(self._aggregate(k, totals, scalar_totals))
            if reset:
                self.denominators = {}
                self.reduce_types = {}
                self.scalars = defaultdict(list)
                self._reset_device_state()
            return results

    @staticmethod
    def _reduce_device():
        """Where the small export tables live: the current CUDA device whenever there is one (an NCCL
        group cannot reduce CPU tensors; the reference uses the platform device, stats_tracker.py:189-201)."""
        if torch.cuda.is_available():
            return torch.device("cuda", torch.cuda.current_device())
        return torch.device("cpu")

    def _scalar_totals(self, keys, reduce_group):
        """{key: (sum, count)} of the SCALAR keys; ONE all-reduce of a [K, 2] table for all of them."""
        scalar_keys = [k for k in keys if self.reduce_types.get(k, ReduceType.SCALAR) == ReduceType.SCALAR]
        if not scalar_keys:
            return {}
        rows = [[float(sum(self.scalars.get(k, []))), float(len(self.scalars.get(k, [])))] for k in scalar_keys]
        if reduce_group is not None:
            buf = torch.tensor(rows, dtype=torch.float64, device=self._reduce_device())
            dist.all_reduce(buf, group=reduce_group)
            rows = buf.cpu().tolist()
        return {k: (r[0], r[1]) for k, r in zip(scalar_keys, rows)}

    def _aggregate(self, key, totals, scalar_totals):
        reduce_type = self.reduce_types.get(key, ReduceType.SCALAR)
        result = {}
        if reduce_type == ReduceType.SCALAR:
            value, cnt = scalar_totals[key]
            result[key] = float(value / cnt)
            return result
        total, count, mn, mx = totals[key]

        def avg():                                # x / d, None when the denominator selected nothing (:247-248)
            return None if count == 0 else total / count

        def finite(x):
            return None if x in (float("inf"), -float("inf")) else float(x)

        if reduce_type == ReduceType.AVG_MIN_MAX:
            result["/".join([key, "avg"])] = avg()
            result["/".join([key, "min"])] = finite(mn)
            result["/".join([key, "max"])] = finite(mx)
        elif reduce_type == ReduceType.AVG:
            result[key] = avg()
        elif reduce_type == ReduceType.SUM:
            result[key] = float(total)
        elif reduce_type == ReduceType.MIN:
            result[key] = finite(mn)
        elif reduce_type == ReduceType.MAX:
            result[key] = finite(mx)
        else:
            raise ValueError(f"Unknown reduce type: {reduce_type}")
        return {k: v for k, v in result.items() if v is not None}


DEFAULT_TRACKER = DistributedStatsTracker()
stat = DEFAULT_TRACKER.stat
denominator = DEFAULT_TRACKER.denominator
export = DEFAULT_TRACKER.export
scope = DEFAULT_TRACKER.scope
scalar = DEFAULT_TRACKER.scalar
record_timing = DEFAULT_TRACKER.record_timing
record = DEFAULT_TRACKER.record

TRACKERS = {"": DEFAULT_TRACKER}
LOCK = Lock()


def get(name: str = ""):
    with LOCK:
        if name not in TRACKERS:
            TRACKERS[name] = DistributedStatsTracker(name)
        return TRACKERS[name]


def export_all(reduce_group=None, reset=True) -> Dict[str, float]:
    """All trackers, in the reference's order (stats_tracker.py:298-317); duplicate keys: last wins."""
    stat = {}
    tracker_keys = list(TRACKERS.keys())
    if reduce_group is not None:
        all_trackers = [None for _ in range(dist.get_world_size(reduce_group))]
        dist.all_gather_object(all_trackers, tracker_keys, group=reduce_group)
        tracker_keys = sorted(set(k for ks in all_trackers for k in ks))
    for tracker_key in tracker_keys:
        stat.update(get(tracker_key).export(reduce_group=reduce_group, reset=reset))
    return stat
