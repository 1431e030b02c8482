"""Lets the UNMODIFIED train engines reach the single-pass actor kernel.

The engines compute the scalar that multiplies a micro-batch's loss only AFTER `loss_fn` has returned
(areal/engine/fsdp_engine.py:578-586: `loss = loss_fn(logits, mb)`, then
`loss_scale = loss_weight_fn(mb) / total_loss_weight * dp_size`, `loss *= loss_scale`, `loss.backward()`),
while `areal_b200_fused_actor_fwd_bwd` writes `dlogits` during the forward and therefore needs that
scalar BEFORE the pass.  Everything the scalar is made of is available earlier, though: the engine calls
`loss_weight_fn` on every micro-batch and all-reduces the sum over its data-parallel group before the
first forward (:518-526).

`install_loss_scale_hook(engine)` wraps the engine's bound `train_batch` (no source edit): the wrapper
passes wrapped `loss_fn` / `loss_weight_fn` to the original method, collects the weights the engine asks
for, obtains the total -- the engine's own all-reduced tensor when the weights can be tracked to it
(`_TrackedWeight`), else by repeating the engine's arithmetic (same tensor ops, same all-reduce group) -- and, for
each micro-batch, announces `loss_weight_fn(mb) / total * dp_size` to `areal_b200.actor.grpo_loss_fn`
through a context variable before calling the user's `loss_fn`.  The announced value is a promise, not a
requirement: the backward of the single-pass path compares it on the device with the scalar that really
arrives and rescales the stored gradient if they differ (areal_b200_rescale_grad_if_needed), so a wrong
or missing announcement costs time, never correctness.
"""
from __future__ import annotations

import contextvars
from typing import Any, Callable, Dict, Optional

import torch
import torch.distributed as dist

# read by areal_b200.actor.grpo_loss_fn when its `loss_scale` argument is None
ANNOUNCED_LOSS_SCALE: contextvars.ContextVar = contextvars.ContextVar("areal_b200_announced_loss_scale", default=None)


def announced_loss_scale():
    """The scale announced for the loss_fn call in progress (a 1-element f32 tensor) or None."""
    return ANNOUNCED_LOSS_SCALE.get()


class _TrackedWeight(torch.Tensor):
    """A loss weight that remembers what is derived from it.  The engine builds its total as
    `torch.stack([loss_weight_fn(mb) ...]).sum().detach().clone().to(float32, device)` and all-reduces THAT tensor in
    place (areal/engine/fsdp_engine.py:518-526).  Every op on a tracked tensor returns a tracked tensor and reports it
    to the announcer, so the last one reported in phase 1 is the very tensor the engine all-reduces -- the announcer
    can read the engine's total (stream-ordered, after the engine's own all-reduce) instead of reducing a second
    time.  Purely an optimisation: if the chain is not recognised, the announcer reduces for itself."""

    _announcer = None

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        ann = None
        for a in args:
            if isinstance(a, _TrackedWeight):
                ann = a._announcer
            elif isinstance(a, (list, tuple)):
                for x in a:
                    if isinstance(x, _TrackedWeight):
                        ann = x._announcer
        with torch._C.DisableTorchFunctionSubclass():
            out = func(*args, **kwargs)
        if ann is not None and isinstance(out, torch.Tensor) and ann.total is None:
            if not isinstance(out, _TrackedWeight):
                out = out.as_subclass(_TrackedWeight)
            out._announcer = ann
            if out.dim() == 0 and out.dtype == torch.float32:     # the total itself (not e.g. `total != 0`)
                ann._last_derived = out
        return out


class LossScaleAnnouncer:
    """Wraps one `(loss_fn, loss_weight_fn)` pair for one `train_batch` call.

    Phase 1 (engine, before any forward): `loss_weight_fn(mb)` for every micro-batch -> collected.
    Phase 2 (engine, per micro-batch): `loss_fn(logits, mb)` -> on the first call the total is formed exactly
    as the engine forms it (stack, sum, float32, all-reduce over the data-parallel group), then the
    scale of this micro-batch is announced around the user's loss_fn."""

    def __init__(self, loss_fn: Callable, loss_weight_fn: Callable, dp_group=None, dp_size: Optional[int] = None,
                 device=None, extra_scale: Optional[Callable[[], float]] = None):
        self.user_loss_fn, self.user_weight_fn = loss_fn, loss_weight_fn
        self.dp_group, self.dp_size, self.device = dp_group, dp_size, device
        self.extra_scale = extra_scale    # e.g. Megatron's optimizer loss scale (megatron_engine.py:900)
        self.weights = []
        self._by_mb = {}
        self._last_derived = None         # see _TrackedWeight
        self.track_engine_total = True
        self.used_engine_total = False    # for tests / diagnostics
        self.total = None
        self.last_announced = None        # for tests / diagnostics
        self.n_announced = 0

    def loss_weight_fn(self, mb):
        w = self.user_weight_fn(mb)
        if self.total is None:            # phase 1: the engine is building total_loss_weight
            self.weights.append(w)
            self._by_mb[id(mb)] = (mb, w)  # (the object is kept so that its id cannot be recycled)
            if self.track_engine_total and isinstance(w, torch.Tensor) and type(w) is torch.Tensor:
                t = w.as_subclass(_TrackedWeight)
                t._announcer = self
                return t
        return w

    def _total(self, like: torch.Tensor) -> torch.Tensor:
        dev = self.device if self.device is not None else like.device
        t = self._last_derived
        if (t is not None and t.dim() == 0 and t.dtype == torch.float32 and t.device == torch.device(dev)
                and len(self.weights) > 0):
            # the engine's own total_loss_weight, already all-reduced by the engine (stream-ordered before this read)
            self.used_engine_total = True
            return t.as_subclass(torch.Tensor)
        # the engine's own expression (areal/engine/fsdp_engine.py:518-526)
        total = torch.stack([torch.as_tensor(w) for w in self.weights]).sum().detach().clone().to(
            dtype=torch.float32, device=dev)
        if dist.is_available() and dist.is_initialized() and self._world() > 1:
            dist.all_reduce(total, group=self.dp_group)
        return total

    def _world(self) -> int:
        if self.dp_size is not None:
            return int(self.dp_size)
        return dist.get_world_size(self.dp_group) if (dist.is_available() and dist.is_initialized()) else 1

    def loss_fn(self, logits, mb):
        if self.total is None:
            if not self.weights:          # an engine that never asked for the weights: nothing to announce
                return self.user_loss_fn(logits, mb)
            self.total = self._total(logits)
        seen = self._by_mb.get(id(mb))    # the weight the engine already asked for in phase 1 (same micro-batch object)
        w = seen[1] if (seen is not None and seen[0] is mb) else self.user_weight_fn(mb)
        scale = torch.as_tensor(w).to(self.total.device) / self.total          # :580
        scale = scale * self._world()                                        # :583 (dp_size)
        if self.extra_scale is not None:
            scale = scale * float(self.extra_scale())
        scale = scale.detach().to(torch.float32).reshape(1)
        self.last_announced = scale
        self.n_announced += 1
        token = ANNOUNCED_LOSS_SCALE.set(scale)
        try:
            return self.user_loss_fn(logits, mb)
        finally:
            ANNOUNCED_LOSS_SCALE.reset(token)


def install_loss_scale_hook(engine, dp_group=None, dp_size: Optional[int] = None,
                            extra_scale: Optional[Callable[[], float]] = None):
    """Wrap `engine.train_batch(input_, loss_fn, loss_weight_fn)` (areal/api/engine_api.py:242-275) so that
    every call runs with an announcer around its two callbacks.  `dp_group` / `dp_size` default to what
    the FSDP engine uses (`engine.dp_group`, `engine.parallel_helper.dp_size`; fsdp_engine.py:526, 583) and
    to `engine.data_parallel_group` / its world size otherwise; `extra_scale()` is one more factor the engine
    applies (the Megatron engine multiplies by its optimizer's loss scale, megatron_engine.py:899-900).  Returns the engine; `engine.train_batch
    .__wrapped__` is the original bound method, `engine.train_batch.last_announcer` the most recent
    announcer."""
    original = engine.train_batch
    if getattr(original, "_areal_b200_hooked", False):
        return engine

    def resolve():
        g = dp_group
        if g is None:
            g = getattr(engine, "dp_group", None)
        if g is None:
            g = getattr(engine, "data_parallel_group", None)
        n = dp_size
        if n is None:
            helper = getattr(engine, "parallel_helper", None)
            n = getattr(helper, "dp_size", None)
        return g, n

    def train_batch(input_: Dict[str, Any], loss_fn: Callable, loss_weight_fn: Callable, *args, **kwargs):
        g, n = resolve()
        ann = LossScaleAnnouncer(loss_fn, loss_weight_fn, dp_group=g, dp_size=n, device=getattr(engine, "device", None),
                                 extra_scale=extra_scale)
        train_batch.last_announcer = ann
        return original(input_, ann.loss_fn, ann.loss_weight_fn, *args, **kwargs)

    train_batch.__wrapped__ = original
    train_batch._areal_b200_hooked = True
    train_batch.last_announcer = None
    engine.train_batch = train_batch
    return engine
