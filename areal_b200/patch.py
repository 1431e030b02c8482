"""Installs areal_b200 behind the reference's own import sites, so the callers do not change.

  install_shims()   registers `cugae`, `realhf._C.cugae`, `realhf._C.interval_op_cuda`,
                    `realhf._C.interval_op` in sys.modules (the names the reference imports at
                    realhf/impl/model/utils/ppo_functional.py:359-362 and
                    realhf/impl/model/nn/flatten_param.py:31,113,162).
  patch_areal()     rebinds gather_logprobs, gather_logprobs_entropy, ppo_actor_loss_fn,
                    ppo_critic_loss_fn in `areal.utils.functional` AND in the modules that bound them
                    with from-imports (areal/engine/ppo/actor.py:16-22, critic.py:11,
                    areal/engine/sft/lm_engine.py:10, recipe/AEnt/actor.py:15-17).
  patch_realhf()    rebinds the GAE / loss / interval entry points of realhf's functional modules.
  patch_areal() also rebinds `Normalization` / `masked_normalization` where they are already imported.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

from . import data as _D
from . import flatten_param as _fp
from . import functional as _F
from . import ppo_functional as _P

_AREAL_NAMES = ("gather_logprobs", "gather_logprobs_entropy", "ppo_actor_loss_fn", "ppo_critic_loss_fn")
_AREAL_IMPORT_SITES = ("areal.utils.functional", "areal.engine.ppo.actor", "areal.engine.ppo.critic",
                       "areal.engine.sft.lm_engine", "recipe.AEnt.actor")


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    return m


def install_shims():
    shim_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "shims")
    if shim_dir not in sys.path:
        sys.path.insert(0, shim_dir)   # makes `cugae` importable and visible to importlib.metadata
    cugae_c = _module("realhf._C.cugae", gae_1d_nolp_misalign=_P.gae_1d_nolp_misalign,
                      gae_2d_olp=_P.gae_2d_olp, gae_2d_nolp=_P.gae_2d_nolp)
    iop_cuda = _module("realhf._C.interval_op_cuda",
                       **{f"{op}_intervals_{dt}": getattr(_fp, f"{op}_intervals_{dt}")
                          for op in ("slice", "set") for dt in ("fp32", "fp16", "bf16")})
    iop = _module("realhf._C.interval_op", merge_intervals=_fp.merge_intervals)
    parent = sys.modules.get("realhf._C")
    if parent is None:
        parent = _module("realhf._C")
        parent.__path__ = []
        sys.modules["realhf._C"] = parent
        if "realhf" in sys.modules:
            setattr(sys.modules["realhf"], "_C", parent)
    for short, mod in (("cugae", cugae_c), ("interval_op_cuda", iop_cuda), ("interval_op", iop)):
        sys.modules[mod.__name__] = mod
        setattr(parent, short, mod)
    return dict(cugae=cugae_c, interval_op_cuda=iop_cuda, interval_op=iop)


def patch_areal(extra_modules=()):
    """Rebind the four functional ops wherever they are already imported; returns the patched module
    names.  Modules not imported yet are patched through `areal.utils.functional` when they import."""
    patched = []
    # the first "next" row: Normalization / masked_normalization (areal/utils/data.py, functional.py)
    for name in ("areal.utils.data", "areal.engine.ppo.actor", "areal.engine.ppo.critic", "areal.utils.functional"):
        mod = sys.modules.get(name)
        if mod is None:
            continue
        if hasattr(mod, "Normalization"):
            setattr(mod, "Normalization", _D.Normalization)
        if hasattr(mod, "masked_normalization"):
            setattr(mod, "masked_normalization", _D.masked_normalization)
        if hasattr(mod, "dynamic_sampling"):
            setattr(mod, "dynamic_sampling", _D.dynamic_sampling)
    # reward construction + GAE of the actor (areal/engine/ppo/actor.py:72-164) as one launch
    mod = sys.modules.get("areal.engine.ppo.actor")
    if mod is not None and hasattr(mod, "PPOActor"):
        from . import actor as _A
        mod.PPOActor.compute_advantages = _A.compute_advantages
        patched.append("areal.engine.ppo.actor.PPOActor.compute_advantages")
    # stats epilogue: the tracker that reduces at record time, and the grpo_loss_fn that feeds it
    st = sys.modules.get("areal.utils.stats_tracker")
    if st is not None:
        from . import stats_tracker as _ST
        for name in ("ReduceType", "DistributedStatsTracker", "DEFAULT_TRACKER", "stat", "denominator", "export",
                     "scope", "scalar", "record_timing", "TRACKERS", "LOCK", "get", "export_all"):
            setattr(st, name, getattr(_ST, name))
        patched.append("areal.utils.stats_tracker")
        if mod is not None and hasattr(mod, "grpo_loss_fn"):
            from . import actor as _A
            mod.grpo_loss_fn = _A.grpo_loss_fn
            patched.append("areal.engine.ppo.actor.grpo_loss_fn")
        smod = sys.modules.get("areal.engine.sft.lm_engine")
        if smod is not None and hasattr(smod, "compute_packed_sft_loss"):
            from . import sft as _S
            smod.compute_packed_sft_loss = _S.compute_packed_sft_loss
            patched.append("areal.engine.sft.lm_engine.compute_packed_sft_loss")
        cmod = sys.modules.get("areal.engine.ppo.critic")
        if cmod is not None and hasattr(cmod, "ppo_loss_fn"):
            from . import critic as _C
            cmod.ppo_loss_fn = _C.ppo_loss_fn
            patched.append("areal.engine.ppo.critic.ppo_loss_fn")
    # AEnt recipe: clamped entropy without the CPU top-k (recipe/AEnt/functional.py, actor.py:21)
    from . import aent_functional as _AE
    for name in ("recipe.AEnt.functional", "recipe.AEnt.actor"):
        m = sys.modules.get(name)
        if m is not None and hasattr(m, "gather_logprobs_clamped_entropy"):
            m.gather_logprobs_clamped_entropy = _AE.gather_logprobs_clamped_entropy
            if hasattr(m, "clamped_softmax_entropy"):
                m.clamped_softmax_entropy = _AE.clamped_softmax_entropy
            for fn in ("aent_grpo_loss_fn", "clamped_entropy_loss_fn"):
                if hasattr(m, fn):
                    setattr(m, fn, getattr(_AE, fn))
            patched.append(name)
    for name in tuple(_AREAL_IMPORT_SITES) + tuple(extra_modules):
        mod = sys.modules.get(name)
        if mod is None:
            if name != "areal.utils.functional":
                continue
            try:
                mod = importlib.import_module(name)
            except Exception:
                continue
        for fn in _AREAL_NAMES:
            if hasattr(mod, fn) or name == "areal.utils.functional":
                setattr(mod, fn, getattr(_F, fn))
        patched.append(name)
    return patched


def patch_realhf():
    patched = []
    m = sys.modules.get("realhf.impl.model.utils.ppo_functional")
    if m is not None:
        for fn in ("cugae1d_nolp_misalign_func", "get_packed_advantages_and_returns", "actor_loss_fn",
                   "critic_loss_fn", "get_packed_rewards", "get_packed_reward_dense"):
            setattr(m, fn, getattr(_P, fn))
        patched.append(m.__name__)
    m = sys.modules.get("realhf.impl.model.utils.functional")
    if m is not None:
        from . import realhf_functional as _RF
        for fn in ("build_shift_one_indices", "build_leave_one_indices", "gather_logprobs", "gather_shifted_log_probs",
                   "gather_packed_shifted_log_probs", "masked_normalization"):
            if hasattr(m, fn):
                setattr(m, fn, getattr(_RF, fn))
        patched.append(m.__name__)
    m = sys.modules.get("realhf.impl.model.modules.rms")
    if m is not None:
        from . import rms as _RMS
        for cls in ("ExponentialRunningMeanStd", "MovingAverageRunningMeanStd"):
            setattr(m, cls, getattr(_RMS, cls))
        patched.append(m.__name__)
    m = sys.modules.get("realhf.impl.model.nn.flatten_param")
    if m is not None:
        for fn in ("slice_intervals", "set_intervals", "merge_intervals"):
            setattr(m, fn, getattr(_fp, fn))
        patched.append(m.__name__)
    install_shims()
    return patched
