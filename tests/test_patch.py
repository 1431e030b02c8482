"""CPU: the drop-in plumbing -- shims are importable under the names the reference imports, the
`cugae` distribution is visible to importlib.metadata (realhf/base/pkg_version.py:7-11), and
patch_areal() rebinds from-import sites."""
import importlib.metadata
import sys
import types


def test_install_shims_and_dist_info():
    from areal_b200 import patch
    mods = patch.install_shims()
    import cugae
    assert callable(cugae.cugae1d_nolp_misalign_func)
    assert importlib.metadata.version("cugae") == "0.1.0"
    from realhf._C.cugae import gae_1d_nolp_misalign, gae_2d_nolp, gae_2d_olp  # noqa: F401
    from realhf._C.interval_op import merge_intervals
    from realhf._C.interval_op_cuda import (set_intervals_bf16, set_intervals_fp16, set_intervals_fp32,  # noqa: F401
                                            slice_intervals_bf16, slice_intervals_fp16, slice_intervals_fp32)
    assert merge_intervals([(0, 2), (2, 5), (7, 9)]) == [(0, 5), (7, 9)]
    assert set(mods) == {"cugae", "interval_op_cuda", "interval_op"}


def test_patch_areal_rebinds_from_import_sites():
    from areal_b200 import functional as F
    from areal_b200 import patch
    saved = {k: sys.modules.get(k) for k in ("areal", "areal.utils", "areal.utils.functional",
                                             "areal.engine", "areal.engine.ppo", "areal.engine.ppo.actor")}
    try:
        fun = types.ModuleType("areal.utils.functional")
        for n in ("gather_logprobs", "gather_logprobs_entropy", "ppo_actor_loss_fn", "ppo_critic_loss_fn"):
            setattr(fun, n, lambda *a, **k: "reference")
        actor = types.ModuleType("areal.engine.ppo.actor")
        actor.gather_logprobs_entropy = fun.gather_logprobs_entropy    # `from ... import ...` binding
        actor.ppo_actor_loss_fn = fun.ppo_actor_loss_fn
        class PPOActor:                                                 # stands for the reference class
            def compute_advantages(self, data):
                return "reference"
        actor.PPOActor = PPOActor
        sys.modules["areal.utils.functional"] = fun
        sys.modules["areal.engine.ppo.actor"] = actor
        patched = patch.patch_areal()
        from areal_b200 import actor as A
        assert PPOActor.compute_advantages is A.compute_advantages
        assert "areal.engine.ppo.actor.PPOActor.compute_advantages" in patched
        assert "areal.utils.functional" in patched and "areal.engine.ppo.actor" in patched
        assert actor.gather_logprobs_entropy is F.gather_logprobs_entropy
        assert actor.ppo_actor_loss_fn is F.ppo_actor_loss_fn
        assert fun.ppo_critic_loss_fn is F.ppo_critic_loss_fn
        assert not hasattr(actor, "ppo_critic_loss_fn")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def test_peer_exchange_needs_a_process_group():
    import pytest
    import torch.distributed as dist
    from areal_b200 import dist as D
    if dist.is_initialized():
        pytest.skip("a process group is already up in this interpreter")
    with pytest.raises(RuntimeError, match="initialised process group"):
        D.PeerExchange(max_tokens=16)
    with pytest.raises(RuntimeError, match="initialised process group"):
        import torch
        D.PeerPackedBuffer(numel=16, dtype=torch.bfloat16)


def test_parallel_group_lookup_without_a_process_group():
    """gather_logprobs' distributed branches stay off unless a group with more than one rank is set or found where
    the reference looks (megatron.core.parallel_state / areal.utils.ulysses, only if already imported)."""
    from areal_b200 import functional as F
    F.reset_parallel_groups()
    assert F._tp_group() is None and F._sp_group() is None
    F.set_parallel_groups(tp_group="not a real group", sp_group="neither")     # no process group: world of one
    assert F._tp_group() is None and F._sp_group() is None
    F.reset_parallel_groups()
    fake = types.ModuleType("megatron.core.parallel_state")
    fake.is_initialized = lambda: True
    fake.get_tensor_model_parallel_world_size = lambda: 4
    fake.get_tensor_model_parallel_group = lambda: "tp-group"
    saved = sys.modules.get("megatron.core.parallel_state")
    sys.modules["megatron.core.parallel_state"] = fake
    try:
        assert F._tp_group() == "tp-group"
        fake.get_tensor_model_parallel_world_size = lambda: 1
        assert F._tp_group() is None
    finally:
        if saved is None:
            del sys.modules["megatron.core.parallel_state"]
        else:
            sys.modules["megatron.core.parallel_state"] = saved


def test_lm_head_entry_points_refuse_cpu_tensors():
    """No CPU fallback on the LM-head path either: CPU tensors raise before any work is done."""
    import pytest
    import torch
    from areal_b200 import _lib
    from areal_b200.lm_head import fused_lm_head_actor_loss, lm_head_logprobs
    h = torch.zeros(4, 16, dtype=torch.bfloat16, requires_grad=True)
    w = torch.zeros(32, 16, dtype=torch.bfloat16, requires_grad=True)
    lab = torch.zeros(4, dtype=torch.long)
    e = torch.zeros(4)
    with pytest.raises(_lib.ArealB200Error, match="CUDA tensors only"):
        fused_lm_head_actor_loss(h, w, lab, 1.0, e, e, e, e > 0, eps_clip=0.2)
    with pytest.raises(_lib.ArealB200Error, match="CUDA tensors only"):
        lm_head_logprobs(h, w, lab)
