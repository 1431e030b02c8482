"""TEST INFRASTRUCTURE ONLY -- loads single files of the UNMODIFIED reference
(`/root/reference`) by file path, with `sys.modules` stubs for the imports this
image lacks (megatron, omegaconf, realhf.base ...).

Used only in the build container to (a) validate the restatement in
`oracle/cpu_oracle.py` and (b) generate the committed golden vectors under
`tests/golden/` (`oracle/gen_golden.py`).  `/root/reference` does not exist on
the GPU box, so nothing in `-m gpu` tests, `smoke()` or `bench.py` imports this.
No reference source is copied: the files are executed where they lie.
"""
import importlib.util
import os
import sys
import types

REF_ROOT = os.environ.get("AREAL_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "areal/utils/functional.py"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _load(path, modname):
    spec = importlib.util.spec_from_file_location(modname, os.path.join(REF_ROOT, path))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_realhf_ppo_functional():
    """realhf/impl/model/utils/ppo_functional.py (pygae1d_nolp_misalign :292-323,
    actor_loss_fn :51-149, critic_loss_fn :161-225)."""
    saved = {k: sys.modules.get(k) for k in ("realhf", "realhf.base", "realhf.base.pkg_version")}
    try:
        pv = _stub("realhf.base.pkg_version", is_available=lambda n: False)
        base = _stub("realhf.base", pkg_version=pv)
        _stub("realhf", base=base)
        return _load("realhf/impl/model/utils/ppo_functional.py", "_ref_ppo_functional")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def load_areal_functional():
    """areal/utils/functional.py, eager (is_npu_available=True skips torch.compile, :38-40)."""
    names = ["megatron", "megatron.core", "megatron.core.parallel_state",
             "megatron.core.tensor_parallel", "areal", "areal.platforms", "areal.utils",
             "areal.utils.mcore", "areal.utils.mcore.functional", "areal.utils.ulysses"]
    saved = {k: sys.modules.get(k) for k in names}
    try:
        mpu = _stub("megatron.core.parallel_state", is_initialized=lambda: False)
        tp = _stub("megatron.core.tensor_parallel")
        core = _stub("megatron.core", parallel_state=mpu, tensor_parallel=tp)
        _stub("megatron", core=core)
        _stub("areal")
        _stub("areal.platforms", is_npu_available=True)
        _stub("areal.utils")
        _stub("areal.utils.mcore")
        _stub("areal.utils.mcore.functional", _VocabParallelEntropy=object)
        _stub("areal.utils.ulysses",
              get_ulysses_sequence_parallel_group=lambda: None,
              get_ulysses_sequence_parallel_world_size=lambda: 1)
        return _load("areal/utils/functional.py", "_ref_areal_functional")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def load_realhf_flatten_param():
    """realhf/impl/model/nn/flatten_param.py (`_slice_intervals_py` :65-75,
    `_set_intervals_py` :78-90); its heavy package imports are stubbed."""
    import logging as _logging
    names = ["realhf", "realhf.api", "realhf.api.core", "realhf.api.core.model_api",
             "realhf.api.core.config", "realhf.base", "realhf.base.constants",
             "realhf.base.logging", "realhf.impl", "realhf.impl.model", "realhf.impl.model.nn",
             "realhf.impl.model.nn.real_llm_base", "realhf.impl.model.nn.real_llm_parallel",
             "realhf._C", "realhf._C.interval_op"]
    saved = {k: sys.modules.get(k) for k in names}
    try:
        for n in ("realhf", "realhf.api", "realhf.api.core", "realhf.impl", "realhf.impl.model",
                  "realhf.impl.model.nn", "realhf.base"):
            m = _stub(n)
            m.__path__ = []
        _stub("realhf.api.core.model_api", ReaLModelConfig=object, ReaLModelParamKeys=object)
        sys.modules["realhf.api.core"].model_api = sys.modules["realhf.api.core.model_api"]
        _stub("realhf.api.core.config", ModelName=object)
        _stub("realhf.base.constants", use_cuda=lambda: False)
        _stub("realhf.base.logging", getLogger=_logging.getLogger)
        sys.modules["realhf.base"].constants = sys.modules["realhf.base.constants"]
        sys.modules["realhf.base"].logging = sys.modules["realhf.base.logging"]
        _stub("realhf.impl.model.nn.real_llm_base", ReaLModelParamKeys=object)
        _stub("realhf.impl.model.nn.real_llm_parallel", get_real_model_param_shape=None,
              intervals_partition_fn=None, shape_partition_fn=None, tp_partition_key=None)
        spec = importlib.util.spec_from_file_location(
            "realhf.impl.model.nn.flatten_param",
            os.path.join(REF_ROOT, "realhf/impl/model/nn/flatten_param.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def run_areal_padded_gae(rewards, values, loss_mask, seq_no_eos_mask, discount, gae_lambda):
    """Executes the GAE loop of `PPOActor.compute_advantages` (areal/engine/ppo/actor.py,
    from 'advantages_reversed = [' to 'data["returns"] = advantages + values', :136-152)
    straight from the reference file -- the class itself cannot be imported here."""
    import textwrap
    import types as _t

    import torch
    with open(os.path.join(REF_ROOT, "areal/engine/ppo/actor.py")) as f:
        lines = f.read().splitlines()
    start = next(i for i, l in enumerate(lines) if "advantages_reversed = [" in l)
    end = next(i for i, l in enumerate(lines) if 'data["returns"] = advantages + values' in l)
    src = textwrap.dedent("\n".join(lines[start:end + 1]))
    bs, max_seqlen = rewards.shape
    ns = dict(torch=torch, rewards=rewards, values=values, loss_mask=loss_mask,
              seq_no_eos_mask=seq_no_eos_mask, bs=bs, max_seqlen=max_seqlen, data={},
              self=_t.SimpleNamespace(discount=discount, gae_lambda=gae_lambda))
    exec(compile(src, "actor.py[gae]", "exec"), ns)
    return ns["advantages"], ns["data"]["returns"]


def load_realhf_functional():
    """realhf/impl/model/utils/functional.py (`build_leave_one_indices` :129-166,
    `gather_packed_shifted_log_probs` :189-234) with its package imports stubbed."""
    import logging as _logging
    names = ["realhf", "realhf.base", "realhf.base.constants", "realhf.base.logging", "realhf.impl",
             "realhf.impl.model", "realhf.impl.model.utils", "realhf.impl.model.utils.padding"]
    saved = {k: sys.modules.get(k) for k in names}
    try:
        for n in ("realhf", "realhf.base", "realhf.impl", "realhf.impl.model", "realhf.impl.model.utils"):
            m = _stub(n)
            m.__path__ = []
        _stub("realhf.base.constants", tensor_parallel_world_size=lambda: 1)
        _stub("realhf.base.logging", getLogger=_logging.getLogger)
        sys.modules["realhf.base"].constants = sys.modules["realhf.base.constants"]
        sys.modules["realhf.base"].logging = sys.modules["realhf.base.logging"]
        _stub("realhf.impl.model.utils.padding", pad_input=None, unpad_input=None)
        return _load("realhf/impl/model/utils/functional.py", "_ref_realhf_functional")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def run_realhf_calc_entropy(logits, cu_seqlens):
    """Executes the body of `calc_entropy` (realhf/impl/model/interface/ppo_interface.py:52-58,
    without its torch.compile decorator) straight from the reference file."""
    import textwrap

    import torch
    rf = load_realhf_functional()
    with open(os.path.join(REF_ROOT, "realhf/impl/model/interface/ppo_interface.py")) as f:
        lines = f.read().splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith("def calc_entropy("))
    end = next(i for i in range(start + 1, len(lines)) if lines[i].startswith("    return entropy"))
    src = textwrap.dedent("\n".join(lines[start:end + 1]))
    ns = dict(torch=torch, build_leave_one_indices=rf.build_leave_one_indices)
    exec(compile(src, "ppo_interface.py[calc_entropy]", "exec"), ns)
    with torch.no_grad():
        return ns["calc_entropy"](logits, cu_seqlens)


def load_areal_normalization():
    """The `Normalization` class of areal/utils/data.py (:1073-1304), extracted by AST and executed
    where it lies (the module itself needs tensordict / einops-heavy imports).  Returns (cls, NormConfig)
    where NormConfig is a stand-in dataclass with the reference's field names and defaults
    (areal/api/cli_args.py:24-59)."""
    import ast
    import dataclasses
    from typing import Optional

    import numpy as np
    import torch
    import torch.distributed as dist
    path = os.path.join(REF_ROOT, "areal/utils/data.py")
    with open(path) as f:
        src = f.read()
    tree = ast.parse(src)
    node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "Normalization")
    code = ast.get_source_segment(src, node)

    @dataclasses.dataclass
    class NormConfig:
        mean_level: Optional[str] = "batch"
        mean_leave1out: bool = False
        std_level: Optional[str] = "batch"
        std_unbiased: bool = True
        eps: float = 1e-5
        group_size: int = 1

    ns = dict(torch=torch, np=np, dist=dist, Optional=Optional, NormConfig=NormConfig)
    exec(compile(code, "data.py[Normalization]", "exec"), ns)
    return ns["Normalization"], NormConfig


def load_areal_pack_tensor_dict():
    """`pack_tensor_dict` (areal/utils/data.py:266-315), extracted by AST and executed where it lies."""
    import ast
    from typing import Any, Dict

    import torch
    import torch.nn.functional as F
    path = os.path.join(REF_ROOT, "areal/utils/data.py")
    with open(path) as f:
        src = f.read()
    tree = ast.parse(src)
    node = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "pack_tensor_dict")
    ns = dict(torch=torch, F=F, Dict=Dict, Any=Any)
    exec(compile(ast.get_source_segment(src, node), "data.py[pack_tensor_dict]", "exec"), ns)
    return ns["pack_tensor_dict"]


def load_areal_compute_advantages():
    """`PPOActor.compute_advantages` (areal/engine/ppo/actor.py:72-164) as a plain function of
    (self, data), extracted by AST and executed where it lies together with `KLEstimator`
    (areal/utils/data.py:1306-1364), `reward_overlong_penalty` (areal/utils/functional.py:376-398) and
    the reference `Normalization`.  Returns (compute_advantages, KLEstimator, Normalization, NormConfig)."""
    import ast
    from typing import Any, Dict, List, Optional

    import torch
    Normalization, NormConfig = load_areal_normalization()

    def segment(path, kind, name, parent=None):
        with open(os.path.join(REF_ROOT, path)) as f:
            src = f.read()
        body = ast.parse(src).body
        if parent is not None:
            body = next(n for n in body if isinstance(n, ast.ClassDef) and n.name == parent).body
        node = next(n for n in body if isinstance(n, kind) and n.name == name)
        node.decorator_list = []
        return ast.unparse(node)

    ns = dict(torch=torch, Dict=Dict, Any=Any, List=List, Optional=Optional, Normalization=Normalization)
    exec(compile(segment("areal/utils/data.py", ast.ClassDef, "KLEstimator"), "data.py[KLEstimator]", "exec"), ns)
    exec(compile(segment("areal/utils/functional.py", ast.FunctionDef, "reward_overlong_penalty"),
                 "functional.py[reward_overlong_penalty]", "exec"), ns)
    exec(compile(segment("areal/engine/ppo/actor.py", ast.FunctionDef, "compute_advantages", parent="PPOActor"),
                 "actor.py[compute_advantages]", "exec"), ns)
    return ns["compute_advantages"], ns["KLEstimator"], Normalization, NormConfig


def load_areal_stats_tracker():
    """areal/utils/stats_tracker.py executed where it lies (stubs: areal.platforms.current_platform,
    areal.utils.logging, areal.utils.datapack.flat2d)."""
    import itertools
    import logging as _logging
    names = ["areal", "areal.platforms", "areal.utils", "areal.utils.logging", "areal.utils.datapack"]
    saved = {k: sys.modules.get(k) for k in names}
    try:
        _stub("areal")
        _stub("areal.platforms", current_platform=types.SimpleNamespace(is_initialized=lambda: False, device_type="cpu"))
        lg = _stub("areal.utils.logging", getLogger=lambda name=None: _logging.getLogger(name))
        dp = _stub("areal.utils.datapack", flat2d=lambda arr: list(itertools.chain(*arr)))
        _stub("areal.utils", logging=lg, datapack=dp)
        return _load("areal/utils/stats_tracker.py", "_ref_areal_stats_tracker")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def load_areal_grpo_loss_fn():
    """`grpo_loss_fn` (areal/engine/ppo/actor.py:313-391) bound to the reference's own
    gather_logprobs_entropy / ppo_actor_loss_fn and stats_tracker.  Returns (fn, stats_tracker module)."""
    import ast
    from typing import Dict

    import torch
    af = load_areal_functional()
    st = load_areal_stats_tracker()
    with open(os.path.join(REF_ROOT, "areal/engine/ppo/actor.py")) as f:
        src = f.read()
    node = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "grpo_loss_fn")
    ns = dict(torch=torch, Dict=Dict, gather_logprobs_entropy=af.gather_logprobs_entropy,
              ppo_actor_loss_fn=af.ppo_actor_loss_fn, stats_tracker=st)
    exec(compile(ast.get_source_segment(src, node), "actor.py[grpo_loss_fn]", "exec"), ns)
    return ns["grpo_loss_fn"], st


def load_realhf_rms():
    """realhf/impl/model/modules/rms.py executed where it lies (stubs: realhf.base.constants with
    current_device() = "cpu", realhf.base.logging)."""
    import logging as _logging
    names = ["realhf", "realhf.base", "realhf.base.constants", "realhf.base.logging"]
    saved = {k: sys.modules.get(k) for k in names}
    try:
        const = _stub("realhf.base.constants", current_device=lambda: "cpu", data_parallel_group=lambda: None)
        lg = _stub("realhf.base.logging", getLogger=lambda name=None: _logging.getLogger(name))
        base = _stub("realhf.base", constants=const, logging=lg)
        _stub("realhf", base=base)
        return _load("realhf/impl/model/modules/rms.py", "_ref_realhf_rms")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def load_aent_functional():
    """`clamped_softmax_entropy` / `_gather_logprobs_clamped_entropy` (recipe/AEnt/functional.py:15-52),
    extracted by AST with the `@torch.compile` decorators dropped and executed where they lie.
    Returns (clamped_softmax_entropy, gather_fn(logits, labels, entropy_clamp, temperature))."""
    import ast

    import torch
    import torch.nn.functional as F
    with open(os.path.join(REF_ROOT, "recipe/AEnt/functional.py")) as f:
        src = f.read()
    ns = dict(torch=torch, F=F)
    for name in ("clamped_softmax_entropy", "_gather_logprobs_clamped_entropy"):
        node = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == name)
        node.decorator_list = []
        exec(compile(ast.unparse(node), f"AEnt/functional.py[{name}]", "exec"), ns)
    return ns["clamped_softmax_entropy"], ns["_gather_logprobs_clamped_entropy"]


def load_aent_grpo_loss_fn():
    """`aent_grpo_loss_fn` + `clamped_entropy_loss_fn` (recipe/AEnt/actor.py:185-278) bound to the reference's own
    `gather_logprobs_clamped_entropy` (recipe/AEnt/functional.py:54-86, Ulysses world size stubbed to 1),
    areal functional and stats_tracker.  Returns (fn, stats_tracker module)."""
    import ast
    from typing import Dict

    import torch
    import torch.nn.functional as F
    af = load_areal_functional()
    st = load_areal_stats_tracker()
    _, inner = load_aent_functional()
    ns = dict(torch=torch, F=F, Dict=Dict, _gather_logprobs_clamped_entropy=inner,
              get_ulysses_sequence_parallel_world_size=lambda: 1, get_ulysses_sequence_parallel_group=lambda: None,
              gather_logprobs_entropy=af.gather_logprobs_entropy, ppo_actor_loss_fn=af.ppo_actor_loss_fn,
              stats_tracker=st)
    for path, name in (("recipe/AEnt/functional.py", "gather_logprobs_clamped_entropy"),
                       ("recipe/AEnt/actor.py", "clamped_entropy_loss_fn"), ("recipe/AEnt/actor.py", "aent_grpo_loss_fn")):
        with open(os.path.join(REF_ROOT, path)) as f:
            src = f.read()
        node = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == name)
        node.decorator_list = []
        exec(compile(ast.unparse(node), f"AEnt[{name}]", "exec"), ns)
    return ns["aent_grpo_loss_fn"], st


def load_areal_critic_loss_fn():
    """`ppo_loss_fn` (areal/engine/ppo/critic.py:81-127) bound to the reference's ppo_critic_loss_fn and
    stats_tracker.  Returns (fn, stats_tracker module)."""
    import ast
    from typing import Dict

    import torch
    af = load_areal_functional()
    st = load_areal_stats_tracker()
    with open(os.path.join(REF_ROOT, "areal/engine/ppo/critic.py")) as f:
        src = f.read()
    node = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "ppo_loss_fn")
    ns = dict(torch=torch, Dict=Dict, ppo_critic_loss_fn=af.ppo_critic_loss_fn, stats_tracker=st)
    exec(compile(ast.get_source_segment(src, node), "critic.py[ppo_loss_fn]", "exec"), ns)
    return ns["ppo_loss_fn"], st


def load_areal_sft_loss_fn():
    """`compute_packed_sft_loss` (areal/engine/sft/lm_engine.py:58-110) bound to the reference's gather_logprobs
    and stats_tracker.  Returns (fn, stats_tracker module)."""
    import ast
    from typing import Any, Dict

    import torch
    af = load_areal_functional()
    st = load_areal_stats_tracker()
    with open(os.path.join(REF_ROOT, "areal/engine/sft/lm_engine.py")) as f:
        src = f.read()
    node = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "compute_packed_sft_loss")
    ns = dict(torch=torch, Dict=Dict, Any=Any, gather_logprobs=af.gather_logprobs, stats_tracker=st)
    exec(compile(ast.get_source_segment(src, node), "lm_engine.py[compute_packed_sft_loss]", "exec"), ns)
    return ns["compute_packed_sft_loss"], st


def load_vocab_parallel(tp_group, tp_rank: int, tp_world: int):
    """The reference's two vocabulary-sharded ops, executed where they lie, for a torch.distributed group
    (gloo on CPU in the build container):
      * `_VocabParallelEntropy` (areal/utils/mcore/functional.py:8-52) -- what gather_logprobs_entropy calls under
        Megatron tensor parallelism (areal/utils/functional.py:91-98); its local `torch.compile` helper runs
        eagerly here (torch.compile is made the identity while the file loads and runs);
      * `vocab_parallel_cross_entropy` (realhf/impl/model/parallelism/tensor_parallel/modules.py:1060-1186), the
        copy of Megatron's `tensor_parallel.vocab_parallel_cross_entropy` that the reference tree carries
        (megatron-core itself, pyproject.toml:49, is not in the tree).
    Returns (VocabParallelEntropy, vocab_parallel_cross_entropy)."""
    import logging as pylogging
    import torch
    tp_dir = os.path.join(REF_ROOT, "realhf/impl/model/parallelism/tensor_parallel")
    mpu = types.ModuleType("megatron.core.parallel_state")
    mpu.get_tensor_model_parallel_group = lambda: tp_group
    mpu.is_initialized = lambda: True
    mpu.get_tensor_model_parallel_world_size = lambda: tp_world
    core = types.ModuleType("megatron.core")
    core.parallel_state = mpu
    constants = types.ModuleType("realhf.base.constants")
    constants.tensor_parallel_group = lambda: tp_group
    constants.tensor_parallel_rank = lambda: tp_rank
    constants.tensor_parallel_world_size = lambda: tp_world
    rlog = types.ModuleType("realhf.base.logging")
    rlog.getLogger = lambda *a, **k: pylogging.getLogger("ref")
    base = types.ModuleType("realhf.base")
    base.constants, base.logging = constants, rlog
    rnd = types.ModuleType("realhf.impl.model.utils.random")
    rnd._initialize_affine_weight_gpu = lambda *a, **k: None
    tp_pkg = types.ModuleType("realhf.impl.model.parallelism.tensor_parallel")
    tp_pkg.__path__ = [tp_dir]
    stubs = {"megatron": types.ModuleType("megatron"), "megatron.core": core, "megatron.core.parallel_state": mpu,
             "realhf": types.ModuleType("realhf"), "realhf.base": base, "realhf.base.constants": constants,
             "realhf.base.logging": rlog, "realhf.impl": types.ModuleType("realhf.impl"),
             "realhf.impl.model": types.ModuleType("realhf.impl.model"),
             "realhf.impl.model.utils": types.ModuleType("realhf.impl.model.utils"),
             "realhf.impl.model.utils.random": rnd,
             "realhf.impl.model.parallelism": types.ModuleType("realhf.impl.model.parallelism"),
             "realhf.impl.model.parallelism.tensor_parallel": tp_pkg}
    for name in ("realhf", "realhf.base", "realhf.impl", "realhf.impl.model", "realhf.impl.model.utils",
                 "realhf.impl.model.parallelism", "megatron"):
        stubs[name].__path__ = []
    saved = {k: sys.modules.get(k) for k in stubs}
    real_compile = torch.compile
    try:
        sys.modules.update(stubs)
        torch.compile = lambda *a, **k: (a[0] if a and callable(a[0]) else (lambda f: f))
        ent_mod = _load("areal/utils/mcore/functional.py", "_ref_areal_mcore_functional")
        ce_mod = _load("realhf/impl/model/parallelism/tensor_parallel/modules.py",
                       "realhf.impl.model.parallelism.tensor_parallel.modules")
    finally:
        # torch.compile stays patched for the caller's run of _VocabParallelEntropy.forward (it calls
        # torch.compile at call time); restore_compile() below puts it back
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    ent_mod._restore_compile = lambda: setattr(torch, "compile", real_compile)
    return ent_mod._VocabParallelEntropy, ce_mod.vocab_parallel_cross_entropy, ent_mod._restore_compile


def load_fsdp_train_batch():
    """`FSDPEngine.train_batch` (areal/engine/fsdp_engine.py:498-606) as a plain function `train_batch(self,
    input_, loss_fn, loss_weight_fn)`: the method's source is cut out of the reference file by its AST node
    and compiled where it lies (the module itself needs peft / torchdata / megatron to import).  The
    globals it touches are the real `torch` / `dist` / `math`; the gradient-norm helper is replaced by
    torch's own clip_grad_norm_ and the Ulysses helpers are never reached with sp_size == 1."""
    import ast
    import math
    from typing import Any, Callable, Dict, List, Optional, Tuple

    import torch
    import torch.distributed as dist
    path = os.path.join(REF_ROOT, "areal/engine/fsdp_engine.py")
    src = open(path).read()
    tree = ast.parse(src)
    node = None
    for cls in tree.body:
        if isinstance(cls, ast.ClassDef) and cls.name == "FSDPEngine":
            for fn in cls.body:
                if isinstance(fn, ast.FunctionDef) and fn.name == "train_batch":
                    node = fn
    assert node is not None, "FSDPEngine.train_batch not found"
    node.decorator_list = []
    mod = ast.Module(body=[node], type_ignores=[])
    ast.fix_missing_locations(mod)

    def fsdp2_clip_grad_norm(params, mesh, max_norm):
        return float(torch.nn.utils.clip_grad_norm_([p for p in params if p.grad is not None] or params, max_norm))

    def unreachable(*a, **k):
        raise AssertionError("Ulysses helper reached with sp_size == 1")
    g = dict(torch=torch, dist=dist, math=math, Any=Any, Callable=Callable, Dict=Dict, List=List, Optional=Optional,
             Tuple=Tuple, fsdp2_clip_grad_norm=fsdp2_clip_grad_norm, set_ulysses_sequence_parallel_group=unreachable,
             ulysses_pad=unreachable, ulysses_pad_and_slice_inputs=unreachable, ulysses_prepare_inputs=unreachable)
    exec(compile(mod, path, "exec"), g)
    return g["train_batch"]
