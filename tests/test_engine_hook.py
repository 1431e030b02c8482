"""The loss-scale hook (areal_b200/engine_hooks.py) against the reference engine's own arithmetic.

CPU: the unmodified `FSDPEngine.train_batch` (areal/engine/fsdp_engine.py:498-606, compiled from the
reference file where it lies, oracle/ref_loader.py:load_fsdp_train_batch) runs on a stand-in engine object
with the hook installed; the scale announced to loss_fn BEFORE each forward must be bit-equal to the
scalar the engine multiplies the loss by AFTER it (captured as the upstream gradient), for one rank and
for two gloo ranks with unequal micro-batches.  GPU: the same loop drives areal_b200.actor.grpo_loss_fn:
gradients equal the two-pass path's and the announced scale is the one that arrives (so the rescale
kernel stays a no-op)."""
import os
import types

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ref_loader as R

needs_ref = pytest.mark.skipif(not R.available(), reason="needs /root/reference (build container)")


class _Capture(torch.autograd.Function):
    """identity on the loss; records the upstream gradient = the engine's loss_scale"""

    @staticmethod
    def forward(ctx, x, sink):
        ctx.sink = sink
        return x.clone()

    @staticmethod
    def backward(ctx, g):
        ctx.sink.append(g.detach().clone())
        return g, None


class _MbList:
    def __init__(self, mbs):
        self.mbs = mbs
        self.padded_mbs = mbs
        self.padding_lengths = [0] * len(mbs)

    def to(self, device):
        return self


def make_engine(train_batch, mbs, weight, dp_group, dp_size, device):
    """The attributes FSDPEngine.train_batch touches (fsdp_engine.py:498-606), nothing else."""
    eng = types.SimpleNamespace()
    eng.optimizer = types.SimpleNamespace(zero_grad=lambda: None, step=lambda: None)
    eng.optimizer_config = types.SimpleNamespace(gradient_clipping=1e9)
    eng.lr_scheduler = types.SimpleNamespace(step=lambda: None, get_last_lr=lambda: [1e-5])
    eng.parallel_helper = types.SimpleNamespace(sp_size=1, dp_size=dp_size)
    eng.is_vision_model = False
    eng.dp_group, eng.sp_group, eng.device, eng.world_mesh = dp_group, None, device, None
    eng.prepare_mb_list = lambda input_: _MbList(mbs)
    eng.model = types.SimpleNamespace(parameters=lambda: [weight],
                                      __call__=None)
    eng.model = _Model(weight)
    eng.train_batch = types.MethodType(train_batch, eng)
    return eng


class _Model:
    def __init__(self, w):
        self.w = w

    def parameters(self):
        return [self.w]

    def __call__(self, **inputs):
        logits = inputs["hidden"] @ self.w            # [T, V]
        return types.SimpleNamespace(logits=logits.unsqueeze(0))


def _run_cpu(rank, world, port, out):
    if world > 1:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from areal_b200 import engine_hooks as EH
        train_batch = R.load_fsdp_train_batch()
        g = torch.Generator().manual_seed(10 + rank)
        w = torch.randn(8, 32, generator=g).requires_grad_()
        mbs = []
        for n in ((5, 9, 14) if rank == 0 else (21, 3)):            # unequal micro-batch counts and sizes
            mbs.append(dict(hidden=torch.randn(n, 8, generator=g), loss_mask=torch.rand(n, generator=g) > 0.4,
                            input_ids=torch.zeros(n, dtype=torch.long)))
            mbs[-1]["loss_mask"][0] = True
        eng = make_engine(train_batch, mbs, w, dist.group.WORLD if world > 1 else None, world, torch.device("cpu"))
        if world == 1:
            dist_all_reduce = dist.all_reduce
            dist.all_reduce = lambda t, group=None: None            # the engine all-reduces unconditionally
        EH.install_loss_scale_hook(eng)
        announced, arrived = [], []

        def loss_fn(logits, mb):
            announced.append(EH.announced_loss_scale().clone())
            loss = (logits.float().logsumexp(-1) * mb["loss_mask"]).sum() / mb["loss_mask"].count_nonzero()
            return _Capture.apply(loss, arrived)
        try:
            stat = eng.train_batch({}, loss_fn=loss_fn, loss_weight_fn=lambda x: x["loss_mask"].count_nonzero())
        finally:
            if world == 1:
                dist.all_reduce = dist_all_reduce
        assert stat["update_successful"] == 1.0
        assert len(announced) == len(arrived) == len(mbs)
        # the announcer read the engine's own all-reduced total (no second all-reduce of its own)
        assert eng.train_batch.last_announcer.used_engine_total
        for a, b in zip(announced, arrived):
            assert a.dtype == torch.float32 and a.shape == (1,)
            assert torch.equal(a.reshape(()), b.reshape(())), (a, b)       # bit-equal to the engine's own scalar
        assert EH.announced_loss_scale() is None                             # nothing leaks out of the call
        # the scales of all micro-batches of all ranks sum to dp_size (a masked mean over the global batch)
        tot = torch.stack(announced).sum()
        if world > 1:
            dist.all_reduce(tot)
        assert abs(float(tot) - world) < 1e-5, float(tot)
        out.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        out.put((rank, traceback.format_exc()))
    finally:
        if world > 1:
            dist.destroy_process_group()


@needs_ref
def test_announced_scale_is_the_engines_scale_one_rank():
    import queue
    q = queue.Queue()
    _run_cpu(0, 1, 0, q)
    assert q.get()[1] == "ok"


@needs_ref
def test_announced_scale_is_the_engines_scale_two_gloo_ranks():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_run_cpu, args=(r, 2, 29561, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", f"rank {rank}: {msg}"


def test_hook_is_idempotent_and_transparent_without_weights():
    from areal_b200 import engine_hooks as EH
    calls = []

    class Eng:
        def train_batch(self, input_, loss_fn, loss_weight_fn):
            calls.append(loss_fn(torch.zeros(2, 3), {"x": 1}))       # an engine that never asks for weights
            return {"ok": 1.0}
    e = Eng()
    EH.install_loss_scale_hook(e)
    first = e.train_batch
    EH.install_loss_scale_hook(e)
    assert e.train_batch is first
    seen = []
    out = e.train_batch({}, lambda lg, mb: seen.append(EH.announced_loss_scale()) or 7, lambda mb: torch.tensor(1))
    assert out == {"ok": 1.0} and calls == [7] and seen == [None]


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_hooked_engine_loop_reaches_the_single_pass_kernel():
    """The reference's train_batch arithmetic (restated: the reference file is not on the GPU box) around
    areal_b200.actor.grpo_loss_fn with the hook installed: same loss and gradient as without the hook, the
    announced scale equals the upstream scalar bit for bit, and the single-pass entry is what ran."""
    import functools
    from areal_b200 import actor as A, engine_hooks as EH, functional as F, stats_tracker as ST
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(3)
    V = 4096
    mbs = []
    for n in (300, 77, 512):
        lg = (torch.randn(n, V, generator=g) * 2).bfloat16().to(dev)
        lab = torch.randint(0, V, (n,), generator=g).to(dev)
        lp = F.gather_logprobs(lg, lab, 1.0)
        mbs.append(dict(logits=lg, input_ids=torch.roll(lab, 1), rolled_input_ids=lab,
                        logprobs=lp + 0.05 * torch.randn(n, generator=g).to(dev),
                        prox_logp=lp + 0.02 * torch.randn(n, generator=g).to(dev),
                        advantages=torch.randn(n, generator=g).to(dev),
                        loss_mask=(torch.rand(n, generator=g) > 0.3).to(dev)))

    class Engine:                                       # fsdp_engine.py:518-526, 571-586 for pre-computed logits
        device = dev
        dp_group = None
        parallel_helper = types.SimpleNamespace(dp_size=1, sp_size=1)

        def train_batch(self, input_, loss_fn, loss_weight_fn):
            total = torch.stack([loss_weight_fn(mb) for mb in input_]).sum().detach().clone().to(
                dtype=torch.float32, device=self.device)
            self.scales = []
            for mb in input_:
                loss = loss_fn(mb["x"], mb)
                loss_scale = loss_weight_fn(mb) / total
                loss_scale *= self.parallel_helper.dp_size
                self.scales.append(loss_scale.clone())
                loss *= loss_scale
                loss.backward()
            return {}

    def run(hook):
        tracker = ST.DistributedStatsTracker()
        eng = Engine()
        if hook:
            EH.install_loss_scale_hook(eng)
        batch = [dict(mb, x=mb["logits"].clone().requires_grad_()) for mb in mbs]
        taken = []
        real = F._FusedActorRowRes.apply

        def spy(*a):
            taken.append(1)
            return real(*a)
        F._FusedActorRowRes.apply = spy
        try:
            eng.train_batch(batch, functools.partial(A.grpo_loss_fn, temperature=1.0, eps_clip=0.2, eps_clip_higher=None,
                                                     c_clip=None, behav_imp_weight_cap=5.0, tracker=tracker),
                            lambda x: x["loss_mask"].count_nonzero())
        finally:
            F._FusedActorRowRes.apply = real
        return eng, [b["x"].grad for b in batch], tracker.export(), len(taken)

    e0, g0, s0, n0 = run(False)
    e1, g1, s1, n1 = run(True)
    assert n0 == 0 and n1 == len(mbs)                   # the hook, and only the hook, reaches the single pass
    ann = e1.train_batch.last_announcer
    assert ann.n_announced == len(mbs)
    assert torch.equal(ann.last_announced.reshape(()), e1.scales[-1].reshape(()).float())
    for a, b in zip(g0, g1):
        assert torch.allclose(a.float(), b.float(), rtol=2 ** -7, atol=1e-10)
    assert set(s0) == set(s1)
    for k in s0:
        assert abs(s0[k] - s1[k]) <= 2e-5 * max(1.0, abs(s0[k])), k
