"""GPU: the whole step (packed GAE, fused single-pass loss + dlogits, compute_advantages, stats record)
is capturable in a CUDA graph -- no hidden host sync, no allocation outside torch's graph pool -- and a
replay on new input CONTENTS gives what the eager calls give."""
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
D = torch.device("cuda:0")


def test_step_replays_from_a_cuda_graph():
    from areal_b200 import functional as F
    from areal_b200 import ppo_functional as P
    from areal_b200 import stats_tracker as ST
    from areal_b200.actor import compute_advantages
    g = torch.Generator(device=D).manual_seed(0)
    T, V, B, S = 512, 32768, 24, 264
    lens = torch.randint(4, 40, (B,), generator=g, device=D)
    lens[0] = T - int(lens[1:].sum())
    cu = torch.zeros(B + 1, dtype=torch.int32, device=D)
    cu[1:] = lens.cumsum(0)
    static = dict(logits=torch.empty(T, V, dtype=torch.bfloat16, device=D), labels=torch.zeros(T, dtype=torch.long, device=D),
                  old=torch.zeros(T, device=D), prox=torch.zeros(T, device=D), adv=torch.zeros(T, device=D),
                  mask=torch.zeros(T, dtype=torch.bool, device=D), rewards=torch.zeros(T, device=D),
                  values=torch.zeros(T + B, device=D), boot=torch.zeros(B, dtype=torch.bool, device=D),
                  p_lm=torch.zeros(B, S, dtype=torch.bool, device=D), p_am=torch.ones(B, S, dtype=torch.bool, device=D),
                  p_lp=torch.zeros(B, S, device=D), p_score=torch.zeros(B, device=D))
    scale = torch.full((1,), 0.25, device=D)
    actor = types.SimpleNamespace(
        config=types.SimpleNamespace(use_decoupled_loss=True, recompute_logprob=True, overlong_reward_penalty=False),
        reward_bias=0.0, reward_scaling=1.0, reward_clip=5.0, kl_ctl=0.1,
        kl_estimator=types.SimpleNamespace(kl_estimator="k2", apply_clamp=True), discount=0.99, gae_lambda=0.95,
        mask_no_eos_with_zero=False, adv_norm=None, reward_norm=None)

    def fill(seed):
        gg = torch.Generator(device=D).manual_seed(seed)
        static["logits"].copy_((torch.randn(T, V, generator=gg, device=D) * 2).bfloat16())
        static["labels"].copy_(torch.randint(0, V, (T,), generator=gg, device=D))
        static["old"].copy_(-torch.rand(T, generator=gg, device=D) * 12 - 5)
        static["prox"].copy_(static["old"] + 0.1 * torch.randn(T, generator=gg, device=D))
        static["adv"].copy_(torch.randn(T, generator=gg, device=D))
        static["mask"].copy_(torch.rand(T, generator=gg, device=D) > 0.3)
        static["rewards"].copy_(torch.randn(T, generator=gg, device=D))
        static["values"].copy_(torch.randn(T + B, generator=gg, device=D))
        static["boot"].copy_(torch.rand(B, generator=gg, device=D) > 0.5)
        static["p_lm"].copy_(torch.rand(B, S, generator=gg, device=D) > 0.3)
        static["p_lp"].copy_(-torch.rand(B, S, generator=gg, device=D))
        static["p_score"].copy_(torch.randn(B, generator=gg, device=D))

    tracker = ST.DistributedStatsTracker()

    def step():
        x = static["logits"].detach().requires_grad_()
        a, r = P.cugae1d_nolp_misalign_func(static["rewards"], static["values"], cu, static["boot"], 0.99, 0.95)
        loss, st = F.fused_actor_loss(x, static["labels"], 1.0, static["prox"], static["old"], static["adv"],
                                      static["mask"], eps_clip=0.2, behav_imp_weight_cap=5.0, loss_scale=scale)
        (loss * scale[0]).backward()
        data = dict(input_ids=static["p_lp"], attention_mask=static["p_am"], loss_mask=static["p_lm"],
                    logprobs=static["p_lp"], rewards=static["p_score"])
        compute_advantages(actor, data)
        tracker.record(dict(valid=static["mask"]), [("valid", dict(entropy=st["entropy"], logp=st["logprobs"]))])
        return [a, r, loss.detach(), st["logprobs"], st["entropy"], x.grad, data["advantages"], data["tot_rewards"]]

    fill(1)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):                     # warm-up off the capture: workspaces, attributes, slots
        for _ in range(2):
            step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    tracker.export()
    # the tracker's slots must exist before capture (they are created on first use)
    tracker.record(dict(valid=static["mask"]), [("valid", dict(entropy=static["old"], logp=static["old"]))])
    tracker.export(reset=False)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        outs = step()
    for seed in (2, 3):
        fill(seed)
        graph.replay()
        torch.cuda.synchronize()
        got = [o.clone() for o in outs]
        want = step()
        torch.cuda.synchronize()
        for i, (a, b) in enumerate(zip(got, want)):
            assert torch.equal(a.view(torch.int16) if a.dtype == torch.bfloat16 else a,
                               b.view(torch.int16) if b.dtype == torch.bfloat16 else b), (seed, i)
