"""Stress: single-pass fused actor loss (row-resident kernel) and the two-pass fused path repeated with
poisoned free memory; every iteration must reproduce the first bit for bit."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from areal_b200 import functional as F
D = torch.device("cuda:0")
T, V = 16384, 151936
g = torch.Generator(device=D).manual_seed(0)
logits = torch.empty(T, V, dtype=torch.bfloat16, device=D)
for i in range(0, T, 2048):
    logits[i:i + 2048] = (torch.randn(2048, V, generator=g, device=D) * 2.0).bfloat16()
labels = torch.randint(0, V, (T,), generator=g, device=D)
old = -torch.rand(T, generator=g, device=D) * 12 - 6
prox = old + 0.1 * torch.randn(T, generator=g, device=D)
adv = torch.randn(T, generator=g, device=D)
mask = torch.rand(T, generator=g, device=D) > 0.2
scale = torch.full((1,), 0.5, device=D)
for mode in ("single", "twopass"):
    ref = None
    bad = 0
    for it in range(int(sys.argv[1]) if len(sys.argv) > 1 else 10):
        junk = torch.empty(T, V, dtype=torch.bfloat16, device=D)
        junk.view(torch.int16).fill_(0x7FC1 if it % 2 else 0x4142)
        small = [torch.full((T,), float("nan"), device=D) for _ in range(12)]
        del junk, small
        x = logits.detach().requires_grad_()
        loss, st = F.fused_actor_loss(x, labels, 1.0, prox, old, adv, mask, eps_clip=0.4, behav_imp_weight_cap=5.0,
                                      loss_scale=scale if mode == "single" else None)
        (loss * scale[0]).backward()
        torch.cuda.synchronize()
        cur = [loss.detach().reshape(1).clone(), st["logprobs"].clone(), st["entropy"].clone(), x.grad]
        if ref is None:
            ref = [c.clone() for c in cur]
            print(mode, "iter 0 loss", float(loss), "nan in grad:", int(torch.isnan(x.grad.float()).sum()), flush=True)
        else:
            same = [bool(torch.equal(a.view(torch.int16) if a.dtype == torch.bfloat16 else a.view(torch.int32),
                                     b.view(torch.int16) if b.dtype == torch.bfloat16 else b.view(torch.int32)))
                    for a, b in zip(cur, ref)]
            if not all(same):
                bad += 1
                print(mode, "iter", it, "MISMATCH", same, flush=True)
        del x, cur
    print(mode, "mismatching iterations:", bad, flush=True)
