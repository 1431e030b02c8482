"""Times fused_actor_loss(single pass) for a list of (cluster_size, n_buffers) on the bench shape."""
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
cfgs = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]] or [(4, 3), (8, 3), (8, 6), (8, 5), (8, 4), (6, 4)]
for C, nb in cfgs:
    ts = []
    try:
        for it in range(14):
            x = logits.detach().requires_grad_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            loss, st = F.fused_actor_loss(x, labels, 1.0, prox, old, adv, mask, eps_clip=0.4, behav_imp_weight_cap=5.0,
                                          loss_scale=scale, cluster_size=C, n_buffers=nb)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
            del x, loss, st
        ts = sorted(ts[4:])
        print(f"C={C} nbuf={nb}: median {ts[len(ts)//2]:.4f} ms  min {ts[0]:.4f}", flush=True)
    except Exception as e:
        print(f"C={C} nbuf={nb}: {str(e)[:150]}", flush=True)
