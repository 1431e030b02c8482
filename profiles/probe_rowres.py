"""Probe: single-pass row-resident fused step vs the two-pass step at the cfg-2 shape."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import functional as F
import bench
dev = torch.device("cuda:0")
T, V = 16384, 151936
b = bench.make_batch(torch, dev, T, V, 0)
lp = F.gather_logprobs(b["logits"], b["labels"], 1.0)
prox = lp + 0.02 * torch.randn_like(lp); old = lp + 0.05 * torch.randn_like(lp)
x = b["logits"].requires_grad_()
scale = torch.full((1,), 0.5, device=dev)
def timeloop(fn, n=20, w=4):
    for _ in range(w): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
def two_pass():
    x.grad = None
    l, _ = F.fused_actor_loss(x, b["labels"], 1.0, prox, old, b["advantages"], b["loss_mask"], eps_clip=0.4, behav_imp_weight_cap=5.0)
    (l * scale).backward()
print("two-pass step ms:", round(timeloop(two_pass), 4))
for (c, nb) in ((3, 2), (4, 2), (4, 3), (5, 3), (6, 3), (8, 3)):
    def one_pass():
        x.grad = None
        l, _ = F.fused_actor_loss(x, b["labels"], 1.0, prox, old, b["advantages"], b["loss_mask"], eps_clip=0.4,
                                  behav_imp_weight_cap=5.0, loss_scale=scale, cluster_size=c, n_buffers=nb)
        (l * scale).backward()
    try:
        ms = timeloop(one_pass)
        print(f"single-pass cluster={c} nbuf={nb}: {ms:.4f} ms  -> {T / ms * 1e3 / 1e6:.2f} M tok/s, {4 * T * V / ms / 1e6:.0f} GB/s of 4TV")
    except Exception as e:
        print(f"single-pass cluster={c} nbuf={nb}: failed {repr(e)[:150]}")
