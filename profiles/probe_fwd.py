"""Probe: per-launch device time of the fused forward / backward at the cfg-2 shape (debug aid)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import functional as F
import bench
dev = torch.device("cuda:0")
b = bench.make_batch(torch, dev, 16384, 151936, 0)
lp = F.gather_logprobs(b["logits"], b["labels"], 1.0)
prox = lp + 0.02 * torch.randn_like(lp); old = lp + 0.05 * torch.randn_like(lp)
x = b["logits"].requires_grad_()
def t(fn, n=12):
    out = []
    for _ in range(n):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); r = fn(); e1.record(); torch.cuda.synchronize(); out.append(round(e0.elapsed_time(e1), 3))
    return out
print("fwd logits_stats (minmax):", t(lambda: F.logits_stats(b["logits"], b["labels"], 1.0)))
print("fwd gather_logprobs_entropy:", t(lambda: F.gather_logprobs_entropy(b["logits"].detach(), b["labels"], 1.0)))
def fused():
    return F.fused_actor_loss(x, b["labels"], 1.0, prox, old, b["advantages"], b["loss_mask"], eps_clip=0.4, behav_imp_weight_cap=5.0)
print("fwd fused:", t(fused))
loss, st = fused()
def bwd():
    x.grad = None
    loss.backward(retain_graph=True)
print("bwd:", t(bwd))
def both():
    x.grad = None
    l, _ = fused(); l.backward()
print("fwd+bwd:", t(both))
import subprocess
print(subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.mem,power.draw,temperature.gpu", "--format=csv,noheader"], capture_output=True, text=True).stdout)
