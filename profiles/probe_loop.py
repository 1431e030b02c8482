"""Probe: device time of the fused forward inside un-synchronised loops (debug aid)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import functional as F, ppo_functional as P
import bench
dev = torch.device("cuda:0")
b = bench.make_batch(torch, dev, 16384, 151936, 0)
lp = F.gather_logprobs(b["logits"], b["labels"], 1.0)
prox = lp + 0.02 * torch.randn_like(lp); old = lp + 0.05 * torch.randn_like(lp)
x = b["logits"].requires_grad_()
def fused():
    return F.fused_actor_loss(x, b["labels"], 1.0, prox, old, b["advantages"], b["loss_mask"], eps_clip=0.4, behav_imp_weight_cap=5.0)
def ev():
    return torch.cuda.Event(enable_timing=True)
def run(name, body, n=16):
    for _ in range(3): body(None)
    torch.cuda.synchronize()
    recs = []
    for _ in range(n):
        r = {}
        body(r); recs.append(r)
    torch.cuda.synchronize()
    keys = [k for k in recs[0] if k.endswith("0")]
    for k in keys:
        base = k[:-1]
        print(name, base, [round(r[base + "0"].elapsed_time(r[base + "1"]), 3) for r in recs])
def A(r):
    if r is not None: r["fwd0"] = ev(); r["fwd0"].record()
    with torch.no_grad(): fused()
    if r is not None: r["fwd1"] = ev(); r["fwd1"].record()
def B(r):
    if r is not None: r["fwd0"] = ev(); r["fwd0"].record()
    l, _ = fused()
    if r is not None: r["fwd1"] = ev(); r["fwd1"].record(); r["bwd0"] = r["fwd1"]
    x.grad = None; l.backward()
    if r is not None: r["bwd1"] = ev(); r["bwd1"].record()
def C(r):
    P.cugae1d_nolp_misalign_func(b["rewards"], b["values"], b["cu_seqlens"], b["bootstrap"], 1.0, 0.95)
    B(r)
def Dd(r):   # keep grad buffer: no free/alloc of the 5 GB dlogits between steps
    if r is not None: r["fwd0"] = ev(); r["fwd0"].record()
    l, _ = fused()
    if r is not None: r["fwd1"] = ev(); r["fwd1"].record(); r["bwd0"] = r["fwd1"]
    l.backward()
    if r is not None: r["bwd1"] = ev(); r["bwd1"].record()
run("A fwd-only loop", A)
run("B fwd,bwd loop", B)
run("C gae,fwd,bwd loop", C)
x.grad = None
run("D fwd,bwd (grad accumulates)", Dd, n=6)
