"""Profiling target: forward + backward launches of the AEnt clamped-entropy kernels (T=2048, V=151936)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import aent_functional as A
D = torch.device("cuda:0")
g = torch.Generator(device=D).manual_seed(0)
x = (torch.randn(2048, 151936, generator=g, device=D) * 2).bfloat16().requires_grad_()
lab = torch.randint(0, 151936, (2048,), generator=g, device=D)
w = torch.randn(2048, generator=g, device=D)
for _ in range(3):
    x.grad = None
    lp, h = A.gather_logprobs_clamped_entropy(x, lab, 0.3, 1.0)
    ((lp + h) * w).sum().backward()
torch.cuda.synchronize()
print("ok")
