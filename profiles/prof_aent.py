"""Profiling target: a few forward launches of the AEnt clamped-entropy kernel (T=2048, V=151936)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import aent_functional as A
D = torch.device("cuda:0")
g = torch.Generator(device=D).manual_seed(0)
x = (torch.randn(2048, 151936, generator=g, device=D) * 2).bfloat16()
lab = torch.randint(0, 151936, (2048,), generator=g, device=D)
for _ in range(3):
    A.gather_logprobs_clamped_entropy(x, lab, 0.3, 1.0)
torch.cuda.synchronize()
print("ok")
