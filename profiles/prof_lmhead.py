"""Profiling target: the tcgen05 LM-head kernel at the bench shape (T=16384, H=1536, V=151936), 3 launches."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import lm_head as LM
D = torch.device("cuda:0")
T, Hd, V = int(os.environ.get("PROF_T", 16384)), 1536, 151936
g = torch.Generator(device=D).manual_seed(3)
h = torch.randn(T, Hd, generator=g, device=D).bfloat16()
w = (torch.randn(V, Hd, generator=g, device=D) * (2.0 / Hd ** 0.5)).bfloat16()
lab = torch.randint(0, V, (T,), generator=g, device=D)
for _ in range(3):
    LM.lm_head_logprobs_tc(h, w, lab, 1.0, with_entropy=True)
torch.cuda.synchronize()
print("ok")
