"""Profiling target: the per-shard partial forward on [131072, 18992] bf16 (configs[3] per GPU), 3 launches."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import dist as D
dev = torch.device("cuda:0")
T, V = 131072, 18992
g = torch.Generator(device=dev).manual_seed(1)
x = torch.empty(T, V, dtype=torch.bfloat16, device=dev)
for i in range(0, T, 8192):
    x[i:i + 8192] = (torch.randn(8192, V, generator=g, device=dev) * 2).bfloat16()
lab = torch.randint(0, V * 8, (T,), generator=g, device=dev)
for _ in range(3):
    D._cuda_partials(x, lab, 0, 1.0)
torch.cuda.synchronize()
print("ok")
