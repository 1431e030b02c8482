"""GPU time of gae_padded_masked (plain padded GAE) at [B, S] from a CUDA-graph replay over a ring of input sets."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import ppo_functional as P  # noqa: E402

dev = "cuda:0"
B, S = int(sys.argv[1]), int(sys.argv[2])
n_sets = 24
sets = [(torch.randn(B, S, device=dev), torch.randn(B, S, device=dev), (torch.rand(B, S, device=dev) > 0.2).float(),
         torch.rand(B, device=dev) > 0.5) for _ in range(n_sets)]


def sweep():
    for r, v, m, ne in sets:
        P.gae_padded_masked(r, v, m, ne, 1.0, 0.95)


for _ in range(3):
    sweep()
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    sweep()
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    g.replay()
e1.record()
torch.cuda.synchronize()
print("plain padded GAE B S", B, S, "cluster env", os.environ.get("AREAL_B200_PADDED_CLUSTER", "default"),
      "gpu us per call %.2f" % (e0.elapsed_time(e1) / 20 / n_sets * 1e3))
