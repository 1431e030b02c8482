"""Profiling target: packed GAE at 2^27 tokens (inputs larger than L2), three calls."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import ppo_functional as P
dev = torch.device("cuda:0")
T = 1 << 27
rng = np.random.RandomState(T % 1000)
lens, left = [], T
while left > 0:
    n = min(left, int(rng.randint(128, 4097))); lens.append(n); left -= n
B = len(lens)
cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
r, v, boot = torch.randn(T, device=dev), torch.randn(T + B, device=dev), torch.rand(B, device=dev) > 0.5
for _ in range(3):
    P.cugae1d_nolp_misalign_func(r, v, cu, boot, 1.0, 0.95)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    P.cugae1d_nolp_misalign_func(r, v, cu, boot, 1.0, 0.95)
e1.record()
torch.cuda.synchronize()
print("us per call %.1f" % (e0.elapsed_time(e1) / 5 * 1e3), "B", B)
