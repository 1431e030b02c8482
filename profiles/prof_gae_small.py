"""Profiling target: packed GAE at 32,768 tokens on distinct (cold) input sets, and the padded GAE at [16, 32768]."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import ppo_functional as P
dev = torch.device("cuda:0")
rng = np.random.RandomState(0)
sets = []
for s in range(24):
    lens, left = [], 32768
    while left > 0:
        n = min(left, int(rng.randint(128, 4097))); lens.append(n); left -= n
    B = len(lens)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
    sets.append((torch.randn(32768, device=dev), torch.randn(32768 + B, device=dev), cu, torch.rand(B, device=dev) > 0.5))
pad = [(torch.randn(16, 32768, device=dev), torch.randn(16, 32768, device=dev), (torch.rand(16, 32768, device=dev) > 0.2).float(),
        torch.rand(16, device=dev) > 0.5) for _ in range(12)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for rep in range(2):
    flush.zero_()
    for r, v, cu, b in sets:
        P.cugae1d_nolp_misalign_func(r, v, cu, b, 1.0, 0.95)
    for r, v, m, ne in pad:
        P.gae_padded_masked(r, v, m, ne, 1.0, 0.95)
torch.cuda.synchronize()
print("ok")
