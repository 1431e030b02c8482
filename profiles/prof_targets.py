"""Profiling target: a few launches of gae1d (large T), interval copy (7B TP-reshard rows) for ncu."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import flatten_param as fp, ppo_functional as P
import bench_kernels as BK
DEV = torch.device("cuda:0")
rng = np.random.RandomState(0)
T = 1 << 26
lens = BK.make_packed(T, rng); B = len(lens)
cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(DEV)
r = torch.randn(T, device=DEV); v = torch.randn(T + B, device=DEV); boot = torch.rand(B, device=DEV) > 0.5
for _ in range(3):
    P.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 0.95)
torch.cuda.synchronize()
rows, base = [], 0
for layer in range(28):
    for width in (3584, 18944):
        for rr in range(3584):
            rows.append((base + rr * width + width // 2, base + (rr + 1) * width))
        base += 3584 * width
iv = np.array(rows, dtype=np.int64); total = int((iv[:, 1] - iv[:, 0]).sum())
src = torch.empty(base, dtype=torch.bfloat16, device=DEV); src.view(torch.int16).random_(-30000, 30000)
ivc = torch.from_numpy(iv).to(DEV)
for _ in range(3):
    fp.slice_intervals_bf16(src, ivc, total, 9472)
torch.cuda.synchronize()
print("ok")
