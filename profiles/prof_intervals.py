"""One slice_intervals call on the 7B tensor-parallel re-shard shape (200,704 row pieces), for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from areal_b200 import flatten_param as fp
D = torch.device("cuda:0")
n_elem = 28 * 3584 * (3584 + 18944)
src = torch.empty(n_elem, dtype=torch.bfloat16, device=D)
src.view(torch.int16).random_(-30000, 30000)
rows, base = [], 0
for layer in range(28):
    for width in (3584, 18944):
        for rr in range(3584):
            rows.append((base + rr * width + width // 2, base + (rr + 1) * width))
        base += 3584 * width
iv = np.array(rows, dtype=np.int64)
total = int((iv[:, 1] - iv[:, 0]).sum())
ivc = torch.from_numpy(iv).to(D)
for _ in range(3):
    out = fp.slice_intervals_bf16(src, ivc, total, 9472)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
out = fp.slice_intervals_bf16(src, ivc, total, 9472)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("bytes moved (read + write):", 4 * total)
