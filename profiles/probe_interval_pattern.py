"""Is the 71 % of the 7B TP-reshard shape (row halves of 3.5 / 18.5 KB, every other half skipped) a per-piece
cost of the copy kernel or the DRAM access pattern?  Same piece sizes and count, three source layouts:
  strided   the real shape: the second half of every matrix row
  dense     the same pieces laid back to back in the source (no gaps; pieces still separate table entries)
  torch     torch's own strided copy of the same halves (one [rows, width/2] view per matrix), for reference"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench_kernels as BK
from areal_b200 import flatten_param as fp
D = BK.DEV
n_elem = 28 * 3584 * (3584 + 18944)
src = torch.empty(n_elem, dtype=torch.bfloat16, device=D)
src.view(torch.int16).random_(-30000, 30000)
strided, dense, views = [], [], []
base = dbase = 0
for layer in range(28):
    for width in (3584, 18944):
        h = width // 2
        for rr in range(3584):
            strided.append((base + rr * width + h, base + (rr + 1) * width))
            dense.append((dbase, dbase + h))
            dbase += h
        views.append((base, width))
        base += 3584 * width
for name, rows in (("strided", strided), ("dense", dense)):
    iv = np.array(rows, dtype=np.int64)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    ivc = torch.from_numpy(iv).to(D)
    ms = BK.timeit(lambda: fp.slice_intervals_bf16(src, ivc, total, 9472), iters=8, warmup=2)
    BK.row(f"slice_intervals, 200,704 pieces, source {name}", ms, 4 * total + 16 * len(rows))
outs = [torch.empty(3584, w // 2, dtype=torch.bfloat16, device=D) for _, w in views]
def torch_copy():
    for (b, w), o in zip(views, outs):
        o.copy_(src[b:b + 3584 * w].view(3584, w)[:, w // 2:])
ms = BK.timeit(torch_copy, iters=8, warmup=2)
BK.row("torch strided copy_ of the same halves (56 launches)", ms, 4 * total)
big = torch.empty(total, dtype=torch.bfloat16, device=D)
ms = BK.timeit(lambda: big.copy_(src[:total]), iters=8, warmup=2)
BK.row("torch contiguous copy_ of the same number of bytes", ms, 4 * total)
