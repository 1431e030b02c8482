"""A/B GPU-side time of small packed-GAE calls: 10 calls captured in a CUDA graph, replayed 50 times."""
import os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code_t = '''
import sys; sys.path.insert(0, %r)
import areal_b200._lib as L
LIBOVR
import numpy as np, torch
from areal_b200 import ppo_functional as P
import bench_kernels as BK
D = torch.device("cuda:0")
rng = np.random.RandomState(0)
for logT in (13, 15, 17):
    T = 1 << logT
    lens = BK.make_packed(T, rng); B = len(lens)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(D)
    r = torch.randn(T, device=D); v = torch.randn(T + B, device=D); boot = torch.rand(B, device=D) > 0.5
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3): P.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 0.95)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10): out = P.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 0.95)
    for _ in range(5): g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): g.replay()
    e1.record(); torch.cuda.synchronize()
    print("TAG", T, round(e0.elapsed_time(e1) / 500 * 1e3, 2), "us per call", flush=True)
''' % root
for tag, lib in (("new", None), ("old", os.path.join(root, "alt", "libareal_b200_old.so")), ("new", None), ("old", os.path.join(root, "alt", "libareal_b200_old.so"))):
    code = code_t.replace("LIBOVR", ("L.LIB_PATH = %r" % lib) if lib else "").replace("TAG", tag)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root)
    print(r.stdout.strip() or r.stderr[-400:], flush=True)
