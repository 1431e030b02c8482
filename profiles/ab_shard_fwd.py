"""A/B of the forward on vocabulary-shard rows ([131072, 18992] bf16, the configs[3] per-GPU shape) and on full
rows, with the current library and with alt/libareal_b200_old.so, alternating, in fresh processes on one box."""
import os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
old = os.path.join(root, "alt", "libareal_b200_old.so")
body = '''
import torch, bench_kernels as BK
from areal_b200 import functional as F, dist as D
dev = BK.DEV
for T, V in ((131072, 18992), (16384, 151936)):
    g = torch.Generator(device=dev).manual_seed(1)
    x = torch.empty(T, V, dtype=torch.bfloat16, device=dev)
    for i in range(0, T, 8192):
        x[i:i + 8192] = (torch.randn(8192, V, generator=g, device=dev) * 2).bfloat16()
    lab = torch.randint(0, V, (T,), generator=g, device=dev)
    ms = BK.timeit(lambda: F.gather_logprobs_entropy(x, lab, 1.0), iters=12, warmup=3)
    print("TAG fwd logp+entropy", T, V, round(ms, 4), round(2.0 * T * V / (ms * 1e-3) / 1e9 / BK.PEAK, 4), flush=True)
    ms = BK.timeit(lambda: D._cuda_partials(x, lab, 0, 1.0), iters=12, warmup=3)
    print("TAG fwd partials   ", T, V, round(ms, 4), round(2.0 * T * V / (ms * 1e-3) / 1e9 / BK.PEAK, 4), flush=True)
    del x
'''
for tag, lib in (("new", None), ("old", old), ("new", None), ("old", old)):
    code = "import sys; sys.path.insert(0, %r); import areal_b200._lib as L\n" % root
    if lib:
        code += "L.LIB_PATH = %r\n" % lib
    code += body.replace("TAG", tag)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root)
    print(r.stdout.strip() or r.stderr[-400:], flush=True)
