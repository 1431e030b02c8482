"""A/B of the narrow forward instance's width (AREAL_B200_FWD_NARROW_W = 2 | 4 consumer warps) on vocabulary-shard rows
[131072, 18992] (configs[3] per GPU at N=8) and [65536, 37984] (N=4), alternating, fresh processes, one box."""
import os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
body = '''
import torch, bench_kernels as BK
from areal_b200 import functional as F, dist as D
dev = BK.DEV
for T, V in ((131072, 18992), (65536, 37984 - 37984 % 8)):
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
for w in ("2", "4", "2", "4"):
    code = "import sys; sys.path.insert(0, %r)\n" % root + body.replace("TAG", "W=" + w)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root,
                       env=dict(os.environ, AREAL_B200_FWD_NARROW_W=w))
    print(r.stdout.strip() or r.stderr[-400:], flush=True)
