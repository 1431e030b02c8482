"""One packed-GAE call at 2^22, 2^26 and 2^27 tokens (for an ncu launch list: prep kernel vs scan kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench_kernels as BK
from areal_b200 import ppo_functional as P
rng = np.random.RandomState(0)
for logT in (22, 26, 27):
    T = 1 << logT
    lens = BK.make_packed(T, rng); B = len(lens)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(BK.DEV)
    r = torch.randn(T, device=BK.DEV); v = torch.randn(T + B, device=BK.DEV); boot = torch.rand(B, device=BK.DEV) > 0.5
    for _ in range(3):
        P.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 0.95)
    torch.cuda.synchronize()
    ms = BK.timeit(lambda: P.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 0.95), iters=10, warmup=2)
    print("T", T, "B", B, "event ms", round(ms, 4), flush=True)
    del r, v
