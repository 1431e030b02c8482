"""gather_logprobs_entropy forward / backward on fp32 and fp16 logits (the secondary dtypes of SURVEY 8 a-5): ms and
fraction of the measured copy rate, CUDA events, inputs larger than L2."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200 import functional as F
dev = "cuda:0"
peak = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
T, V = 8192, 151936
g = torch.Generator(device=dev).manual_seed(0)
lab = torch.randint(0, V, (T,), generator=g, device=dev)
w = torch.randn(T, generator=g, device=dev)
for dt in (torch.bfloat16, torch.float16, torch.float32):
    x = torch.empty(T, V, dtype=dt, device=dev)
    for i in range(0, T, 1024):
        x[i:i + 1024] = (torch.randn(1024, V, generator=g, device=dev) * 2).to(dt)
    es = x.element_size()
    def fwd():
        with torch.no_grad():
            return F.gather_logprobs_entropy(x, lab, 1.0)
    xg = x.requires_grad_()
    def fb():
        xg.grad = None
        lp, h = F.gather_logprobs_entropy(xg, lab, 1.0)
        (lp * w).sum().backward()
    res = {}
    for name, fn, nbytes in (("fwd", fwd, es * T * V), ("fwd+bwd", fb, 3 * es * T * V)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        res[name] = (round(ms, 3), round(nbytes / (ms * 1e-3) / 1e9 / peak, 3))
    print(str(dt), res, flush=True)
    del x, xg
    torch.cuda.empty_cache()
