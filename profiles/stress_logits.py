"""Stress: full-size logits fwd + bwd repeated with the allocator's free memory poisoned in between;
every iteration must reproduce the first one bit for bit."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from areal_b200 import functional as F
D = torch.device("cuda:0")
T, V = 16384, 151936
g = torch.Generator(device=D).manual_seed(0)
logits = torch.empty(T, V, dtype=torch.bfloat16, device=D)
for i in range(0, T, 2048):
    logits[i:i + 2048] = (torch.randn(2048, V, generator=g, device=D) * 2.0).bfloat16()
labels = torch.randint(0, V, (T,), generator=g, device=D)
w = torch.randn(T, generator=g, device=D)
w[torch.rand(T, generator=g, device=D) < 0.3] = 0.0
logits.requires_grad_()
ref = None
bad = 0
for it in range(int(sys.argv[1]) if len(sys.argv) > 1 else 12):
    # poison whatever the caching allocator will hand out next
    junk = torch.empty(T, V, dtype=torch.bfloat16, device=D)
    junk.view(torch.int16).fill_(0x7FC1 if it % 2 else 0x4142)
    small = [torch.full((T,), float("nan"), device=D) for _ in range(8)]
    del junk, small
    lp, ent = F.gather_logprobs_entropy(logits, labels, 1.0)
    (grad,) = torch.autograd.grad((lp * w).sum(), logits)
    torch.cuda.synchronize()
    cur = (lp.detach().clone(), ent.detach().clone(), grad)
    if ref is None:
        ref = (cur[0], cur[1], grad.clone())
        z = int(torch.count_nonzero(grad[w == 0]))
        print("iter 0: nonzero in zero rows:", z, "nan:", int(torch.isnan(grad.float()).sum()), flush=True)
    else:
        same = [bool(torch.equal(a.view(torch.int16) if a.dtype == torch.bfloat16 else a.view(torch.int32),
                                 b.view(torch.int16) if b.dtype == torch.bfloat16 else b.view(torch.int32)))
                for a, b in zip(cur, ref)]
        if not all(same):
            bad += 1
            diff = (grad.view(torch.int16) != ref[2].view(torch.int16))
            rows = torch.nonzero(diff.any(-1)).flatten()[:8].tolist()
            print("iter", it, "MISMATCH", same, "rows", rows, "count", int(diff.sum()), flush=True)
            if rows:
                r = rows[0]
                cols = torch.nonzero(diff[r]).flatten()
                print("  row", r, "w", float(w[r]), "ncols", len(cols), "first cols", cols[:6].tolist(), "last", cols[-3:].tolist(),
                      "vals", grad[r, cols[:4]].tolist(), "ref", ref[2][r, cols[:4]].tolist(), flush=True)
    del grad, cur
print("done, mismatching iterations:", bad)
