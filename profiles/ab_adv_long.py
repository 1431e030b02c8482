"""A/B of compute_advantages on few long rows: one cluster of CTAs per row (default) against one CTA per row
(AREAL_B200_PADDED_CLUSTER=0, read once per process -> run this twice).  Prints the GPU time of the calls replayed
from a CUDA graph over a ring of input sets, and the kernels of one call (torch profiler)."""
import os
import sys
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200.actor import compute_advantages  # noqa: E402

dev = "cuda:0"
B, S = (int(x) for x in (sys.argv[1:3] if len(sys.argv) > 2 else (16, 32768)))
actor = types.SimpleNamespace(
    config=types.SimpleNamespace(use_decoupled_loss=True, recompute_logprob=True, overlong_reward_penalty=False),
    reward_bias=0.0, reward_scaling=1.0, reward_clip=20.0, kl_ctl=0.1,
    kl_estimator=types.SimpleNamespace(kl_estimator="k1", apply_clamp=True), discount=1.0, gae_lambda=0.95,
    mask_no_eos_with_zero=False, adv_norm=None, reward_norm=None)
ring = []
for _ in range(16):
    lens = torch.randint(S // 2, S + 1, (B,), device=dev)
    am = torch.arange(S, device=dev)[None, :] < lens[:, None]
    ring.append(dict(input_ids=torch.zeros(B, S, dtype=torch.long, device=dev), attention_mask=am,
                     loss_mask=am & (torch.rand(B, S, device=dev) > 0.2), logprobs=-torch.rand(B, S, device=dev),
                     ref_logp=-torch.rand(B, S, device=dev), values=torch.randn(B, S, device=dev),
                     rewards=torch.randn(B, device=dev)))


def sweep():
    for d in ring:
        compute_advantages(actor, dict(d))


for _ in range(3):
    sweep()
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    sweep()
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    g.replay()
e1.record()
torch.cuda.synchronize()
print("cluster env", os.environ.get("AREAL_B200_PADDED_CLUSTER", "default"), "B S", B, S,
      "gpu us per call %.2f" % (e0.elapsed_time(e1) / 20 / len(ring) * 1e3))
from torch.profiler import ProfilerActivity, profile  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    sweep()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=8, max_name_column_width=70))
