"""compute_advantages (fused rewards + padded GAE, one launch) on many rows: eager ms per call, CUDA events.
AREAL_B200_PADDED_THREADS = 64 | 128 | 256 picks the CTA width (read once per process)."""
import os
import sys
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from areal_b200.actor import compute_advantages  # noqa: E402

dev = "cuda:0"
actor = types.SimpleNamespace(
    config=types.SimpleNamespace(use_decoupled_loss=True, recompute_logprob=True, overlong_reward_penalty=False),
    reward_bias=0.0, reward_scaling=1.0, reward_clip=20.0, kl_ctl=0.1,
    kl_estimator=types.SimpleNamespace(kl_estimator="k1", apply_clamp=True), discount=1.0, gae_lambda=0.95,
    mask_no_eos_with_zero=False, adv_norm=None, reward_norm=None)
for B, S in ((2048, 8192), (512, 4096), (4096, 4096), (256, 16384)):
    lens = torch.randint(2, S + 1, (B,), device=dev)
    am = torch.arange(S, device=dev)[None, :] < lens[:, None]
    base = dict(input_ids=torch.zeros(B, S, dtype=torch.long, device=dev), attention_mask=am,
                loss_mask=am & (torch.rand(B, S, device=dev) > 0.2), logprobs=-torch.rand(B, S, device=dev),
                ref_logp=-torch.rand(B, S, device=dev), values=torch.randn(B, S, device=dev),
                rewards=torch.randn(B, device=dev))
    for _ in range(3):
        compute_advantages(actor, dict(base))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        compute_advantages(actor, dict(base))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("threads", os.environ.get("AREAL_B200_PADDED_THREADS", "256"), "B S", B, S, "us %.1f" % (ms * 1e3),
          "GB/s %.0f" % (42.0 * B * S / (ms * 1e-3) / 1e9))
    del base, am
