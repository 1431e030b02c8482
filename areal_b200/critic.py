"""Host-side mirror of the critic's loss function, `ppo_loss_fn` (areal/engine/ppo/critic.py:81-127):
same arguments, same keys recorded under the same denominators, same returned loss.  The value loss and
its per-token statistics are one launch (`areal_b200.functional.ppo_critic_loss_fn`), the nine tracker
keys one more (`stats_tracker.record`); no `[T]` temporaries for `clip_mask.float()` / `torch.where`."""
from __future__ import annotations

from typing import Dict

import torch


def ppo_loss_fn(value: torch.Tensor, input_data: Dict, eps_clip: float, tracker=None):
    from . import functional as F
    from . import stats_tracker as ST
    from .actor import _all_tokens
    tracker = ST.DEFAULT_TRACKER if tracker is None else tracker
    value = value.squeeze(-1).float()
    old_value = input_data["values"].float()
    target_value = input_data["returns"].float()
    loss_mask = input_data["loss_mask"].bool()
    loss, stat = F.ppo_critic_loss_fn(value=value, old_value=old_value, target_value=target_value,
                                      value_eps_clip=eps_clip, loss_mask=loss_mask)
    v_d = value.detach()
    tracker.record(
        dict(n_tokens=_all_tokens(value.shape[0], value.device), n_valid_tokens=loss_mask,
             clipped_tokens=stat["clip_mask"]),
        [("n_valid_tokens", dict(new_value=v_d, old_value=old_value, target_value=target_value,
                                 critic_loss=stat["loss"], clip_ratio=stat["clip_mask"])),
         # torch.where(clip_mask, x, 0.0) under the `clipped_tokens` denominator selects the clipped entries of x
         ("clipped_tokens", dict(clipped_new_value=v_d, clipped_old_value=old_value))])
    return loss
