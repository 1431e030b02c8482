"""torchrun / plain-python worker: `export_all` over an NCCL group when only SCALAR keys were recorded after
a reset -- the pattern of the reference's example loop (examples/math/gsm8k_grpo.py:198-247: ppo_update
exports and resets, then `record_timing` scalars, then `export_all(reduce_group=<NCCL group>)`).  An NCCL
group cannot reduce CPU tensors, so the scalar table has to live on the GPU.  Works at any world size
(world 1: plain `python tests/run_tracker_nccl.py`)."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def check_scalar_only_export(group=None):
    from areal_b200 import stats_tracker as ST
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dev = torch.device("cuda", torch.cuda.current_device())
    tr = ST.DistributedStatsTracker("t")
    m = torch.ones(8, dtype=torch.bool, device=dev)
    tr.denominator(n=m)
    tr.stat(denominator="n", x=torch.full((8,), float(rank), device=dev))
    first = tr.export(reduce_group=group)                       # exports and RESETS (device totals dropped)
    assert abs(first["t/x/avg"] - (world - 1) / 2) < 1e-6, first
    tr.scalar(a=1.0 + rank, b=10.0)
    tr.scalar(a=3.0 + rank)
    with tr.record_timing("step"):
        pass
    out = tr.export(reduce_group=group)                         # scalars only, after a reset, NCCL group
    want_a = sum((1.0 + r) + (3.0 + r) for r in range(world)) / (2 * world)
    assert abs(out["t/a"] - want_a) < 1e-9 and abs(out["t/b"] - 10.0) < 1e-9, out
    assert "timeperf/step" in out and out["timeperf/step"] >= 0.0


def main():
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    check_scalar_only_export(dist.group.WORLD)
    dist.barrier()
    if rank == 0:
        print("TRACKER_NCCL_OK", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
