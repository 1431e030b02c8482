"""CPU: host logic of the stats-tracker mirror (validation, scopes, scalar keys, export arithmetic and
the world_size-2 export protocol under gloo).  The reductions themselves need the GPU
(tests/test_gpu_stats.py)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from areal_b200 import stats_tracker as ST

INF = float("inf")


def test_validation_matches_reference_messages():
    tr = ST.DistributedStatsTracker()
    with pytest.raises(ValueError, match="must be a pytorch bool tensor"):
        tr.denominator(n=torch.ones(3))
    with pytest.raises(ValueError, match="must be non-empty"):
        tr.denominator(n=torch.ones(0, dtype=torch.bool))
    with pytest.raises(ValueError, match="should be a pytorch float tensor"):
        tr.stat(denominator="n", x=torch.ones(3, dtype=torch.float64))
    with pytest.raises(ValueError, match="Denominator `n` does not exist"):
        tr.stat(denominator="n", x=torch.ones(3))
    with pytest.raises(ValueError, match="Cannot use the scalar reduce type"):
        tr.stat(denominator="n", reduce_type=ST.ReduceType.SCALAR, x=torch.ones(3))
    with pytest.raises(ValueError, match="reduce_type must be a ReduceType enum"):
        tr._set_reduce_type("k", "avg")
    with pytest.raises(_lib_error()):
        tr.denominator(n=torch.ones(3, dtype=torch.bool))          # CPU tensor: no fallback


def _lib_error():
    from areal_b200._lib import ArealB200Error
    return ArealB200Error


def test_scopes_scalars_and_timing():
    tr = ST.DistributedStatsTracker("root/")
    with tr.scope("a"):
        tr.scalar(lr=1.0)
        tr.scalar(lr=3.0)
        with tr.scope("/b/"):
            tr.scalar(x=5)
        with tr.disable_scope():
            tr.scalar(y=7)
    with tr.record_timing("step"):
        pass
    out = tr.export()
    assert out["root/a/lr"] == 2.0 and out["root/a/b/x"] == 5.0 and out["y"] == 7.0
    assert out["timeperf/step"] >= 0.0
    assert tr.export() == {}                                       # reset


def _inject(tr, key, rtype, vec, denominator=None):
    if tr._acc is None:
        tr._acc = torch.tensor(ST._IDENTITY, dtype=torch.float64).repeat(8, 1)
    slot = tr._next_slot
    tr._next_slot += 1
    tr._slot[key] = slot
    tr._acc[slot] = torch.tensor(vec, dtype=torch.float64)
    tr.reduce_types[key] = rtype
    if denominator:
        tr.denominators[key] = denominator


def test_export_arithmetic_single_rank():
    tr = ST.DistributedStatsTracker()
    _inject(tr, "n", ST.ReduceType.SUM, (5, 5, 1, 1))
    _inject(tr, "x", ST.ReduceType.AVG_MIN_MAX, (10.0, 4, -2.0, 6.0), "n")
    _inject(tr, "empty", ST.ReduceType.AVG_MIN_MAX, (0.0, 0, INF, -INF), "n")
    _inject(tr, "s", ST.ReduceType.SUM, (3.5, 2, 0, 0), "n")
    _inject(tr, "lo", ST.ReduceType.MIN, (0, 1, -4.0, -4.0), "n")
    _inject(tr, "hi", ST.ReduceType.MAX, (0, 1, 9.0, 9.0), "n")
    _inject(tr, "a", ST.ReduceType.AVG, (9.0, 3, 0, 0), "n")
    out = tr.export(reset=False)
    assert out == {"n": 5.0, "x/avg": 2.5, "x/min": -2.0, "x/max": 6.0, "s": 3.5, "lo": -4.0, "hi": 9.0, "a": 3.0}
    assert tr.export(key="x") == {"x/avg": 2.5, "x/min": -2.0, "x/max": 6.0}
    assert "x/avg" not in tr.export()                               # its totals were restarted


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        tr = ST.DistributedStatsTracker()
        _inject(tr, "n", ST.ReduceType.SUM, (4 + rank, 4 + rank, 1, 1))
        _inject(tr, "x", ST.ReduceType.AVG_MIN_MAX, (10.0 * (rank + 1), 4 + rank, -1.0 - rank, 2.0 + rank), "n")
        if rank == 1:                                               # a key only one rank recorded
            _inject(tr, "only1", ST.ReduceType.AVG_MIN_MAX, (6.0, 3, 1.0, 3.0), "n")
        tr.scalar(lr=1.0 + rank)
        q.put((rank, tr.export(reduce_group=dist.group.WORLD)))
    finally:
        dist.destroy_process_group()


def test_export_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    exp = {"n": 9.0, "x/avg": 30.0 / 9.0, "x/min": -2.0, "x/max": 3.0, "only1/avg": 2.0, "only1/min": 1.0,
           "only1/max": 3.0, "lr": 1.5}
    for r in (0, 1):
        assert res[r].keys() == exp.keys()
        for k, v in exp.items():
            assert abs(res[r][k] - v) < 1e-12, (r, k, res[r][k], v)
