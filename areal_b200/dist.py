"""Multi-GPU plumbing of the hot path: one process per GPU, `torch.distributed` (NCCL over
NVLink 5 / NVSwitch) for the few bytes that must cross ranks.

Two ways the path shards:

* by packed sequence (data parallel, BASELINE.json configs[2]): sequences are independent for GAE and
  rows are independent for the logits pass, so ranks exchange NO tensors; only the masked-mean
  normalisation needs the global token count -- one all-reduce of two floats, the same exchange the
  reference engines do for their loss weight (areal/engine/fsdp_engine.py:518-526).
* by vocabulary (tensor parallel, configs[3]): every rank streams its `[T, V/R]` shard once
  (areal_b200_logits_partial_fwd), ranks exchange `[4, T]` f32 partials with ONE all-gather, and each
  combines locally in a fixed order (areal_b200_vocab_combine) -- the reference needs three
  all-reduces for the entropy (areal/utils/mcore/functional.py:16-31) plus three inside Megatron's
  vocab_parallel_cross_entropy.  The backward is local to the shard.
"""
from __future__ import annotations

import ctypes
from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from . import _lib
from ._lib import call, ptr, stream_ptr


# --------------------------------------------------------------------------------------------------
# batch (sequence) sharding
# --------------------------------------------------------------------------------------------------
def partition_sequences(seqlens: Sequence[int], n_parts: int) -> List[List[int]]:
    """Token-balanced assignment of whole sequences to ranks (longest-first greedy, like the
    reference's allocate_balanced_mbs, areal/utils/data.py:240-263).  Returns, per rank, the
    sequence indices in ascending order; every sequence appears exactly once."""
    if n_parts <= 0:
        raise ValueError("n_parts must be positive")
    order = sorted(range(len(seqlens)), key=lambda i: (-int(seqlens[i]), i))
    loads = [0] * n_parts
    parts: List[List[int]] = [[] for _ in range(n_parts)]
    for i in order:
        r = min(range(n_parts), key=lambda k: (loads[k], k))
        parts[r].append(i)
        loads[r] += int(seqlens[i])
    return [sorted(p) for p in parts]


def take_sequences(x: torch.Tensor, cu_seqlens: torch.Tensor, seq_ids: Sequence[int], extra_per_seq: int = 0):
    """Packed sub-batch holding the given sequences of `x` (dim 0 packed by `cu_seqlens`).
    `extra_per_seq=1` is the "misaligned" layout of the GAE values (one bootstrap slot per
    sequence: sequence i occupies [cu[i]+i, cu[i+1]+i+1)).  Returns (packed, new_cu_seqlens)."""
    cu = cu_seqlens.tolist()
    pieces, lens = [], []
    for i in seq_ids:
        s = cu[i] + i * extra_per_seq
        e = cu[i + 1] + (i + 1) * extra_per_seq
        pieces.append(x[s:e])
        lens.append(cu[i + 1] - cu[i])
    out = torch.cat(pieces, 0) if pieces else x[:0]
    new_cu = torch.zeros(len(lens) + 1, dtype=cu_seqlens.dtype, device=cu_seqlens.device)
    if lens:
        new_cu[1:] = torch.tensor(lens, dtype=cu_seqlens.dtype, device=cu_seqlens.device).cumsum(0)
    return out, new_cu


def global_loss_scale(local_count: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """(local_count / global_count, global_count) with ONE all-reduce of one float.  Multiply the
    local masked-mean loss by the first value and sum over ranks to get the global masked mean."""
    total = local_count.detach().to(torch.float32).reshape(1).clone()
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(total, group=group)
    total = torch.clamp(total, min=1.0)
    return local_count.to(torch.float32) / total[0], total[0]


def all_reduce_loss(local_loss: torch.Tensor, local_count: torch.Tensor, group=None) -> torch.Tensor:
    """Global masked-mean loss from per-rank masked means: one all-reduce of (numerator, count)."""
    buf = torch.stack([local_loss.detach().float() * local_count.float(), local_count.float()])
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(buf, group=group)
    return buf[0] / torch.clamp(buf[1], min=1.0)


# --------------------------------------------------------------------------------------------------
# vocab sharding
# --------------------------------------------------------------------------------------------------
def gather_partials(parts: torch.Tensor, group=None) -> torch.Tensor:
    """[4, T] per rank -> [R, 4, T] in rank order (one all-gather; deterministic combine order)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return parts.unsqueeze(0)
    parts = parts.contiguous()
    out = torch.empty((world * parts.shape[0],) + tuple(parts.shape[1:]), dtype=parts.dtype, device=parts.device)
    dist.all_gather_into_tensor(out, parts, group=group)
    return out.view((world,) + tuple(parts.shape))


def _cuda_partials(shard: torch.Tensor, labels: torch.Tensor, vocab_start: int, temperature: float,
                   raw_min: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`raw_min` (f32 [T], optional): filled with the minimum of the shard's raw logits per row, out of the same pass."""
    T, V = shard.shape
    dev = shard.device
    parts = torch.empty(4, T, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        call("areal_b200_logits_partial_fwd", ptr(shard), _lib.DTYPE_CODE[shard.dtype], T, V,
             shard.stride(0) if T > 1 else V, ptr(labels), int(vocab_start), float(temperature), ptr(parts), ptr(raw_min),
             stream_ptr(dev))
    return parts


def _cuda_combine(all_parts: torch.Tensor, temperature: float):
    R, _, T = all_parts.shape
    dev = all_parts.device
    f32 = dict(dtype=torch.float32, device=dev)
    logp, ent, row_max, ln_s = (torch.empty(T, **f32) for _ in range(4))
    with torch.cuda.device(dev):
        call("areal_b200_vocab_combine", ptr(all_parts), R, T, float(temperature), ptr(logp), ptr(ent),
             ptr(row_max), ptr(ln_s), stream_ptr(dev))
    return logp, ent, row_max, ln_s


class PeerExchange:
    """NVLink peer-memory exchange for the vocab-sharded forward: every rank's `[4,T]` partials live in a
    buffer that all ranks of `group` have mapped (torch symmetric memory does the rendezvous and the
    mapping -- plumbing), and `areal_b200_vocab_combine_p2p` reads them where they are.  Replaces the
    all-gather + combine pair by one launch; results are bit-identical.

    One instance per (group, stream of use); `max_tokens` bounds T.  All ranks must call `forward` the
    same number of times (the step counter is the handshake's epoch)."""

    TIMEOUT_CYCLES = 20_000_000_000          # ~10 s of SM clocks: a peer that never shows up is an error, not a hang

    def __init__(self, max_tokens: int, group=None, device=None):
        import torch.distributed._symmetric_memory as symm
        if not dist.is_initialized():
            raise RuntimeError("PeerExchange needs an initialised process group")
        self.group = dist.group.WORLD if group is None else group
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        if self.world > 16:
            raise ValueError("PeerExchange supports at most 16 ranks")
        self.max_tokens = int(max_tokens)
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device = dev
        nbytes = int(_lib.lib().areal_b200_vocab_p2p_buffer_bytes(self.max_tokens))
        try:
            symm.enable_symm_mem_for_group(self.group.group_name)
        except Exception:
            pass                              # newer torch enables groups on demand
        self.buf = symm.empty(nbytes, dtype=torch.uint8, device=dev)
        self.buf.zero_()
        self.handle = symm.rendezvous(self.buf, self.group)
        self._peers = (ctypes.c_void_p * self.world)(*[int(p) for p in self.handle.buffer_ptrs])
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self.epoch = 0
        torch.cuda.synchronize(dev)
        dist.barrier(self.group)              # every buffer is zeroed before anybody's first flag lands

    def parts_slot(self, epoch: int, T: int) -> torch.Tensor:
        """The `[4,T]` f32 window of this rank's buffer that step `epoch` writes."""
        off = 256 + (epoch & 1) * 16 * self.max_tokens
        return self.buf[off: off + 16 * T].view(torch.float32).view(4, T)

    def forward(self, shard: torch.Tensor, labels: torch.Tensor, vocab_start: int, temperature: float):
        T, V = shard.shape
        if T > self.max_tokens:
            raise ValueError(f"T={T} exceeds the exchange buffer ({self.max_tokens} tokens)")
        dev = self.device
        self.epoch += 1
        parts = self.parts_slot(self.epoch, T)
        f32 = dict(dtype=torch.float32, device=dev)
        logp, ent, row_max, ln_s = (torch.empty(T, **f32) for _ in range(4))
        with torch.cuda.device(dev):
            if T:
                call("areal_b200_logits_partial_fwd", ptr(shard), _lib.DTYPE_CODE[shard.dtype], T, V,
                     shard.stride(0) if T > 1 else V, ptr(labels), int(vocab_start), float(temperature), ptr(parts),
                     None, stream_ptr(dev))
            call("areal_b200_vocab_combine_p2p", self._peers, self.rank, self.world, T, self.max_tokens, self.epoch,
                 float(temperature), self.TIMEOUT_CYCLES, ptr(logp), ptr(ent), ptr(row_max), ptr(ln_s),
                 ptr(self.status), stream_ptr(dev))
        return logp, ent, row_max, ln_s

    def check(self):
        """Host sync: raises if some step timed out waiting for a peer."""
        if int(self.status.item()) != 0:
            raise RuntimeError("areal_b200 peer exchange: a rank did not publish its partials in time")


class FusedPeerExchange:
    """The vocab-sharded forward as ONE launch (areal_b200_logits_partial_fwd_push): the streaming kernel's row epilogue
    stores a row's four partial numbers into EVERY rank's exchange buffer over NVLink (plain 16-byte stores into
    peer-mapped symmetric memory), a CTA that has finished its rows announces their number to every rank, and once a
    rank's counter says that all R*T rows are in, the same CTAs combine them from local memory.  The transfer rides
    under the streaming pass: no all-gather, no second launch, no remote reads.  Same bits as `PeerExchange` and as the
    all-gather path.

    One instance per (group, stream of use); `max_tokens` bounds T.  All ranks must call `forward` the same number of
    times with the same T (the step counter picks the buffer slot; the expected row count is cumulative)."""

    TIMEOUT_CYCLES = 20_000_000_000          # ~10 s of SM clocks: a rank that never delivers is an error, not a hang

    def __init__(self, max_tokens: int, group=None, device=None):
        import torch.distributed._symmetric_memory as symm
        if not dist.is_initialized():
            raise RuntimeError("FusedPeerExchange needs an initialised process group")
        self.group = dist.group.WORLD if group is None else group
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        if self.world > 16:
            raise ValueError("FusedPeerExchange supports at most 16 ranks")
        self.max_tokens = int(max_tokens)
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device = dev
        nbytes = int(_lib.lib().areal_b200_vocab_push_buffer_bytes(self.max_tokens, self.world))
        try:
            symm.enable_symm_mem_for_group(self.group.group_name)
        except Exception:
            pass                              # newer torch enables groups on demand
        self.buf = symm.empty(nbytes, dtype=torch.uint8, device=dev)
        self.buf.zero_()
        self.handle = symm.rendezvous(self.buf, self.group)
        self._peers = (ctypes.c_void_p * self.world)(*[int(p) for p in self.handle.buffer_ptrs])
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self.epoch = 0
        self.rows_target = [0, 0]             # per slot: rows delivered by all ranks over all steps so far
        torch.cuda.synchronize(dev)
        dist.barrier(self.group)              # every buffer is zeroed before anybody's first row lands

    def forward(self, shard: torch.Tensor, labels: torch.Tensor, vocab_start: int, temperature: float,
                raw_min: Optional[torch.Tensor] = None):
        T, V = shard.shape
        if T > self.max_tokens:
            raise ValueError(f"T={T} exceeds the exchange buffer ({self.max_tokens} tokens)")
        dev = self.device
        self.epoch += 1
        slot = self.epoch & 1
        self.rows_target[slot] = (self.rows_target[slot] + self.world * T) & 0xFFFFFFFF
        f32 = dict(dtype=torch.float32, device=dev)
        logp, ent, row_max, ln_s = (torch.empty(T, **f32) for _ in range(4))
        with torch.cuda.device(dev):
            call("areal_b200_logits_partial_fwd_push", ptr(shard), _lib.DTYPE_CODE[shard.dtype], T, V,
                 shard.stride(0) if T > 1 else V, ptr(labels), int(vocab_start), float(temperature), self._peers,
                 self.rank, self.world, self.max_tokens, self.epoch, self.rows_target[slot], self.TIMEOUT_CYCLES,
                 ptr(logp), ptr(ent), ptr(row_max), ptr(ln_s), ptr(raw_min), ptr(self.status), stream_ptr(dev))
        return logp, ent, row_max, ln_s

    def check(self):
        """Host sync: raises if some step timed out waiting for a rank's rows."""
        if int(self.status.item()) != 0:
            raise RuntimeError("areal_b200 fused peer exchange: a rank did not deliver its rows in time")


def vocab_parallel_forward(shard: torch.Tensor, labels: torch.Tensor, vocab_start: int, temperature: float,
                           group=None, partials_fn: Optional[Callable] = None,
                           combine_fn: Optional[Callable] = None):
    """partial kernel -> all-gather -> combine.  `partials_fn` / `combine_fn` default to the CUDA
    kernels; tests inject CPU stand-ins to exercise the collective protocol under gloo."""
    parts = (partials_fn or _cuda_partials)(shard, labels, vocab_start, temperature)
    allp = gather_partials(parts, group)
    return (combine_fn or _cuda_combine)(allp, temperature)


class _VocabParallelLogprobsEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, shard, labels, vocab_start, temperature, group, exchange=None, with_minmax=False):
        ctx.in_shape = shard.shape
        if shard.dim() != 2 or shard.stride(1) != 1:
            shard = shard.reshape(-1, shard.shape[-1]).contiguous()
        _lib.require_cuda(shard, labels)
        lab = labels.reshape(-1).long().contiguous()
        local_min = local_max = None
        if with_minmax:
            # the shard's own extremes come out of the partial kernel: parts[0] is the shard maximum
            local_min = torch.empty(shard.shape[0], dtype=torch.float32, device=shard.device)
            parts = _cuda_partials(shard, lab, vocab_start, temperature, raw_min=local_min)
            local_max = parts[0].clone()
            logp, ent, row_max, ln_s = _cuda_combine(gather_partials(parts, group), temperature)
        elif exchange is not None:
            logp, ent, row_max, ln_s = exchange.forward(shard, lab, vocab_start, temperature)
        else:
            logp, ent, row_max, ln_s = vocab_parallel_forward(shard, lab, vocab_start, temperature, group)
        ctx.save_for_backward(shard, lab, row_max, ln_s, ent)
        ctx.vocab_start, ctx.temperature = int(vocab_start), float(temperature)
        ctx.set_materialize_grads(False)
        if with_minmax:
            ctx.mark_non_differentiable(local_min, local_max)
            return logp, ent, local_min, local_max
        return logp, ent

    @staticmethod
    def backward(ctx, g_logp, g_ent, *unused):
        shard, lab, row_max, ln_s, ent = ctx.saved_tensors
        if g_logp is None and g_ent is None:
            return None, None, None, None, None, None, None
        from .functional import _f32, _logits_backward
        gl = None if g_logp is None else _f32(g_logp.reshape(-1))
        gh = None if g_ent is None else _f32(g_ent.reshape(-1))
        d = _logits_backward(shard, lab, ctx.temperature, row_max, ln_s, ent, gl, gh, vocab_start=ctx.vocab_start)
        return d.view(ctx.in_shape), None, None, None, None, None, None


def vocab_parallel_logprobs_entropy(shard_logits: torch.Tensor, labels: torch.Tensor, temperature: float = 1.0,
                                    group=None, vocab_start: Optional[int] = None,
                                    exchange: Optional[PeerExchange] = None, with_minmax: bool = False):
    """Log-prob of `labels` (GLOBAL vocabulary ids) and entropy from vocab-sharded logits
    `[T, V/R]`; replaces the Megatron-TP branch of gather_logprobs_entropy
    (areal/utils/functional.py:91-98).  `vocab_start` defaults to rank * shard_width.  Every rank gets
    the full `[T]` results (as after the reference's all-reduces); gradients flow to the local shard.
    `exchange`: a `PeerExchange` -- the partials are then read from the peers' memory over NVLink by the
    combine kernel itself instead of being all-gathered first (same bits, one launch less).
    `with_minmax=True`: also returns the row minimum and maximum of THIS rank's shard (what the reference logs as
    vocab_min_logits / vocab_max_logits under tensor parallelism), from the same pass over the shard."""
    if vocab_start is None:
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        vocab_start = rank * shard_logits.shape[-1]
    return _VocabParallelLogprobsEntropy.apply(shard_logits, labels, vocab_start, temperature, group, exchange, with_minmax)


# --------------------------------------------------------------------------------------------------
# parameter reallocation: slice straight into a peer's staging buffer over NVLink
# --------------------------------------------------------------------------------------------------
class PeerPackedBuffer:
    """Packed staging buffers that every rank of `group` has mapped (torch symmetric memory does the
    rendezvous -- plumbing).  The reference moves re-sharded parameters as slice_intervals -> broadcast of
    the packed buffer -> set_intervals (realhf/impl/model/nn/real_llm_api.py:745-805).  Here the sender's
    `slice_intervals` kernel writes its pieces directly into the RECEIVER's buffer with plain streaming
    stores over NVLink (`slice_into`), so there is no sender-side staging buffer and no collective; the
    receiver then runs `set_from_local` on its own copy.  `fence()` is the hand-off between the two: a device-side
    epoch handshake in stream order (no host synchronisation), or a host barrier on request.

    `numel` elements of `dtype` per rank.  One instance per group; every rank constructs it together."""

    def __init__(self, numel: int, dtype: torch.dtype, group=None, device=None):
        import torch.distributed._symmetric_memory as symm
        if not dist.is_initialized():
            raise RuntimeError("PeerPackedBuffer needs an initialised process group")
        self.group = dist.group.WORLD if group is None else group
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device, self.numel, self.dtype = dev, int(numel), dtype
        try:
            symm.enable_symm_mem_for_group(self.group.group_name)
        except Exception:
            pass                              # newer torch enables groups on demand
        self.buf = symm.empty(max(self.numel, 1), dtype=dtype, device=dev)
        self.handle = symm.rendezvous(self.buf, self.group)
        self._ptrs = [int(p) for p in self.handle.buffer_ptrs]
        # device-side fence: the epoch flags of the vocabulary exchange (a 256-byte symmetric header, no payload)
        self._flags = symm.empty(int(_lib.lib().areal_b200_vocab_p2p_buffer_bytes(0)), dtype=torch.uint8, device=dev)
        self._flags.zero_()
        self._flag_handle = symm.rendezvous(self._flags, self.group)
        self._flag_ptrs = (ctypes.c_void_p * self.world)(*[int(p) for p in self._flag_handle.buffer_ptrs])
        self._status = torch.zeros(1, dtype=torch.int32, device=dev)
        self._epoch = 0
        torch.cuda.synchronize(dev)
        dist.barrier(self.group)

    def local(self, numel: Optional[int] = None) -> torch.Tensor:
        """This rank's staging buffer (what the peers write into)."""
        return self.buf[: self.numel if numel is None else int(numel)]

    def slice_into(self, peer: int, src: torch.Tensor, intervals_cuda: torch.Tensor, total_size: int,
                   offset: int = 0) -> None:
        """`slice_intervals(src, intervals, total_size)` with the packed result written at element
        `offset` of rank `peer`'s staging buffer (this rank's own buffer when peer == rank)."""
        from .flatten_param import _check, _intervals_cuda
        _check(src.is_cuda and src.is_contiguous() and src.dim() == 1, "src must be a contiguous 1-D CUDA tensor")
        _check(src.dtype == self.dtype, "Source and destination tensors must have the same dtype")
        _check(0 <= peer < self.world, "peer rank out of range")
        total_size, offset = int(total_size), int(offset)
        _check(offset >= 0 and offset + total_size <= self.numel, "packed range exceeds the staging buffer")
        iv = _intervals_cuda(None, intervals_cuda, src.device)
        n = iv.size(0)
        dst = ctypes.c_void_p(self._ptrs[peer] + offset * src.element_size())
        with torch.cuda.device(self.device):
            nbytes = _lib.lib().areal_b200_interval_workspace_bytes(n)
            ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            call("areal_b200_slice_intervals", ptr(src), src.numel(), dst, total_size, ptr(iv), n,
                 src.element_size(), ptr(ws), nbytes, stream_ptr(self.device))

    def set_from_local(self, dst: torch.Tensor, intervals_cuda: torch.Tensor, total_size: int, offset: int = 0) -> None:
        """`set_intervals(local staging[offset : offset + total_size], dst, intervals)` on this rank."""
        from .flatten_param import set_intervals_any
        set_intervals_any(self.buf[int(offset): int(offset) + int(total_size)], dst, intervals_cuda)

    FENCE_TIMEOUT_CYCLES = 20_000_000_000     # ~10 s of SM clocks

    def fence(self, host: bool = False) -> None:
        """Everything `slice_into` enqueued so far on this rank's current stream has landed in the peers'
        buffers, and every rank has got here: the receivers may read, the senders may overwrite.

        Default: a DEVICE-side fence, stream-ordered, no host synchronisation -- one tiny launch that release-stores
        this step's epoch into every peer's flag word and waits (bounded) until every peer's epoch has arrived here
        (the handshake of `areal_b200_vocab_combine_p2p`, called with no rows).  What follows on this stream sees the
        peers' writes; all ranks must call it the same number of times.  `check()` reports a rank that never came.
        `host=True`: drain the stream and meet at a `dist.barrier` instead."""
        if host:
            torch.cuda.current_stream(self.device).synchronize()
            dist.barrier(self.group)
            return
        self._epoch += 1
        with torch.cuda.device(self.device):
            call("areal_b200_vocab_combine_p2p", self._flag_ptrs, self.rank, self.world, 0, 0, self._epoch, 1.0,
                 self.FENCE_TIMEOUT_CYCLES, None, None, None, None, ptr(self._status), stream_ptr(self.device))

    def check(self):
        """Host sync: raises if a device-side fence timed out waiting for a rank."""
        if int(self._status.item()) != 0:
            raise RuntimeError("areal_b200 peer buffer: a rank did not reach the fence in time")
