"""Head-sharded partition of the TriangleMix prefill path across the GPUs of one box.

Heads never interact inside attention, so rank r of P owns KV heads
[r*Hkv/P, (r+1)*Hkv/P) and the Hq/Hkv query heads of each -- the split the
reference computes per tensor-parallel rank (vllm/model_executor/models/qwen3.py:85-98:
``num_heads = total/tp``, ``num_kv_heads = max(1, total_kv/tp)``).  The block index is
head-independent and rebuilt per rank; the only collective on the path is one
all-gather of the [T, Hq/P, D] output shard (the analogue of ``gather_output`` in the
reference's column-parallel layers, vllm/model_executor/layers/linear.py:559).

Host logic only (works with any torch.distributed backend; the CPU tests use gloo).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class HeadPartition:
    rank: int
    world: int
    num_heads: int          # query heads on this rank
    num_kv_heads: int       # KV heads on this rank
    q_head_start: int       # first global query head owned
    kv_head_start: int      # first global KV head owned
    kv_replicas: int        # ranks sharing one KV head (1 unless world > total_kv_heads)

    @property
    def q_slice(self) -> slice:
        return slice(self.q_head_start, self.q_head_start + self.num_heads)

    @property
    def kv_slice(self) -> slice:
        return slice(self.kv_head_start, self.kv_head_start + self.num_kv_heads)


def head_partition(rank: int, world: int, total_num_heads: int, total_num_kv_heads: int) -> HeadPartition:
    """Per-rank head counts exactly as reference qwen3.py:85-98 computes them.

    world <= total_kv_heads: KV heads are partitioned (world must divide them);
    world >  total_kv_heads: KV heads are replicated (total_kv_heads must divide world),
    each rank keeping one KV head and total_num_heads/world query heads of it.
    """
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    if total_num_heads % world != 0:
        raise ValueError("total_num_heads %d not divisible by world %d" % (total_num_heads, world))
    if total_num_heads % total_num_kv_heads != 0:
        raise ValueError("query heads %d not a multiple of KV heads %d" % (total_num_heads, total_num_kv_heads))
    hq = total_num_heads // world
    if total_num_kv_heads >= world:
        if total_num_kv_heads % world != 0:
            raise ValueError("total_num_kv_heads %d not divisible by world %d" % (total_num_kv_heads, world))
        hkv, rep = total_num_kv_heads // world, 1
        kv_start = rank * hkv
    else:
        if world % total_num_kv_heads != 0:
            raise ValueError("world %d not a multiple of total_num_kv_heads %d" % (world, total_num_kv_heads))
        hkv, rep = 1, world // total_num_kv_heads
        kv_start = rank // rep
    return HeadPartition(rank, world, hq, hkv, rank * hq, kv_start, rep)


def shard_heads(x: torch.Tensor, part: HeadPartition, kind: str) -> torch.Tensor:
    """Slice a global [T, H, D] tensor down to this rank's heads (kind = 'q' or 'kv')."""
    sl = part.q_slice if kind == "q" else part.kv_slice
    return x[:, sl, :]


def all_gather_heads(out_shard: torch.Tensor, gathered: Optional[torch.Tensor] = None, group=None,
                     async_op: bool = False):
    """All-gather [T, Hq/P, D] output shards into [P, T, Hq/P, D] (rank-major, one contiguous
    message per rank: what ncclAllGather produces).  Returns (gathered, work_or_None)."""
    world = dist.get_world_size(group)
    shard = out_shard.contiguous()
    if gathered is None:
        gathered = torch.empty((world,) + tuple(shard.shape), dtype=shard.dtype, device=shard.device)
    work = dist.all_gather_into_tensor(gathered.view(-1), shard.view(-1), group=group, async_op=async_op)
    return gathered, work


def heads_from_gathered(gathered: torch.Tensor) -> torch.Tensor:
    """[P, T, Hq/P, D] -> the global [T, Hq, D] view order (a permuted view, no copy)."""
    p, t, h, d = gathered.shape
    return gathered.permute(1, 0, 2, 3).reshape(t, p * h, d)


class OverlappedGather:
    """Per-layer output all-gather that runs behind the next layer's attention.

    Holds ``depth`` (output shard, gathered) buffer pairs.  ``buffer(i)`` hands layer i its output
    buffer, first waiting for the gather that last used it; ``launch(i)`` starts the asynchronous
    all-gather of that buffer (``async_op=True``: on NCCL it runs on the communicator's stream while
    the caller's stream goes on with layer i+1); ``finish()`` waits for everything in flight.  With
    depth 2 the only exposed gather is the last layer's (and whatever is longer than a layer).
    """

    def __init__(self, shard_shape, dtype, device, depth: int = 2, group=None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.outs = [torch.empty(shard_shape, dtype=dtype, device=device) for _ in range(depth)]
        self.gathered = [torch.empty((self.world,) + tuple(shard_shape), dtype=dtype, device=device)
                         for _ in range(depth)]
        self.pending = [None] * depth

    def buffer(self, i: int) -> torch.Tensor:
        b = i % len(self.outs)
        if self.pending[b] is not None:       # the previous gather out of this buffer must be done
            self.pending[b].wait()
            self.pending[b] = None
        return self.outs[b]

    def launch(self, i: int) -> None:
        b = i % len(self.outs)
        self.pending[b] = dist.all_gather_into_tensor(self.gathered[b].view(-1), self.outs[b].view(-1),
                                                      group=self.group, async_op=True)

    def result(self, i: int) -> torch.Tensor:
        """The gathered [P, ...] tensor of layer i (waits for its gather)."""
        b = i % len(self.outs)
        if self.pending[b] is not None:
            self.pending[b].wait()
            self.pending[b] = None
        return self.gathered[b]

    def finish(self) -> None:
        for b, w in enumerate(self.pending):
            if w is not None:
                w.wait()
                self.pending[b] = None


class PeerGather:
    """The same per-layer output all-gather as OverlappedGather, without NCCL: every rank PUSHES its
    [T, Hq/P * D] shard into the other ranks' gather buffers over NVLink / NVSwitch peer memory.

    The gather buffers [P, T, Hq/P * D] are symmetric-memory allocations (torch.distributed._symmetric_memory:
    cuMem peer mappings), the attention kernels write a layer's output straight into slot `rank` of this rank's
    buffer (no staging copy), and `launch(i)` issues P - 1 peer copies on a side stream -- copy engines, not SMs,
    so the next layer's kernels keep all 148 SMs and the transfer overlaps them -- followed by a device-side
    barrier, so that "gather i done" means every rank's buffer holds all P shards.
    Measured against ncclAllGather at 8 x B200: profiles/r02_peer_gather.md.
    Interface = OverlappedGather's (buffer / launch / result / finish).
    """

    def __init__(self, shard_shape, dtype, device, depth: int = 2, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.shape = (self.world,) + tuple(shard_shape)
        self.dtype = dtype
        self.gathered, self.handles = [], []
        for _ in range(depth):
            t = symm_mem.empty(self.shape, dtype=dtype, device=device)
            self.handles.append(symm_mem.rendezvous(t, self.group.group_name))
            self.gathered.append(t)
        self.outs = [g[self.rank] for g in self.gathered]          # the kernels' output buffers
        self.peer_slots = [[h.get_buffer(p, self.shape, dtype)[self.rank] if p != self.rank else None
                            for p in range(self.world)] for h in self.handles]
        self.stream = torch.cuda.Stream(device=device)
        self.ready = [torch.cuda.Event() for _ in range(depth)]    # layer output written (compute stream)
        self.done = [None] * depth                                 # pushes + barrier finished (side stream)

    def buffer(self, i: int) -> torch.Tensor:
        b = i % len(self.outs)
        if self.done[b] is not None:          # our pushes out of this buffer must have read it
            torch.cuda.current_stream().wait_event(self.done[b])
            self.done[b] = None
        return self.outs[b]

    def launch(self, i: int) -> None:
        b = i % len(self.outs)
        self.ready[b].record()
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(self.ready[b])
            # start with the right-hand neighbour so that at any moment the P ranks target P different peers
            for k in range(1, self.world):
                p = (self.rank + k) % self.world
                self.peer_slots[b][p].copy_(self.outs[b], non_blocking=True)
            self.handles[b].barrier(channel=b)
            ev = torch.cuda.Event()
            ev.record()
        self.done[b] = ev

    def result(self, i: int) -> torch.Tensor:
        b = i % len(self.outs)
        if self.done[b] is not None:
            torch.cuda.current_stream().wait_event(self.done[b])
        return self.gathered[b]

    def finish(self) -> None:
        for b, ev in enumerate(self.done):
            if ev is not None:
                torch.cuda.current_stream().wait_event(ev)
                self.done[b] = None


def make_gather(kind: str, shard_shape, dtype, device, depth: int = 2, group=None):
    """'peer' (NVLink peer-memory push, falls back to NCCL when symmetric memory is unavailable) or 'nccl'.
    Returns (gather object, kind actually used)."""
    if kind == "peer":
        try:
            return PeerGather(shard_shape, dtype, device, depth, group), "peer"
        except Exception as e:      # noqa: BLE001 -- no symmetric memory on this box / build
            import sys
            print("PeerGather unavailable (%s: %s); using ncclAllGather" % (type(e).__name__, e), file=sys.stderr)
    return OverlappedGather(shard_shape, dtype, device, depth, group), "nccl"


class RowParallelOProj:
    """o_proj under head sharding with the reduce-scatter fused into the GEMM.

    Reference: `output, _ = self.o_proj(attn_output)` (vllm/model_executor/models/qwen3.py:190) is a
    RowParallelLinear (vllm/model_executor/layers/linear.py:1197-1359): every rank multiplies its
    [T, Hq/P * D] attention shard with its [hidden, Hq/P * D] weight shard and the partial results are
    all-reduced (:1359).  Here rank r owns the token rows [r * T/P, (r + 1) * T/P): the GEMM's epilogue
    TMA-stores every partial tile straight into the owner's staging buffer [P, T/P, hidden] (slot = source
    rank) over NVLink peer memory -- the transfer rides behind the math tile by tile -- a device barrier
    follows, and the owner adds its P partials in fp32 (ops.sum_partials).  `forward` returns the rank's
    [T/P, hidden] rows (reduce-scatter); `all_gather=True` pushes them to every rank (the all-reduce value).
    Needs T % (128 * P) == 0 (token rows are padded by the caller otherwise)."""

    def __init__(self, weight_shard: torch.Tensor, tokens: int, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.weight = weight_shard
        self.hidden = weight_shard.shape[0]
        if tokens % (128 * self.world) != 0:
            raise ValueError("tokens (%d) must be a multiple of 128 * world (%d)" % (tokens, 128 * self.world))
        self.tokens, self.rows = tokens, tokens // self.world
        self.staging = symm_mem.empty((self.world, self.rows, self.hidden), dtype=torch.bfloat16, device=device)
        self.handle = symm_mem.rendezvous(self.staging, self.group.group_name)
        self.dest_ptrs = [self.handle.get_buffer(p, self.staging.shape, torch.bfloat16).data_ptr()
                          for p in range(self.world)]
        self.full = None

    def forward(self, attn_shard: torch.Tensor, all_gather: bool = False) -> torch.Tensor:
        from . import ops
        assert attn_shard.shape[0] == self.tokens
        self.handle.barrier(channel=0)                 # every rank is done reading its staging buffer
        ops.linear_scatter(attn_shard, self.weight, self.dest_ptrs, self.rank, self.world, self.rows)
        self.handle.barrier(channel=1)                 # every rank's partial tiles have landed everywhere
        mine = ops.sum_partials(self.staging)
        if not all_gather:
            return mine
        if self.full is None:
            self.full = PeerGather((self.rows, self.hidden), torch.bfloat16, mine.device, depth=1, group=self.group)
        self.full.buffer(0).copy_(mine)
        self.full.launch(0)
        return self.full.result(0).view(self.tokens, self.hidden)
