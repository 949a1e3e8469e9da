"""Column-parallel sharding of the W4A4 MixedQuant linear over the GPUs of one box.

The path shards along N (out_features): rows of the packed weight (codes, group scales, salient slice) and the bias
are independent per output column, while every rank needs the full quantised activation (K is never split, so the
permutation and the per-group scales are untouched).  Rank r owns output columns [r*S, min((r+1)*S, N)) with
S = ceil(N / world) rounded up to a multiple of 8 (16-byte rows for the epilogue's vector stores).

The only collective is the all-gather that reassembles ``y`` where the consumer needs the full width
(``gather_output=True``); layers whose consumer is column-local (q/k/v -> per-head attention, gate/up -> SiLU*mul)
keep their output sharded.  The reference has no multi-GPU code at all (its only touchpoint is
``device_map="auto"``, ppl_eval.py:70 / run_experiments.py:147), so there is nothing to mirror here; the plumbing is
``torch.distributed`` (NCCL on the GPUs, gloo in the CPU tests -- this file never launches a kernel itself).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_size(n_out: int, world: int) -> int:
    """Rows per rank: ceil(N / world) rounded up to a multiple of 8."""
    if world < 1:
        raise ValueError(f"world size must be >= 1, got {world}")
    s = (n_out + world - 1) // world
    return (s + 7) // 8 * 8


def shard_range(n_out: int, world: int, rank: int) -> Tuple[int, int]:
    """[lo, hi) of the output columns owned by ``rank`` (empty for trailing ranks when N is small)."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of size {world}")
    s = shard_size(n_out, world)
    lo = min(rank * s, n_out)
    return lo, min(lo + s, n_out)


def shard_rows(t: Optional[torch.Tensor], n_out: int, world: int, rank: int, dim: int = 0) -> Optional[torch.Tensor]:
    """The slice of a per-output-channel tensor (dimension ``dim`` has >= N entries) that ``rank`` owns."""
    if t is None:
        return None
    lo, hi = shard_range(n_out, world, rank)
    return t.narrow(dim, lo, hi - lo).contiguous()


def shard_scales(scales: Optional[torch.Tensor], n_out: int, world: int, rank: int) -> Optional[torch.Tensor]:
    """Group-major scale table [G, ld >= N] -> [G, ld_shard] with ld_shard a multiple of 4 (TMA row alignment)."""
    if scales is None:
        return None
    lo, hi = shard_range(n_out, world, rank)
    n = hi - lo
    ld = max((n + 3) // 4 * 4, 4)
    out = torch.zeros((scales.shape[0], ld), dtype=scales.dtype, device=scales.device)
    out[:, :n] = scales[:, lo:hi]
    return out


def gather_columns(y_local: torch.Tensor, n_out: int, group=None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """All-gather column shards [M, n_r] (n_r <= S) into the full [M, N].

    Every rank contributes a [M, S] block (zero padded on the right when its shard is narrower); the gathered
    [world, M, S] buffer is then laid out as [M, world*S] and trimmed to N.  Bytes received per rank:
    (world-1) * M * S * itemsize."""
    world = dist.get_world_size(group)
    if world == 1:
        return y_local[:, :n_out]
    S = shard_size(n_out, world)
    M = y_local.shape[0]
    if y_local.shape[1] != S or not y_local.is_contiguous():
        blk = torch.zeros((M, S), dtype=y_local.dtype, device=y_local.device)
        blk[:, : y_local.shape[1]] = y_local
    else:
        blk = y_local
    buf = torch.empty((world, M, S), dtype=y_local.dtype, device=y_local.device)
    dist.all_gather_into_tensor(buf.view(world * M, S), blk, group=group)
    full = buf.permute(1, 0, 2).reshape(M, world * S)
    if out is not None:
        out.copy_(full[:, :n_out])
        return out
    return full[:, :n_out]


class BlockedColumns:
    """The output of a column-parallel layer after the all-gather, left in the layout the collective produces:
    ``buf`` [world, M, S], rank r's column block at buf[r] (columns [r*S, r*S + S) of the full [M, n_cols] tensor; the
    last block is zero padded beyond n_cols).  The next layer's row quantiser reads this layout in place
    (ops.quant_rows -> mq_quant_rows_blocked: its bulk copies reassemble each row in shared memory), so the gathered
    activations are never re-laid-out in HBM; ``dense()`` makes the [M, n_cols] tensor for any other consumer.

    The gather may still be in flight: ``wait()`` (called by both consumers) makes the current stream wait for it."""

    def __init__(self, buf: torch.Tensor, n_cols: int, work=None):
        if buf.dim() != 3 or not buf.is_contiguous():
            raise ValueError("BlockedColumns wants a contiguous [world, M, S] buffer")
        if not (buf.shape[0] - 1) * buf.shape[2] < n_cols <= buf.shape[0] * buf.shape[2]:
            raise ValueError(f"{buf.shape[0]} blocks of {buf.shape[2]} columns do not tile {n_cols} columns")
        self.buf, self.n_cols, self._work = buf, int(n_cols), work

    # what W4A4Linear.forward reads of its input
    @property
    def shape(self):
        return torch.Size((self.buf.shape[1], self.n_cols))

    @property
    def dtype(self):
        return self.buf.dtype

    @property
    def device(self):
        return self.buf.device

    def dim(self):
        return 2

    # identity of the underlying storage (the activation cache of fake_quant keys on these)
    def data_ptr(self):
        return self.buf.data_ptr()

    def stride(self):
        return self.buf.stride()

    def is_inference(self):
        return self.buf.is_inference()

    @property
    def _version(self):
        return self.buf._version

    def wait(self):
        if self._work is not None:
            self._work.wait()          # stream-level: the current stream waits for the collective's stream
            self._work = None
        return self

    def dense(self) -> torch.Tensor:
        self.wait()
        world, M, S = self.buf.shape
        if world == 1:
            return self.buf[0, :, : self.n_cols]
        return self.buf.permute(1, 0, 2).reshape(M, world * S)[:, : self.n_cols]


ASYNC_GATHER = True     # layers launch their blocked gathers asynchronously (False: serialised, for the comm split)
COMM_EVENTS = None      # set to a list to collect (start_event, end_event, bytes_received) per gather (NCCL stream)


def gather_buffer(M: int, n_out: int, world: int, dtype, device) -> torch.Tensor:
    """[world, M, S] destination of gather_columns_blocked; a layer writes its own block (buf[rank]) in place."""
    return torch.empty((world, M, shard_size(n_out, world)), dtype=dtype, device=device)


def gather_columns_blocked(y_local: torch.Tensor, n_out: int, group=None, buf: Optional[torch.Tensor] = None,
                           async_op: Optional[bool] = None) -> BlockedColumns:
    """All-gather column shards [M, n_r] into a BlockedColumns, without the re-layout copy of gather_columns.

    If ``y_local`` already is ``buf[rank]`` (the GEMM wrote its block in place) the collective runs in place; otherwise
    the shard is copied into its block first.  With ``async_op`` the collective runs on the process group's own
    stream and the caller's stream only waits for it when the result is consumed (BlockedColumns.wait), so the next
    independent GEMM overlaps the transfer.  Bytes received per rank: (world - 1) * M * S * itemsize."""
    if async_op is None:
        async_op = ASYNC_GATHER
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    M = y_local.shape[0]
    S = shard_size(n_out, world)
    if buf is None:
        buf = gather_buffer(M, n_out, world, y_local.dtype, y_local.device)
    mine = buf[rank]
    if y_local.data_ptr() != mine.data_ptr() or y_local.stride() != mine.stride():
        n = y_local.shape[1]
        mine[:, :n].copy_(y_local)
        if n < S:
            mine[:, n:].zero_()
    if world == 1:
        return BlockedColumns(buf, n_out)
    # NCCL gathers in place (send buffer = its own slot of the receive buffer); gloo (CPU tests) wants them disjoint
    src = mine if y_local.is_cuda else mine.clone()
    ev = None
    if COMM_EVENTS is not None and y_local.is_cuda:
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record()
    work = dist.all_gather_into_tensor(buf.view(world * M, S), src, group=group, async_op=async_op)
    if ev is not None:
        # timed on the caller's stream around a blocking wait: only meaningful with async_op=False (bench comm split)
        if not async_op:
            ev[1].record()
            COMM_EVENTS.append((ev[0], ev[1], (world - 1) * M * S * y_local.element_size()))
    return BlockedColumns(buf, n_out, work if async_op else None)


def allreduce_colmax(key: torch.Tensor, group=None) -> torch.Tensor:
    """max over data-parallel ranks of a per-column statistic: what the literal dynamic sort (fake_quant.py:113, a
    reduction over ALL tokens of the batch) needs when the tokens of one batch are spread over ranks."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(key, op=dist.ReduceOp.MAX, group=group)
    return key


def split_sequences(n_seq: int, world: int, rank: int) -> Tuple[int, int]:
    """Data-parallel prefill: contiguous block of sequences for ``rank`` (sizes differ by at most one)."""
    base, rem = divmod(n_seq, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def layer_plan_llama() -> List[Tuple[str, bool]]:
    """(linear name, gather_output) for one column-parallel Llama decoder layer: outputs are gathered only where the
    consumer needs the full width: v->o and up->down feed K of the next linear, o/down feed the
    residual stream; q/k feed per-head attention and gate feeds SiLU*mul, both column-local."""
    return [("q_proj", False), ("k_proj", False), ("v_proj", True), ("o_proj", True),
            ("gate_proj", False), ("up_proj", True), ("down_proj", True)]
