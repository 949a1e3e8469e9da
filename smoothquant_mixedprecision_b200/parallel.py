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


class PeerGather:
    """SM-free replacement for the NCCL all-gather of gather_columns_blocked on one NVLink box.

    Gather buffers [world, M, S] and a flag block live in peer-mapped memory (torch.distributed._symmetric_memory =
    cuMem allocations mapped into every rank of the group); the transfer is mq_peer_gather (include/mq.h): copy-engine
    pulls over NVLink ordered by stream memory operations, issued on a side stream, so it runs underneath the GEMMs
    of the next row chunk instead of queueing behind them for SMs.

    Buffers come from a ring of ``ring`` buffers per (M, S, dtype): the BlockedColumns a layer returns is valid until
    ``ring`` more gathers of the same shape have been issued (``dense()`` copies).  Every rank must issue the same
    sequence of gathers."""

    def __init__(self, group=None, ring: int = 4, lanes: Optional[int] = None, mode: str = "pull"):
        import ctypes as C
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        self._C, self._symm, self._lib = C, symm, _lib
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.ring = max(2, int(ring))
        # pull: every rank copies its peers' blocks in (mq_peer_gather).  push: every rank copies its block out to every
        # peer (mq_peer_push).  mcast: push with ONE copy to the multicast address of the buffers (NVSwitch replicates).
        if mode not in ("pull", "push", "mcast"):
            raise ValueError(f"Invalid peer gather mode: {mode}")
        self.mode = mode
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.seq = 0
        self.rings = {}                       # (M, S, dtype) -> {"slots": [...], "next": i}
        # copy streams pulling in parallel; measured on 8 B200s: 1, 2, 4, 7 lanes move 326 / 350 / 311 / 304 GB/s per
        # rank -- the pulls are not copy-engine bound, so one lane (fewest host calls) is the default
        self.lanes = max(1, min(self.world - 1, 1 if lanes is None else int(lanes)))
        self.streams = [torch.cuda.Stream(device=self.dev, priority=-1) for _ in range(self.lanes)]
        self._lane_ptrs = (C.c_uint64 * self.lanes)(*[s.cuda_stream for s in self.streams])
        self.max_buffers = 124                # rows of the consumed[buffer id] table behind the 4 fixed flag rows
        self.n_buffers = 0
        self.flags = symm.empty((4 + self.max_buffers) * self.world, dtype=torch.int32, device=self.dev)
        self.flags.zero_()
        self._fh = symm.rendezvous(self.flags, self.group)
        self._flag_ptrs = (C.c_uint64 * self.world)(*[int(p) for p in self._fh.buffer_ptrs])
        torch.cuda.synchronize(self.dev)
        self._fh.barrier()                    # nobody publishes a flag before everyone has zeroed theirs
        torch.cuda.synchronize(self.dev)

    def _slot(self, M, S, dtype):
        key = (M, S, dtype)
        ring = self.rings.get(key)
        if ring is None:
            ring = self.rings[key] = {"slots": [], "next": 0}
        if len(ring["slots"]) < self.ring:    # collective: every rank allocates the same buffers in the same order
            buf = self._symm.empty((self.world, M, S), dtype=dtype, device=self.dev)
            h = self._symm.rendezvous(buf, self.group)
            ptrs = (self._C.c_uint64 * self.world)(*[int(p) for p in h.buffer_ptrs])
            mc = 0
            if self.mode == "mcast":
                if not h.has_multicast_support or not int(h.multicast_ptr):
                    raise RuntimeError("peer gather mode 'mcast': no multicast mapping for the gather buffer")
                mc = int(h.multicast_ptr)
            if self.n_buffers >= self.max_buffers:
                raise RuntimeError(f"peer gather: more than {self.max_buffers} ring buffers (distinct gather shapes)")
            slot = {"buf": buf, "h": h, "ptrs": ptrs, "last": 0, "prev": 0, "mc": mc, "id": self.n_buffers}
            self.n_buffers += 1
            ring["slots"].append(slot)
            return slot
        slot = ring["slots"][ring["next"]]
        ring["next"] = (ring["next"] + 1) % self.ring
        return slot

    def buffer(self, M: int, n_out: int, dtype) -> dict:
        """Next gather buffer of the ring for this shape; the current stream is made to wait until every peer has
        copied this rank's block of the buffer's previous gather out, so the GEMM may overwrite it."""
        slot = self._slot(M, shard_size(n_out, self.world), dtype)
        slot["prev"] = slot["last"]
        if slot["last"]:
            st = torch.cuda.current_stream(self.dev).cuda_stream
            if self.mode == "pull":
                self._lib.check(self._lib.lib.mq_peer_wait_pulled(self._flag_ptrs, self.rank, self.world, slot["last"],
                                                                  st), "mq_peer_wait_pulled")
            else:       # everything queued on this stream so far -- the readers of the buffer's last gather -- is done
                self._lib.check(self._lib.lib.mq_peer_publish_consumed(self._flag_ptrs, self.rank, self.world,
                                                                       slot["last"], slot["id"], st),
                                "mq_peer_publish_consumed")
        return slot

    def gather(self, slot: dict, n_out: int, async_op: bool = True) -> "BlockedColumns":
        """All-gather `slot` (this rank's block already written on the current stream)."""
        buf = slot["buf"]
        cur = torch.cuda.current_stream(self.dev)
        self.seq += 1
        slot["last"] = self.seq
        block = buf.shape[1] * buf.shape[2] * buf.element_size()
        ev = None
        if not async_op and COMM_EVENTS is not None:      # serialised: timed on the caller's stream (bench comm split)
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record(cur)
        start = torch.cuda.Event()
        start.record(cur)
        for s in self.streams:
            s.wait_event(start)
        if self.mode == "pull":
            self._lib.check(self._lib.lib.mq_peer_gather(slot["ptrs"], self._flag_ptrs, self.rank, self.world, block,
                                                         self.seq, self._lane_ptrs, self.lanes), "mq_peer_gather")
            done = []
            for s in self.streams:
                e = torch.cuda.Event()
                e.record(s)
                done.append(e)
            out = BlockedColumns(buf, n_out, _EventWork(done, self.dev))
        else:
            self._lib.check(self._lib.lib.mq_peer_push(slot["ptrs"], self._flag_ptrs, slot["mc"], self.rank, self.world,
                                                       block, self.seq, slot["prev"], slot["id"], self._lane_ptrs,
                                                       self.lanes),
                            "mq_peer_push")
            out = BlockedColumns(buf, n_out, _ArrivedWork(self, self.seq))
        if not async_op:
            out.wait()
            if ev is not None:
                ev[1].record(cur)
                COMM_EVENTS.append((ev[0], ev[1], (self.world - 1) * block))
        return out


class _ArrivedWork:
    """wait() = the current stream waits (stream memory operations, no kernel) until every peer's block of gather `seq`
    has landed in this rank's buffer."""

    def __init__(self, peer, seq):
        self.peer, self.seq = peer, seq

    def wait(self):
        p = self.peer
        st = torch.cuda.current_stream(p.dev).cuda_stream
        p._lib.check(p._lib.lib.mq_peer_wait_arrived(p._flag_ptrs, p.rank, p.world, self.seq, st), "mq_peer_wait_arrived")


class _EventWork:
    """wait() = the current stream waits for an event (the interface BlockedColumns expects of a collective's Work)."""

    def __init__(self, events, dev):
        self.events, self.dev = events, dev

    def wait(self):
        cur = torch.cuda.current_stream(self.dev)
        for e in self.events:
            cur.wait_event(e)


_PEER = {}              # id(group) -> PeerGather
TRANSPORT = "nccl"      # "peer": gathers of blocked-output layers go through PeerGather (set by enable_peer_gather)


def enable_peer_gather(group=None, ring: int = 4, lanes: Optional[int] = None, mode: str = "pull") -> PeerGather:
    """Collective: set up the peer-memory transport for `group` (one NVLink box) and make blocked-output layers use
    it.  Raises if symmetric memory cannot be established; the NCCL path stays available (TRANSPORT = "nccl")."""
    global TRANSPORT
    key = id(group) if group is not None else 0
    if key not in _PEER or _PEER[key].mode != mode:
        _PEER[key] = PeerGather(group, ring, lanes, mode)
    TRANSPORT = "peer"
    return _PEER[key]


def peer_gather_for(group=None) -> Optional[PeerGather]:
    if TRANSPORT != "peer":
        return None
    return _PEER.get(id(group) if group is not None else 0)


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
    """(linear name, gather_output) of the column-parallel plan bench.py times: a BYTE-COUNT PROXY of one decoder
    layer, not a recipe for a real attention block.  The bench chains the 7 linears directly (v -> o, o -> gate/up,
    up -> down), so the four tensors a pure column-parallel layer consumes at full width (attention
    output [M, hidden], o_proj output, MLP intermediate [M, inter], down_proj output) are stood in for by the outputs
    of v, o, up and down; q / k / gate stay column-local.  In a real model q, k and v are sharded over the same heads,
    attention runs on the local heads and it is the ATTENTION OUTPUT that is gathered in front of o_proj (v_proj then
    has gather_output=False), and act(gate) * up is formed on the local columns before it is gathered."""
    return [("q_proj", False), ("k_proj", False), ("v_proj", True), ("o_proj", True),
            ("gate_proj", False), ("up_proj", True), ("down_proj", True)]
