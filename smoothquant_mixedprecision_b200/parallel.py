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
