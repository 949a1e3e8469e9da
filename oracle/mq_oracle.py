"""CPU oracle for the W4A4 MixedQuant linear path  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may
import this module.  The product path (``smoothquant_mixedprecision_b200``) never imports it and has no CPU
fallback.

It restates, out of place and with the hidden intermediate results (integer codes, group scales, permutation)
exposed, the algorithm of the reference's ``smoothquant/fake_quant.py`` and ``smoothquant/smooth.py``.  Every
function names the reference lines it follows (paths relative to ``/root/reference``).

Parity status: PINNED.  ``tests/golden/make_golden.py`` imports the real reference in the build container, runs
its quantisers / ``W4A4Linear`` / ``smooth_lm`` on seeded inputs and stores inputs + outputs under
``tests/golden/*.pt``; ``tests/test_oracle_golden.py`` checks this restatement against those files bit for bit.
Unpinned by construction (no reference code exists, README prose only): the ``mean3std`` and ``argmax`` sort
keys -- see ``sort_key``.

Arithmetic model: everything is done in the tensor's own dtype by torch CPU ops, exactly like the
reference:  scale = clamp(absmax, 1e-5) / qmax ;  q = round_half_even(x / scale) ;  out = q * scale, each op
rounding to the dtype.  ``qmax = 2**(n_bits-1) - 1`` and there is no clamp of the codes (|q| <= qmax holds by
construction).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn.functional as F

__all__ = [
    "GroupQuant", "qmax_of", "group_quant", "per_row_quant", "per_tensor_quant", "stable_argsort",
    "sort_key", "act_quant_group_sort", "weight_quant_group_sort", "select_salient", "nonsalient_columns",
    "w4a4_weight_reference", "w4a4_forward_reference", "fused_partition", "fused_layer_oracle",
    "smooth_scales", "int4_pack", "int4_unpack",
]


def qmax_of(n_bits: int) -> int:
    # fake_quant.py:13, :46, :94, :138, :189   q_max = 2 ** (n_bits - 1) - 1
    return 2 ** (n_bits - 1) - 1


@dataclass
class GroupQuant:
    """Decomposed result of one per-group abs-max quantisation over the last dim of a 2-D tensor."""
    codes: torch.Tensor    # int8  [R, G*g]   integer codes in [-qmax, qmax]; padded tail columns are 0
    scales: torch.Tensor   # dtype [R, G]     group scales, in the input dtype (exact values the reference uses)
    deq: torch.Tensor      # dtype [R, C]     codes*scales in the input dtype, un-padded == reference output
    group_size: int
    cols: int


@torch.no_grad()
def group_quant(t: torch.Tensor, n_bits: int, group_size: int) -> GroupQuant:
    """Natural-order per-group abs-max quantiser.

    Follows quantize_activation_per_group_absmax (fake_quant.py:78-101) and, identically,
    quantize_weight_per_group_absmax (fake_quant.py:30-53): zero-pad the last dim to a multiple of
    ``group_size`` (:84-86), view as [R, G, g] (:89), scale = absmax.clamp_(1e-5).div_(qmax) (:92-94),
    t.div_(scale).round_().mul_(scale) (:97), drop the padding (:100).  Out of place: the input is not touched.
    """
    assert t.dim() == 2
    R, C = t.shape
    g = group_size
    G = (C + g - 1) // g
    tp = F.pad(t, (0, G * g - C)) if G * g != C else t.clone()
    tg = tp.view(R, G, g)
    scales = tg.abs().max(dim=-1, keepdim=True)[0]
    scales = scales.clamp(min=1e-5).div(qmax_of(n_bits))          # rounded to t.dtype, like clamp_/div_
    q = tg.div(scales).round()                                    # quotient rounded to t.dtype BEFORE round (F4)
    deq = q.mul(scales).view(R, G * g)[:, :C].contiguous()
    codes = q.view(R, G * g).to(torch.float32).to(torch.int8)
    return GroupQuant(codes=codes, scales=scales.view(R, G), deq=deq, group_size=g, cols=C)


@torch.no_grad()
def per_row_quant(t: torch.Tensor, n_bits: int) -> GroupQuant:
    """One group spanning the whole row: quantize_activation_per_token_absmax (fake_quant.py:57-64) and
    quantize_weight_per_channel_absmax (fake_quant.py:10-16).  Out of place (the reference mutates)."""
    return group_quant(t.reshape(-1, t.shape[-1]), n_bits, t.shape[-1])


@torch.no_grad()
def per_tensor_quant(t: torch.Tensor, n_bits: int) -> torch.Tensor:
    """quantize_activation_per_tensor_absmax (fake_quant.py:68-75) / quantize_weight_per_tensor_absmax (:20-26)."""
    s = t.abs().max().clamp(min=1e-5).div(qmax_of(n_bits))
    return t.div(s).round().mul(s)


def stable_argsort(key: torch.Tensor, descending: bool = False) -> torch.Tensor:
    """Canonical ordering used by the new implementation: ties broken by ascending index.
    The reference calls torch.argsort without stable=True (fake_quant.py:116, :167, :268) whose tie order is
    implementation-defined; on tie-free keys both agree."""
    return torch.argsort(key, stable=True, descending=descending)


@torch.no_grad()
def sort_key(x2d: torch.Tensor, kind: str = "max") -> torch.Tensor:
    """Column statistic the channels are sorted by (ascending).

    "max"      : x.abs().max(dim=0)                      fake_quant.py:113 / :164   -- PINNED by reference code
    "mean3std" : mean|x| + 3*std|x| over rows (unbiased) README.md:36               -- parity UNPINNED (prose only)
    "argmax"   : index of the first row attaining max|x| README.md:37               -- parity UNPINNED (prose only)
    "none"     : identity order
    """
    a = x2d.abs()
    if kind == "max":
        return a.max(dim=0).values
    if kind == "mean3std":
        af = a.float()
        mean = af.mean(dim=0)
        if af.shape[0] > 1:
            var = (af - mean).pow(2).sum(dim=0) / (af.shape[0] - 1)
        else:
            var = torch.zeros_like(mean)
        return mean + 3.0 * var.sqrt()
    if kind == "argmax":
        return a.float().argmax(dim=0).to(torch.float32)
    if kind == "none":
        return torch.arange(x2d.shape[1], dtype=torch.float32)
    raise ValueError(f"Invalid sort key: {kind}")


@torch.no_grad()
def act_quant_group_sort(t: torch.Tensor, n_bits: int, group_size: int,
                         perm: Optional[torch.Tensor] = None):
    """quantize_activation_per_group_absmax_sort (fake_quant.py:105-154) with the permutation made explicit.

    key = |t|.max over ALL rows (:113); perm = argsort(key) ascending (:116); gather columns (:120); pad,
    group, quantise (:124-142); un-pad (:145-147); inverse permutation (:150-151).
    Returns (dequantised tensor in original column order and shape, GroupQuant of the permuted matrix, perm).
    """
    shape = t.shape
    t2 = t.reshape(-1, shape[-1])
    if perm is None:
        perm = stable_argsort(sort_key(t2, "max"))
    gq = group_quant(t2[:, perm], n_bits, group_size)
    inv = torch.argsort(perm)
    return gq.deq[:, inv].reshape(shape), gq, perm


@torch.no_grad()
def weight_quant_group_sort(w: torch.Tensor, n_bits: int, group_size: int,
                            perm: Optional[torch.Tensor] = None):
    """quantize_weight_per_group_absmax_sort (fake_quant.py:157-207): same algorithm on w [N, K]; the key is the
    per-input-column abs-max over the N output rows (:164)."""
    return act_quant_group_sort(w, n_bits, group_size, perm)


def select_salient(importance: Optional[torch.Tensor], salient_prop: float) -> Optional[torch.Tensor]:
    """W4A4Linear.__init__ (fake_quant.py:265-270): argsort(importance, descending)[:max(1, int(p*K))].
    Canonical stable order on ties."""
    if importance is None or not salient_prop or salient_prop <= 0:
        return None
    order = stable_argsort(importance, descending=True)
    n = max(1, int(salient_prop * len(order)))
    return order[:n]


def nonsalient_columns(K: int, salient: Optional[torch.Tensor]) -> torch.Tensor:
    """Ascending list of the columns kept by the boolean mask of fake_quant.py:293-294."""
    mask = torch.ones(K, dtype=torch.bool)
    if salient is not None:
        mask[salient] = False
    return torch.nonzero(mask).flatten()


@torch.no_grad()
def w4a4_weight_reference(w: torch.Tensor, salient: Optional[torch.Tensor], n_bits: int, group_size: int,
                          weight_quant: str = "per_group", perm: Optional[torch.Tensor] = None) -> torch.Tensor:
    """W4A4Linear.from_float weight path (fake_quant.py:347-365): keep a copy of the salient columns (:347),
    fake-quantise the WHOLE weight (:348-361), write the FP originals back (:363-365)."""
    if weight_quant == "per_group":
        wq, _, _ = weight_quant_group_sort(w, n_bits, group_size, perm)
    elif weight_quant == "per_channel":
        wq = per_row_quant(w, n_bits).deq
    elif weight_quant == "per_tensor":
        wq = per_tensor_quant(w, n_bits)
    else:
        raise ValueError(f"Invalid weight_quant: {weight_quant}")
    wq = wq.clone()
    if salient is not None:
        wq[:, salient] = w[:, salient]
    return wq


@torch.no_grad()
def w4a4_forward_reference(x: torch.Tensor, weight_q: torch.Tensor, bias: Optional[torch.Tensor],
                           salient: Optional[torch.Tensor], n_bits: int, group_size: int,
                           act_quant: str = "per_group", quantize_output: bool = False,
                           act_perm: Optional[torch.Tensor] = None, out_perm: Optional[torch.Tensor] = None,
                           return_qx: bool = False):
    """W4A4Linear.forward (fake_quant.py:279-322), literal semantics ("reference" mode).

    * 3-D input flattened to [M, K] (:281-287)
    * non-salient columns compacted with a boolean mask, quantised with the dynamic max-sort group quantiser,
      scattered back (:291-301); without salient channels the whole row is quantised (:304)
    * y = F.linear(q_x, weight, bias) (:306)
    * optional output quantiser that re-uses the K-sized mask on y's columns (:309-316)
    """
    shape = x.shape
    if x.dim() == 3:
        x2 = x.reshape(-1, shape[-1])
    elif x.dim() == 2:
        x2 = x
    else:
        raise ValueError(f"Unsupported input shape: {shape}")

    def quant(t, perm):
        if act_quant == "per_group":
            return act_quant_group_sort(t, n_bits, group_size, perm)[0]
        if act_quant == "per_token":
            return per_row_quant(t, n_bits).deq
        if act_quant == "per_tensor":
            return per_tensor_quant(t, n_bits)
        raise ValueError(f"Invalid act_quant: {act_quant}")

    K = x2.shape[-1]
    if salient is not None:
        keep = nonsalient_columns(K, salient)
        q_x = x2.clone()
        if keep.numel() > 0:
            q_x[:, keep] = quant(x2[:, keep], act_perm)
    else:
        q_x = quant(x2, act_perm)
    y = F.linear(q_x, weight_q, bias)
    if quantize_output:
        if salient is not None:
            q_y = y.clone()
            if keep.numel() > 0:
                q_y[:, keep] = quant(y[:, keep], out_perm)      # IndexError when N != K, like the reference
        else:
            q_y = quant(y, out_perm)
    else:
        q_y = y
    out = q_y.view(shape[0], shape[1], -1) if x.dim() == 3 else q_y
    return (out, q_x) if return_qx else out


@dataclass
class FusedPartition:
    """The ONE static K-partition shared by both operands in mode="fused" (SURVEY 8(a'))."""
    salient: torch.Tensor     # int64 [Ks]  salient columns, ascending
    perm: torch.Tensor        # int64 [Kn]  original column index of permuted position j (non-salient only)
    group_size: int

    @property
    def groups(self) -> int:
        return (self.perm.numel() + self.group_size - 1) // self.group_size


def fused_partition(K: int, salient: Optional[torch.Tensor], key: Optional[torch.Tensor],
                    group_size: int) -> FusedPartition:
    """Static permutation: non-salient columns (ascending original index) stably sorted by ``key`` ascending --
    the same ordering rule as fake_quant.py:113-120, with an offline key instead of the per-call one."""
    keep = nonsalient_columns(K, salient)
    if key is None:
        perm = keep
    else:
        perm = keep[stable_argsort(key[keep].float())]
    sal = torch.sort(salient).values if salient is not None else torch.empty(0, dtype=torch.long)
    return FusedPartition(salient=sal, perm=perm, group_size=group_size)


@torch.no_grad()
def fused_layer_oracle(x2: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], part: FusedPartition,
                       n_bits: int = 4, exact_int: bool = True):
    """Composed oracle for mode="fused":
        y = a2(x[:, perm]) @ a4(W[:, perm])^T + x[:, sal] @ W[:, sal]^T + b
    with a2 = quantize_activation_per_group_absmax (fake_quant.py:78-101) and a4 =
    quantize_weight_per_group_absmax (fake_quant.py:30-53).  Evaluated in float64 from the integer codes and
    scales (exact_int) so that it is the mathematically exact value the fused GEMM approximates.
    Returns (y float64 [M, N], act GroupQuant, weight GroupQuant)."""
    g = part.group_size
    aq = group_quant(x2[:, part.perm], n_bits, g)
    wq = group_quant(w[:, part.perm], n_bits, g)
    M, N, G = x2.shape[0], w.shape[0], part.groups
    y = torch.zeros(M, N, dtype=torch.float64)
    if exact_int:
        ac = aq.codes.view(M, G, g).to(torch.float64)
        wc = wq.codes.view(N, G, g).to(torch.float64)
        for gi in range(G):
            p = ac[:, gi, :] @ wc[:, gi, :].t()
            y += p * (aq.scales[:, gi].double()[:, None] * wq.scales[:, gi].double()[None, :])
    else:
        y += aq.deq.double() @ wq.deq.double().t()
    if part.salient.numel() > 0:
        y += x2[:, part.salient].double() @ w[:, part.salient].double().t()
    if bias is not None:
        y += bias.double().view(1, -1)
    return y, aq, wq


@torch.no_grad()
def smooth_scales(act_scales: torch.Tensor, fc_weights, alpha: float = 0.5) -> torch.Tensor:
    """smooth_ln_fcs / smooth_ln_fcs_llama_like (smooth.py:26-39, :56-67):
    s = clamp(act^alpha / clamp(max_fcs max_n |W|, 1e-5)^(1-alpha), 1e-5) in the weight dtype."""
    dtype = fc_weights[0].dtype
    a = act_scales.to(dtype=dtype)
    ws = torch.cat([w.abs().max(dim=0, keepdim=True)[0] for w in fc_weights], dim=0)
    ws = ws.max(dim=0)[0].clamp(min=1e-5)
    return (a.pow(alpha) / ws.pow(1 - alpha)).clamp(min=1e-5).to(dtype)


def int4_pack(codes: torch.Tensor) -> torch.Tensor:
    """Two's-complement nibbles, even column in the low nibble: the storage format of the weight packer."""
    assert codes.dtype == torch.int8 and codes.shape[-1] % 2 == 0
    u = codes.to(torch.int16) & 0xF
    return (u[..., 0::2] | (u[..., 1::2] << 4)).to(torch.uint8)


def int4_unpack(packed: torch.Tensor) -> torch.Tensor:
    p = packed.to(torch.int16)
    lo = ((p & 0xF) ^ 8) - 8
    hi = (((p >> 4) & 0xF) ^ 8) - 8
    return torch.stack([lo, hi], dim=-1).reshape(*packed.shape[:-1], -1).to(torch.int8)
