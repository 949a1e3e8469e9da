"""Tensor-level wrappers over the C ABI (include/mq.h).  PyTorch is used for device memory and streams only; every
function launches hand-written sm_100a kernels on the caller's current CUDA stream and never synchronises.

All inputs must be CUDA tensors: there is no CPU fallback.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from ._lib import lib, check as _check

_DTYPES = {torch.float16: _lib.MQ_F16, torch.bfloat16: _lib.MQ_BF16, torch.float32: _lib.MQ_F32}
STAT_KINDS = {"max": _lib.MQ_STAT_ABSMAX, "absmax": _lib.MQ_STAT_ABSMAX, "mean": _lib.MQ_STAT_ABSMEAN,
              "absmean": _lib.MQ_STAT_ABSMEAN, "mean3std": _lib.MQ_STAT_MEAN3STD, "argmax": _lib.MQ_STAT_ARGMAX}


# instrumentation for bench.py: number of kernel-launching C-ABI calls, and (optionally) CUDA-event pairs around every
# mixed GEMM launch on the launching stream.  Neither synchronises.
COUNTERS = {"launch_calls": 0}
GEMM_EVENTS = None          # set to a list to collect (start_event, end_event, ops_int8, ops_f16)
QUANT_EVENTS = None         # set to a list to collect (start_event, end_event, algorithmic_bytes) per mq_quant_rows


def check(rc: int, what: str = "") -> None:
    COUNTERS["launch_calls"] += 1
    _check(rc, what)


def _round_up(v: int, m: int) -> int:
    return (v + m - 1) // m * m


def _dtype_code(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype}: expected float16, bfloat16 or float32") from None


def _need_cuda(*ts) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("smoothquant_mixedprecision_b200 runs on CUDA tensors only (no CPU fallback); "
                               f"got a tensor on {t.device}")


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _rows2d(x: torch.Tensor) -> torch.Tensor:
    """[*, K] -> [R, K] view with unit inner stride (copy only if the layout forces it)."""
    x2 = x.reshape(-1, x.shape[-1])
    if x2.stride(-1) != 1 or (x2.shape[0] > 1 and x2.stride(0) < x2.shape[1]):
        x2 = x2.contiguous()
    return x2


def _ld(x2: torch.Tensor) -> int:
    return x2.stride(0) if x2.shape[0] > 1 else max(x2.shape[1], 1)


def _i32(t: Optional[torch.Tensor], device) -> Optional[torch.Tensor]:
    if t is None:
        return None
    return t.to(device=device, dtype=torch.int32).contiguous()


# ---------------------------------------------------------------------------------------------- K0 / K0b
def colstat(x: torch.Tensor, kind: str = "max") -> torch.Tensor:
    """Per-column statistic over all rows of x [*, K] -> float32 [K]  (fake_quant.py:113, :164; run_experiments.py:60)."""
    _need_cuda(x)
    x2 = _rows2d(x)
    M, K = x2.shape
    out = torch.empty(K, dtype=torch.float32, device=x.device)
    k = STAT_KINDS[kind]
    nbytes = lib.mq_colstat_workspace_bytes(K, k)
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        check(lib.mq_colstat(x2.data_ptr(), M, K, _ld(x2), _dtype_code(x2), k, out.data_ptr(), ws.data_ptr(),
                             ws.numel(), _stream(x)), "mq_colstat")
    return out


def argsort_stable(key: torch.Tensor, subset: Optional[torch.Tensor] = None, descending: bool = False) -> torch.Tensor:
    """Stable argsort of float32 keys (ties by ascending index).  With ``subset`` (ascending int32 column ids) only
    those columns are ordered and their ids are returned (fake_quant.py:116, :167, :268, :293-300)."""
    _need_cuda(key, subset)
    key = key.to(torch.float32).contiguous()
    subset = _i32(subset, key.device)
    n = key.numel() if subset is None else subset.numel()
    out = torch.empty(n, dtype=torch.int32, device=key.device)
    with torch.cuda.device(key.device):
        check(lib.mq_argsort_stable(key.data_ptr(), _ptr(subset), n, int(descending), out.data_ptr(), None, 0,
                                    _stream(key)), "mq_argsort_stable")
    return out


# ---------------------------------------------------------------------------------------------- K1 / P0
@dataclass
class QuantizedRows:
    """Output of the fused row quantiser: what the mixed GEMM consumes for one operand."""
    codes: torch.Tensor            # int8  [R, ld_codes]  codes in [-qmax, qmax], zero beyond Kn
    scales: torch.Tensor           # fp32  [G, ld_scales] group-major
    sal: Optional[torch.Tensor]    # dtype [R, ld_sal]    salient columns (zero padded) [+ dequantised tail groups]
    rows: int
    Kn: int
    Ks: int                        # columns of `sal` the 16-bit MMA contracts (salient, pad to 8, hybrid tail)
    group_size: int
    int_groups: Optional[int] = None   # groups held as integer codes (None: all of them)
    n_salient: Optional[int] = None    # true salient column count (Ks when there is no hybrid tail)

    @property
    def groups(self) -> int:
        return self.scales.shape[0] if self.int_groups is None else self.int_groups


def quant_rows(x: torch.Tensor, perm: Optional[torch.Tensor], sal: Optional[torch.Tensor], group_size: int,
               n_bits: int = 4, smooth: Optional[torch.Tensor] = None, Kn: Optional[int] = None,
               int_groups: Optional[int] = None) -> QuantizedRows:
    """codes/scales of x[:, perm] per group of ``group_size`` plus the salient columns x[:, sal] in x's dtype.
    perm=None means the identity over the first Kn columns (fake_quant.py:78-101 / :30-53 on x[:, perm]).
    ``int_groups`` < G (hybrid engine): only the first ``int_groups`` groups leave as codes, the rest leave
    dequantised behind the salient columns of ``sal`` (mq_quant_rows_split)."""
    blocked = hasattr(x, "buf") and hasattr(x, "n_cols")      # parallel.BlockedColumns: [world, M, S] gather layout
    if blocked:
        bc, x = x, x.buf
        bc.wait()
        _need_cuda(x, perm, sal, smooth)
        if smooth is not None:
            raise ValueError("a blocked input takes no smoothing divisor; call .dense() first")
        x2 = x
        nb, R, S = x2.shape
        K = bc.n_cols
    else:
        _need_cuda(x, perm, sal, smooth)
        x2 = _rows2d(x)
        R, K = x2.shape
    dev = x.device
    perm = _i32(perm, dev)
    sal = _i32(sal, dev)
    if Kn is None:
        Kn = perm.numel() if perm is not None else K
    Ks = 0 if sal is None else sal.numel()
    g = int(group_size)
    G = (Kn + g - 1) // g
    Gi = G if int_groups is None else int(int_groups)
    if not 0 <= Gi <= G:
        raise ValueError(f"int_groups={Gi} outside [0, {G}]")
    tail = (G - Gi) * g
    if blocked and tail:
        raise ValueError("a blocked input takes no hybrid split; call .dense() first")
    if tail and x2.dtype not in (torch.float16, torch.bfloat16):
        raise TypeError("the hybrid tail is stored in the 16-bit compute dtype; cast the input first")
    ld_codes = max(_round_up(Gi * g, 128), 128)
    ld_scales = max(_round_up(R, 4), 4)
    ks_total = (_round_up(Ks, 8) + tail) if tail else Ks
    ld_sal = _round_up(ks_total, 8)
    codes = torch.empty((R, ld_codes), dtype=torch.int8, device=dev)
    scales = torch.empty((max(Gi, 1), ld_scales), dtype=torch.float32, device=dev)
    sal_out = torch.empty((R, ld_sal), dtype=x2.dtype, device=dev) if ks_total > 0 else None
    if smooth is not None:
        smooth = smooth.to(device=dev, dtype=x2.dtype).contiguous()
    ev = None
    if QUANT_EVENTS is not None:
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record(torch.cuda.current_stream(dev))
    with torch.cuda.device(dev):
        if blocked:
            check(lib.mq_quant_rows_blocked(x2.data_ptr(), R, K, nb, S, R * S, S, _dtype_code(x2), _ptr(perm), Kn,
                                            _ptr(sal), Ks, g, n_bits, codes.data_ptr(), ld_codes, scales.data_ptr(),
                                            ld_scales, _ptr(sal_out), ld_sal, _stream(x)), "mq_quant_rows_blocked")
        else:
            check(lib.mq_quant_rows_split(x2.data_ptr(), R, K, _ld(x2), _dtype_code(x2), _ptr(smooth), _ptr(perm), Kn,
                                          _ptr(sal), Ks, g, n_bits, Gi, codes.data_ptr(), ld_codes,
                                          scales.data_ptr(), ld_scales, _ptr(sal_out), ld_sal, _stream(x)),
                  "mq_quant_rows_split")
    if ev is not None:
        ev[1].record(torch.cuda.current_stream(dev))
        # algorithmic bytes: x read once, codes + fp32 scales + salient copy written once
        QUANT_EVENTS.append((ev[0], ev[1], R * K * x2.element_size() + R * Gi * g + Gi * R * 4
                             + R * ld_sal * x2.element_size()))
    return QuantizedRows(codes=codes, scales=scales[:Gi], sal=sal_out, rows=R, Kn=Kn, Ks=ks_total, group_size=g,
                         int_groups=Gi, n_salient=Ks)


def fakequant_rows(x: torch.Tensor, perm: Optional[torch.Tensor], group_size: int, n_bits: int = 4,
                   Kn: Optional[int] = None, inplace: bool = False) -> torch.Tensor:
    """Reference-shaped fake quantisation: code*scale at the original column of every permuted position, x itself
    elsewhere (fake_quant.py:105-154, :157-207, :57-64, :10-16, :301)."""
    _need_cuda(x, perm)
    x2 = _rows2d(x)
    R, K = x2.shape
    perm = _i32(perm, x.device)
    if Kn is None:
        Kn = perm.numel() if perm is not None else K
    if inplace:
        if x2.data_ptr() != x.data_ptr():
            raise RuntimeError("in-place fake quantisation needs a tensor with contiguous rows")
        out = x2
    else:
        out = torch.empty((R, K), dtype=x2.dtype, device=x.device)
    with torch.cuda.device(x.device):
        check(lib.mq_fakequant_rows(x2.data_ptr(), R, K, _ld(x2), _dtype_code(x2), _ptr(perm), Kn, int(group_size),
                                    n_bits, out.data_ptr(), _ld(out), _stream(x)), "mq_fakequant_rows")
    if inplace and not x.is_inference():
        torch.autograd.graph.increment_version(x)      # written through a raw pointer: tell torch (and the act cache)
    return x if inplace else out.view(x.shape)


def fakequant_tensor(x: torch.Tensor, n_bits: int = 4, inplace: bool = False) -> torch.Tensor:
    """One global scale (fake_quant.py:68-75, :20-26)."""
    _need_cuda(x)
    xc = x if x.is_contiguous() else x.contiguous()
    if inplace and xc.data_ptr() != x.data_ptr():
        raise RuntimeError("in-place fake quantisation needs a contiguous tensor")
    out = xc if inplace else torch.empty_like(xc)
    ws = torch.empty(16, dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        check(lib.mq_fakequant_tensor(xc.data_ptr(), xc.numel(), _dtype_code(xc), n_bits, out.data_ptr(),
                                      ws.data_ptr(), 16, _stream(x)), "mq_fakequant_tensor")
    if inplace and not x.is_inference():
        torch.autograd.graph.increment_version(x)
    return x if inplace else out.view(x.shape)


def pack_int4(codes: torch.Tensor, cols: Optional[int] = None) -> torch.Tensor:
    """int8 codes [R, ld] -> uint8 [R, cols/2], two's-complement nibbles, even column in the low nibble."""
    _need_cuda(codes)
    assert codes.dtype == torch.int8 and codes.dim() == 2 and codes.stride(1) == 1
    R = codes.shape[0]
    cols = codes.shape[1] if cols is None else cols
    packed = torch.empty((R, cols // 2), dtype=torch.uint8, device=codes.device)
    with torch.cuda.device(codes.device):
        check(lib.mq_pack_int4(codes.data_ptr(), R, cols, _ld(codes), packed.data_ptr(), max(cols // 2, 1),
                               _stream(codes)), "mq_pack_int4")
    return packed


def unpack_int4(packed: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    _need_cuda(packed)
    assert packed.dtype == torch.uint8 and packed.dim() == 2 and packed.stride(1) == 1
    R, half = packed.shape
    if out is None:
        out = torch.empty((R, half * 2), dtype=torch.int8, device=packed.device)
    with torch.cuda.device(packed.device):
        check(lib.mq_unpack_int4(packed.data_ptr(), R, half * 2, _ld(packed), out.data_ptr(), _ld(out),
                                 _stream(packed)), "mq_unpack_int4")
    return out


NORM_KINDS = {"rms": 1, "layer": 2}


def norm_quant_rows(x: torch.Tensor, norm_weight: torch.Tensor, norm_bias: Optional[torch.Tensor], eps: float,
                    kind: str, perm: Optional[torch.Tensor], sal: Optional[torch.Tensor], group_size: int,
                    n_bits: int = 4, Kn: Optional[int] = None, want_y: bool = False):
    """RMSNorm / LayerNorm fused into the row quantiser (mq_norm_quant_rows): x [*, K] is read once,
    the normalised row only ever lives in shared memory.  Returns the QuantizedRows of norm(x) (what quant_rows would
    return for the normalised tensor, bit for bit given the same normalised values) and, with ``want_y``, the
    normalised rows themselves.  Anchors: smooth.py:126-141 (which norm feeds which linears), fake_quant.py:78-101."""
    _need_cuda(x, norm_weight, norm_bias, perm, sal)
    if kind not in NORM_KINDS:
        raise ValueError(f"Invalid norm kind: {kind}")
    x2 = _rows2d(x)
    if x2.dtype not in (torch.float16, torch.bfloat16):
        raise TypeError("norm_quant_rows computes on float16 / bfloat16 rows")
    R, K = x2.shape
    dev = x.device
    perm, sal = _i32(perm, dev), _i32(sal, dev)
    if Kn is None:
        Kn = perm.numel() if perm is not None else K
    Ks = 0 if sal is None else sal.numel()
    g = int(group_size)
    G = (Kn + g - 1) // g
    w = norm_weight.to(device=dev, dtype=x2.dtype).contiguous()
    b = None if norm_bias is None else norm_bias.to(device=dev, dtype=x2.dtype).contiguous()
    ld_codes = max(_round_up(G * g, 128), 128)
    ld_scales = max(_round_up(R, 4), 4)
    ld_sal = _round_up(Ks, 8)
    codes = torch.empty((R, ld_codes), dtype=torch.int8, device=dev)
    scales = torch.empty((max(G, 1), ld_scales), dtype=torch.float32, device=dev)
    sal_out = torch.empty((R, ld_sal), dtype=x2.dtype, device=dev) if Ks > 0 else None
    y = torch.empty((R, K), dtype=x2.dtype, device=dev) if want_y else None
    ev = None
    if QUANT_EVENTS is not None:
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record(torch.cuda.current_stream(dev))
    with torch.cuda.device(dev):
        check(lib.mq_norm_quant_rows(x2.data_ptr(), R, K, _ld(x2), _dtype_code(x2), NORM_KINDS[kind], w.data_ptr(),
                                     _ptr(b), float(eps), _ptr(perm), Kn, _ptr(sal), Ks, g, n_bits, codes.data_ptr(),
                                     ld_codes, scales.data_ptr(), ld_scales, _ptr(sal_out), ld_sal, _ptr(y), K,
                                     _stream(x)), "mq_norm_quant_rows")
    if ev is not None:
        ev[1].record(torch.cuda.current_stream(dev))
        QUANT_EVENTS.append((ev[0], ev[1], R * K * x2.element_size() + R * G * g + G * R * 4
                             + R * ld_sal * x2.element_size() + (R * K * x2.element_size() if want_y else 0)))
    a = QuantizedRows(codes=codes, scales=scales[:G], sal=sal_out, rows=R, Kn=Kn, Ks=Ks, group_size=g, int_groups=G,
                      n_salient=Ks)
    return (a, y) if want_y else a


# ---------------------------------------------------------------------------------------------- K2
def mixed_gemm(a: Optional[QuantizedRows], w: Optional[QuantizedRows], bias: Optional[torch.Tensor],
               a_sal: Optional[torch.Tensor] = None, w_sal: Optional[torch.Tensor] = None,
               Ks: Optional[int] = None, out_dtype: Optional[torch.dtype] = None,
               out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = sum_g a_scale*w_scale*(a_codes . w_codes) + a_sal . w_sal + bias   (replaces F.linear, fake_quant.py:306).

    ``a`` / ``w`` carry codes, scales and (optionally) the salient slices; pass ``a=None, w=None`` with explicit
    ``a_sal`` / ``w_sal`` for the pure 16-bit GEMM of mode="reference".  ``out``: a [M, >= N rounded up to 8] tensor of
    the compute dtype with unit column stride to write into (a column-parallel layer's own block of the gather
    buffer)."""
    if a is not None:
        a_sal = a.sal if a_sal is None else a_sal
        w_sal = w.sal if w_sal is None else w_sal
        if Ks is None:
            Ks = a.Ks
        M, N, G, g = a.rows, w.rows, a.groups if a.Kn > 0 else 0, a.group_size
        if G != (w.groups if w.Kn > 0 else 0) or (G > 0 and g != w.group_size) or a.Ks != w.Ks:
            raise ValueError("activation / weight group partitions differ")
        dev = a.codes.device
    else:
        M, N, G, g = a_sal.shape[0], w_sal.shape[0], 0, 32
        if Ks is None:
            Ks = a_sal.shape[1]
        dev = a_sal.device
    _need_cuda(a_sal, w_sal, bias)
    if Ks and Ks > 0:
        dt = a_sal.dtype
        if w_sal.dtype != dt:
            raise TypeError("salient slices of activation and weight must share a dtype")
    else:
        dt = out_dtype or (bias.dtype if bias is not None else torch.float16)
        Ks = 0
    if dt not in (torch.float16, torch.bfloat16):
        raise TypeError(f"mixed_gemm computes in float16 / bfloat16, got {dt}")
    if bias is not None:
        bias = bias.reshape(-1).to(device=dev, dtype=dt).contiguous()
    if out is None:
        y = torch.empty((M, _round_up(N, 8)), dtype=dt, device=dev)
    else:
        if out.dim() != 2 or out.shape[0] != M or out.shape[1] < _round_up(N, 8) or out.stride(1) != 1 \
                or out.dtype != dt or out.device != dev or out.stride(0) % 8 or out.data_ptr() % 16:
            raise ValueError(f"mixed_gemm: out must be a [{M}, >= {_round_up(N, 8)}] {dt} tensor with unit column "
                             "stride and 16-byte aligned rows")
        y = out
    ev = None
    if GEMM_EVENTS is not None:
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record(torch.cuda.current_stream(dev))
    with torch.cuda.device(dev):
        check(lib.mq_mixed_gemm(
            _ptr(a.codes) if G else None, a.codes.stride(0) if G else 0,
            _ptr(a.scales) if G else None, a.scales.stride(0) if G else 0,
            _ptr(a_sal) if Ks else None, a_sal.stride(0) if Ks else 0,
            _ptr(w.codes) if G else None, w.codes.stride(0) if G else 0,
            _ptr(w.scales) if G else None, w.scales.stride(0) if G else 0,
            _ptr(w_sal) if Ks else None, w_sal.stride(0) if Ks else 0,
            _ptr(bias), M, N, G, g, Ks, _DTYPES[dt], y.data_ptr(), y.stride(0),
            torch.cuda.current_stream(dev).cuda_stream), "mq_mixed_gemm")
    if ev is not None:
        ev[1].record(torch.cuda.current_stream(dev))
        # algorithmic work: the true quantised columns (zero padding of the last group is not work) on the integer
        # pipe, the salient columns (+ hybrid tail) on the 16-bit pipe
        kn_int = min(a.Kn, G * g) if a is not None else 0
        ks_true = (a.n_salient if a is not None and a.n_salient is not None else Ks)
        if a is not None and G * g < a.Kn:
            ks_true += a.Kn - G * g
        GEMM_EVENTS.append((ev[0], ev[1], 2.0 * M * N * kn_int, 2.0 * M * N * ks_true))
    return y[:, :N]


# ---------------------------------------------------------------------------------------------- K2d (decode, int4)
DECODE_MAX_TOKENS = 64      # what mq_decode_gemm accepts
DECODE_PREFER_TOKENS = 32   # up to here it beats the prefill kernel on every measured shape (profiles/r02_decode_*);
                            # at 33..64 tokens the 128-row tcgen05 tiles are already faster when the int8 codes are resident
DECODE_EVENTS = None        # set to a list to collect (start_event, end_event, weight_bytes) per mq_decode_gemm


@dataclass
class DecodeWeights:
    """INT4-resident weight operand of the decode kernel (mq_pack_decode layout; the group scales travel inside the
    nibble stream, in the compute dtype)."""
    nib: torch.Tensor              # uint8: per (row tile, 256-column block) 2048 B of nibbles + the block's scales
    sal: Optional[torch.Tensor]    # uint8: salient slice (16-bit values) in mma fragment order
    rows: int
    groups: int
    group_size: int
    Ks: int                        # salient columns
    dtype: torch.dtype

    def nbytes(self) -> int:
        return self.nib.numel() + (0 if self.sal is None else self.sal.numel())

    @property
    def record_bytes(self) -> int:
        return 2048 + 32 * max(1, 256 // self.group_size)


def decode_supported(group_size: int, int_groups: Optional[int], groups: int, n_bits: int) -> bool:
    return group_size >= 64 and (group_size & (group_size - 1)) == 0 and groups >= 1 \
        and (int_groups is None or int_groups == groups) and n_bits <= 4


def pack_decode(w: QuantizedRows, dtype: Optional[torch.dtype] = None) -> DecodeWeights:
    """int8 codes [N, ld] + fp32 group scales + salient [N, ld_sal] of the prefill operand -> the stream the decode
    kernel reads.  ``dtype``: the 16-bit compute dtype (default: that of the salient slice, else float16)."""
    _need_cuda(w.codes, w.sal)
    N, G, g = w.rows, w.groups, w.group_size
    Ks = 0 if w.sal is None else (w.n_salient if w.n_salient is not None else w.Ks)
    dt = dtype or (w.sal.dtype if w.sal is not None else torch.float16)
    import ctypes as C
    nb, sb = C.c_size_t(0), C.c_size_t(0)
    _check(lib.mq_decode_weight_bytes(N, G, g, Ks, C.byref(nb), C.byref(sb)), "mq_decode_weight_bytes")
    dev = w.codes.device
    nib = torch.empty(nb.value, dtype=torch.uint8, device=dev)
    sal = torch.empty(sb.value, dtype=torch.uint8, device=dev) if Ks else None
    with torch.cuda.device(dev):
        check(lib.mq_pack_decode(w.codes.data_ptr(), w.codes.stride(0), w.scales.data_ptr(), w.scales.stride(0),
                                 _ptr(w.sal) if Ks else None, w.sal.stride(0) if Ks else 0, N, G, g, Ks, _DTYPES[dt],
                                 nib.data_ptr(), _ptr(sal), torch.cuda.current_stream(dev).cuda_stream), "mq_pack_decode")
    return DecodeWeights(nib=nib, sal=sal, rows=N, groups=G, group_size=g, Ks=Ks, dtype=dt)


def decode_gemm(a: QuantizedRows, w: DecodeWeights, bias: Optional[torch.Tensor],
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y for M <= 64 tokens from int4-resident weights (same contract as mixed_gemm; fake_quant.py:306)."""
    M, N, G, g = a.rows, w.rows, w.groups, w.group_size
    if M > DECODE_MAX_TOKENS:
        raise ValueError(f"decode_gemm handles at most {DECODE_MAX_TOKENS} tokens, got {M}")
    if a.groups != G or a.group_size != g or (a.n_salient if a.n_salient is not None else a.Ks) != w.Ks:
        raise ValueError("activation / weight group partitions differ")
    dev = a.codes.device
    dt = w.dtype
    if w.Ks and a.sal.dtype != dt:
        raise TypeError("salient slices of activation and weight must share a dtype")
    if bias is not None:
        bias = bias.reshape(-1).to(device=dev, dtype=dt).contiguous()
    if out is None:
        y = torch.empty((M, _round_up(N, 8)), dtype=dt, device=dev)
    else:
        if out.dim() != 2 or out.shape[0] != M or out.shape[1] < _round_up(N, 8) or out.stride(1) != 1 \
                or out.dtype != dt or out.device != dev:
            raise ValueError(f"decode_gemm: out must be a [{M}, >= {_round_up(N, 8)}] {dt} tensor, unit column stride")
        y = out
    ev = None
    if DECODE_EVENTS is not None:
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record(torch.cuda.current_stream(dev))
    with torch.cuda.device(dev):
        check(lib.mq_decode_gemm(a.codes.data_ptr(), a.codes.stride(0), a.scales.data_ptr(), a.scales.stride(0),
                                 _ptr(a.sal) if w.Ks else None, a.sal.stride(0) if w.Ks else 0,
                                 w.nib.data_ptr(), _ptr(w.sal), _ptr(bias),
                                 M, N, G, g, w.Ks, _DTYPES[dt], y.data_ptr(), y.stride(0),
                                 torch.cuda.current_stream(dev).cuda_stream), "mq_decode_gemm")
    if ev is not None:
        ev[1].record(torch.cuda.current_stream(dev))
        DECODE_EVENTS.append((ev[0], ev[1], w.nbytes()))
    return y[:, :N]
