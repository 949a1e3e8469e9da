"""Module-swap API of the W4A4 MixedQuant linear path, B200-native.

Mirrors the public surface of the reference's ``smoothquant/fake_quant.py`` (names, argument order, defaults and
exception types) so that ``from smoothquant.fake_quant import W4A4Linear, quantize_model, ...`` keeps working, but
every tensor operation on the path is a hand-written sm_100a kernel reached through the C ABI of ``include/mq.h``
(``ops`` -> ctypes -> ``libmq.so``).  There is no CPU fallback: tensors must live on a CUDA device.

Two execution modes (SURVEY.md 8(a')):

``mode="reference"``  literal semantics of the reference layer.  Activations: the non-salient columns are compacted,
    sorted by their per-call column abs-max and grouped (fake_quant.py:291-301 -> :105-154); weights: ALL K columns
    are sorted by the weight's own column abs-max, grouped, and the salient columns restored afterwards
    (fake_quant.py:347-365).  The two operands use different K partitions, so the contraction runs on dequantised
    16-bit values (tcgen05 kind::f16).  Kernels: K0 column stats -> K0b stable argsort -> K1 fake-quantise -> K2.

``mode="fused"``  the fast path named by BASELINE.json: ONE static permutation of the non-salient columns shared by
    both operands, activations quantised to integer codes + per-(token, group) scales by K1, weights packed offline
    by the same kernel, contraction as tcgen05 kind::i8 per K group with an fp32 rescale plus the salient slice as
    kind::f16 in the same CTA (K2).  Codes / scales are bit-exact against the reference's natural-order group
    quantisers (fake_quant.py:78-101, :30-53) applied to ``x[:, perm]`` / ``W[:, perm]``.

``torch.argsort`` in the reference is unstable under ties; this implementation always orders ties by
ascending index and accepts an explicit permutation where parity against a recorded reference run is wanted.
"""
from __future__ import annotations

import threading
from typing import Dict, Optional

import torch
from torch import nn

from . import ops, parallel

__all__ = [
    "quantize_weight_per_channel_absmax", "quantize_weight_per_tensor_absmax", "quantize_weight_per_group_absmax",
    "quantize_activation_per_token_absmax", "quantize_activation_per_tensor_absmax",
    "quantize_activation_per_group_absmax", "quantize_activation_per_group_absmax_sort",
    "quantize_weight_per_group_absmax_sort", "W4A4Linear", "quantize_opt", "quantize_llama_like", "quantize_model",
    "save_packed_model", "load_packed_model", "fold_mlp_permutation",
]

SORT_KINDS = ("max", "mean3std", "argmax", "none")
MODES = ("reference", "fused")
_ACT_QUANTS = ("per_token", "per_tensor", "per_group")
_WEIGHT_QUANTS = ("per_channel", "per_tensor", "per_group")


# ------------------------------------------------------------------------------------------- quantiser functions
# Same names / positional arguments as the reference.  Each returns the fake-quantised (dequantised) tensor in the
# input dtype.  Like the reference, the per-token / per-tensor / per-channel variants overwrite their argument
# (in-place div_/round_/mul_, fake_quant.py:15, :25, :63, :74); the per-group variants never touch it.

@torch.no_grad()
def quantize_weight_per_channel_absmax(w, n_bits):
    """fake_quant.py:10-16 -- one scale per output channel (row); in place."""
    return ops.fakequant_rows(w, None, w.shape[-1], n_bits, inplace=True)


@torch.no_grad()
def quantize_weight_per_tensor_absmax(w, n_bits):
    """fake_quant.py:20-26 -- one global scale; in place."""
    return ops.fakequant_tensor(w, n_bits, inplace=True)


@torch.no_grad()
def quantize_weight_per_group_absmax(w, n_bits, group_size=128):
    """fake_quant.py:30-53 -- natural-order groups of ``group_size`` input columns per output row."""
    return ops.fakequant_rows(w, None, group_size, n_bits)


@torch.no_grad()
def quantize_activation_per_token_absmax(t, n_bits):
    """fake_quant.py:57-64 -- one scale per token (row of the flattened tensor); in place."""
    return ops.fakequant_rows(t, None, t.shape[-1], n_bits, inplace=True)


@torch.no_grad()
def quantize_activation_per_tensor_absmax(t, n_bits):
    """fake_quant.py:68-75 -- one global scale; in place."""
    return ops.fakequant_tensor(t, n_bits, inplace=True)


@torch.no_grad()
def quantize_activation_per_group_absmax(t, n_bits, group_size=128):
    """fake_quant.py:78-101 -- natural-order groups along the last dimension."""
    return ops.fakequant_rows(t, None, group_size, n_bits)


def _dynamic_perm(t2d: torch.Tensor, subset: Optional[torch.Tensor] = None, kind: str = "max") -> torch.Tensor:
    """Ascending stable argsort of the column statistic over ALL rows (fake_quant.py:113-116, :164-167)."""
    if kind == "none":
        return subset if subset is not None else torch.arange(t2d.shape[-1], dtype=torch.int32, device=t2d.device)
    return ops.argsort_stable(ops.colstat(t2d, kind), subset=subset)


@torch.no_grad()
def quantize_activation_per_group_absmax_sort(t, n_bits, group_size=128, perm=None):
    """fake_quant.py:105-154 -- sort channels by max |t| over every row, group, quantise, undo the permutation.
    ``perm`` (optional, extension) replays a recorded permutation instead of computing one."""
    if perm is None:
        perm = _dynamic_perm(t.reshape(-1, t.shape[-1]))
    else:
        _validate_columns(perm, t.shape[-1], "perm")
    return ops.fakequant_rows(t, perm, group_size, n_bits)


@torch.no_grad()
def quantize_weight_per_group_absmax_sort(w, n_bits, group_size=128, perm=None):
    """fake_quant.py:157-207 -- the same algorithm on w [N, K]; the key is the per-input-column max over rows."""
    if perm is None:
        perm = _dynamic_perm(w)
    else:
        _validate_columns(perm, w.shape[-1], "perm")
    return ops.fakequant_rows(w, perm, group_size, n_bits)


# ------------------------------------------------------------------------------------------- the layer
# Quantised-activation sharing between layers fed by the SAME tensor with the same K partition (q/k/v, gate/up: the
# reference's calibration hooks give them identical statistics, fake_quant.py:396-455): K1 runs once, the others reuse
# its output.  One entry per (device, CUDA stream): a hit can only happen on the stream that produced the entry, so
# the consumer is stream-ordered behind the producer without an event.  An entry is identified by the identity of the
# tensor object the caller passed (kept alive, so neither its id nor its address can be recycled), its version counter
# when it has one (inference tensors do not: there the caller must not mutate a tensor between the calls of one
# group, which no decoder layer does) and the partition id.  Walkers tell each layer how many layers share its input
# (`_share_n`); the entry is dropped when the last of them has consumed it, so nothing stays pinned after a model
# forward.  Layers built one by one (`_share_n` unset) keep the classic behaviour: the entry lives until the next
# miss on that stream, or until clear_act_cache().
_ACT_CACHE = {"hits": 0, "entries": {}}
_ACT_LOCK = threading.Lock()


def clear_act_cache():
    with _ACT_LOCK:
        _ACT_CACHE["entries"].clear()


# Share of the K groups that run as integer codes (tcgen05 kind::i8 + per-group fp32 rescale on the CUDA cores); the
# rest leave the quantiser DEQUANTISED (code * scale rounded to the 16-bit type, the reference's fake-quantised value)
# and run through the 16-bit tensor path together with the salient columns.  The rescale bounds the integer path for
# small groups.  Measured on a B200 at 32768 x 4096 x 4096, 5 % salient (profiles/r01_hybrid_sweep.txt):
# g = 64: 1.46 ms all-integer, 0.88 ms all-dequantised, monotone in between; g = 128: 0.83 ms all-integer, 0.80 ms at
# 0.7 (but the quantiser then writes 2-byte tails and K = 11008 loses), 0.89 ms all-dequantised; g = 256: all-integer
# wins everywhere.  So groups below 128 default to the 16-bit path (the packed weight then holds 2 bytes per
# element; pass int_fraction=1.0 to keep 1-byte codes, e.g. for weight-stream-bound decode), 128 and up to integers.
def default_int_fraction(g):
    return 0.0 if g < 128 else 1.0


def _int_groups(G, g, frac):
    if frac is None:
        frac = default_int_fraction(g)
    if not 0.0 <= frac <= 1.0:
        raise ValueError(f"int_fraction must be in [0, 1], got {frac}")
    gi = int(round(G * frac))
    if gi >= G:                             # everything on the integer engine: no tail, the last block is zero padded
        return G
    step = max(1, 128 // g)                 # a hybrid split must leave whole 128-column K blocks of integer groups
    return max(0, gi // step * step)


def _version_of(t):
    try:
        return None if t.is_inference() else t._version
    except RuntimeError:            # no version counter: identity of the object is all there is
        return None


def _shared_act(x_orig, xin, mod, producer=None):
    """K1 output for `xin` (= x_orig as a 2-D tensor of the compute dtype), shared within a group of layers.
    ``producer(perm, sal)``: what to run instead of ops.quant_rows on a miss (the fused norm + quantiser)."""
    pid = getattr(mod, "_partition_id", None)
    share = getattr(mod, "_share_n", None)
    dev = xin.device
    slot = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    sig = (_version_of(x_orig), xin.data_ptr(), tuple(xin.shape), xin.stride(), xin.dtype, pid)
    if pid is not None:
        with _ACT_LOCK:
            e = _ACT_CACHE["entries"].get(slot)
            if e is not None and e["x"] is x_orig and e["sig"] == sig:
                _ACT_CACHE["hits"] += 1
                a = e["a"]
                if e["left"] is not None:
                    e["left"] -= 1
                    if e["left"] <= 0:
                        del _ACT_CACHE["entries"][slot]
                return a
    perm, sal = mod.exec_partition()
    if producer is not None:
        a = producer(perm, sal)
    else:
        a = ops.quant_rows(xin, perm, sal, mod.group_size, mod.quant_bits,
                           Kn=0 if mod.perm is None else mod.perm.numel(), int_groups=mod.int_groups)
    if pid is not None and (share is None or share > 1):
        with _ACT_LOCK:
            _ACT_CACHE["entries"][slot] = {"x": x_orig, "xin": xin, "sig": sig, "a": a,
                                           "left": None if share is None else share - 1}
    return a


class NormedRows:
    """Lazy output of a FusedNorm: the norm's INPUT plus the norm itself.  A mode="fused" W4A4Linear consumes it
    through mq_norm_quant_rows (norm + quantise in one pass over x); anything else calls
    ``materialize()``.  Carries just enough of the tensor interface (shape / size / dim / dtype / device) for the
    attention and MLP blocks of transformers' Llama / Mistral / OPT to pass it on to their linears."""

    def __init__(self, x, norm, kind):
        self.x, self.norm, self.kind = x, norm, kind
        self._y = None

    @property
    def shape(self):
        return self.x.shape

    @property
    def dtype(self):
        return self.x.dtype

    @property
    def device(self):
        return self.x.device

    def dim(self):
        return self.x.dim()

    def size(self, *a):
        return self.x.size(*a)

    def is_inference(self):
        return self.x.is_inference()

    @property
    def _version(self):
        return self.x._version

    @property
    def eps(self):
        return float(getattr(self.norm, "variance_epsilon", getattr(self.norm, "eps", 1e-6)))

    def materialize(self):
        if self._y is None:
            self._y = self.norm(self.x)
        return self._y


class FusedNorm(nn.Module):
    """Stands in for the RMSNorm / LayerNorm in front of a group of mode="fused" W4A4 linears that share one K
    partition (q/k/v behind input_layernorm, gate/up behind post_attention_layernorm; smooth.py:126-141): returns a
    NormedRows for 16-bit CUDA inputs, the plain norm output otherwise."""

    def __init__(self, norm, kind):
        super().__init__()
        _check_choice(kind, ("rms", "layer"), "norm kind")
        self.norm, self.kind = norm, kind

    def forward(self, x):
        if x.is_cuda and x.dtype in (torch.float16, torch.bfloat16) and x.shape[-1] % 8 == 0 and x.dim() in (2, 3) \
                and (self.kind == "rms" or tuple(self.norm.normalized_shape) == (x.shape[-1],)):
            return NormedRows(x, self.norm, self.kind)
        return self.norm(x)


def fuse_norm_(parent, norm_attr, consumers, kind):
    """Replace ``parent.<norm_attr>`` by a FusedNorm when every consumer is a packed mode="fused" W4A4Linear with the
    same partition on the integer engine (otherwise the norm would be recomputed per consumer).  Returns True if
    fused."""
    norm = getattr(parent, norm_attr, None)
    if norm is None or isinstance(norm, FusedNorm) or getattr(norm, "weight", None) is None:
        return False
    ok = all(isinstance(c, W4A4Linear) and c.mode == "fused" and c._norm_fusable() for c in consumers)
    if not ok or len({getattr(c, "_partition_id", None) for c in consumers}) != 1:
        return False
    setattr(parent, norm_attr, FusedNorm(norm, kind))
    return True


def _validate_columns(idx, K, what):
    """A column list installed into a layer (permutation, salient set) must hold distinct ids in [0, K): the
    quantiser kernels turn it into 16-bit shared-memory offsets without a bounds check.  Offline / parity-replay
    paths only (one device sync)."""
    if idx is None or idx.numel() == 0:
        return
    i64 = torch.as_tensor(idx).long()
    lo, hi = int(i64.min()), int(i64.max())
    if lo < 0 or hi >= K:
        raise ValueError(f"{what}: column ids must lie in [0, {K}), got [{lo}, {hi}]")
    if int(torch.unique(i64).numel()) != i64.numel():
        raise ValueError(f"{what}: duplicate column ids")


def _check_choice(value, allowed, what):
    if value not in allowed:
        raise ValueError(f"Invalid {what}: {value}")


PACKED_FORMAT = 1      # version tag of packed_state / save_packed_model


class W4A4Linear(nn.Module):
    """Drop-in for the reference's ``W4A4Linear`` (fake_quant.py:209-374).

    Constructor arguments, attribute names (``in_features out_features group_size weight bias salient_indices
    act_quant_name weight_quant_name output_quant_name``), ``from_float`` and ``__repr__`` follow the reference;
    ``mode`` / ``sort`` / ``sort_key`` are new, defaulted keyword arguments.
    """

    def __init__(self, in_features, out_features, bias=True, act_quant="per_token", quantize_output=False,
                 importance=None, salient_prop=0, quant_bits=4, group_size=128, mode="reference", sort="max",
                 device=None, dtype=torch.float16):
        super().__init__()
        _check_choice(act_quant, _ACT_QUANTS, "act_quant")          # ValueError like fake_quant.py:256
        _check_choice(mode, MODES, "mode")
        _check_choice(sort, SORT_KINDS, "sort")
        self.in_features = in_features
        self.out_features = out_features
        self.group_size = group_size
        self.quant_bits = quant_bits
        self.mode = mode
        self.sort = sort
        self.act_quant_name = act_quant
        self.output_quant_name = act_quant if quantize_output else "None"
        self.weight_quant_name = "None"
        # the reference allocates (and later discards) an fp16 randn weight here (fake_quant.py:227-235); an empty
        # placeholder of the same shape / dtype keeps `.weight.shape` meaningful without the wasted RNG pass
        self.register_buffer("_weight", torch.empty((out_features, in_features), dtype=dtype, device=device))
        self.register_buffer("bias", torch.zeros((1, out_features), dtype=dtype, device=device) if bias else None)
        # fused-mode operands (filled by from_float)
        for name in ("w_codes", "w_scales", "w_sal", "perm", "sal_sorted", "keep", "out_perm", "sal_exec"):
            self.register_buffer(name, None)
        self.input_folded = False           # the producer already emits [perm columns | salient columns] (fold_mlp_permutation)
        self.w_decode = None                # ops.DecodeWeights: INT4-resident operand of the decode kernel (prepare_decode)
        self.int_groups = None              # fused mode: groups kept as integer codes (hybrid engine), set by the packer
        self.int_fraction = None
        self.salient_indices = None
        self.tp_rank, self.tp_world, self.tp_group, self.tp_gather_output = 0, 1, None, True
        self.full_out_features = out_features
        if importance is not None and salient_prop > 0:
            # fake_quant.py:265-270 -- top max(1, int(p*K)) channels by importance, canonical tie order
            imp = torch.as_tensor(importance).float()
            order = torch.argsort(imp, descending=True, stable=True)
            self.salient_indices = order[: max(1, int(salient_prop * len(order)))]

    # ---- reference-compatible `weight` attribute: a stored buffer in reference mode, a lazily dequantised view of
    # the packed operands in fused mode
    @property
    def weight(self):
        if self.mode == "fused" and (self.w_codes is not None or self.w_decode is not None):
            return self.dequantized_weight()
        return self._weight

    @weight.setter
    def weight(self, value):
        self._buffers["_weight"] = value

    def to(self, *args, **kwargs):
        # fake_quant.py:272-277 moves weight and bias explicitly; buffers registered here move with nn.Module.to
        super().to(*args, **kwargs)
        return self

    def _device_indices(self, device):
        """int32 device copies of the salient list (sorted ascending) and of its complement, cached per device (the
        reference re-uploads a CPU LongTensor on every forward)."""
        if self.salient_indices is None:
            return None, None
        if self.sal_sorted is None or self.sal_sorted.device != device:
            sal = torch.sort(self.salient_indices.to(torch.int64)).values
            mask = torch.ones(self.in_features, dtype=torch.bool)
            mask[sal.cpu()] = False
            self.sal_sorted = sal.to(device=device, dtype=torch.int32)
            self.keep = torch.nonzero(mask).flatten().to(device=device, dtype=torch.int32)
        return self.sal_sorted, self.keep

    def exec_partition(self):
        """(perm, sal) the activation quantiser runs with.  Normally the layer's own lists (original column ids); when
        the producer of x already emits the columns as [perm | salient] (fold_mlp_permutation) the permutation is the
        identity over the first Kn columns (None: K1's contiguous mode, no gather) and the salient set is the tail."""
        if not self.input_folded:
            return self.perm, self.sal_sorted
        return None, self.sal_exec

    @torch.no_grad()
    def fold_input_(self):
        """Declare that x arrives as x_orig[:, concat(perm, sal_sorted)].  Returns that column order (int64) for the
        producer to adopt.  The packed weight is untouched: it is stored in permuted order already."""
        if self.mode != "fused" or self.w_codes is None:
            raise ValueError("only packed mode='fused' layers can fold their input permutation")
        if self.input_folded:
            raise RuntimeError("input permutation already folded")
        dev = self.w_codes.device
        kn = self.perm.numel()
        parts = [self.perm.long()]
        if self.sal_sorted is not None:
            parts.append(self.sal_sorted.long())
        sigma = torch.cat(parts)
        if sigma.numel() != self.in_features:
            raise ValueError("permutation + salient set do not cover every input column")
        ks = self.in_features - kn
        self.sal_exec = torch.arange(kn, self.in_features, dtype=torch.int32, device=dev) if ks else None
        self.input_folded = True
        self._partition_id = ("folded",) + tuple(self._partition_id)
        return sigma

    @torch.no_grad()
    def permute_outputs_(self, sigma):
        """Emit y[:, sigma] instead of y: reorder the rows of the packed weight (codes, group scales, salient slice)
        and the bias.  Each output element keeps its own dot product, so values are bit-identical, only their column
        changes.  Composes with an earlier call."""
        if self.mode != "fused" or self.w_codes is None:
            raise ValueError("only packed mode='fused' layers can permute their outputs")
        if self.tp_world != 1:
            raise RuntimeError("permute the outputs before sharding")
        if self.output_quant_name != "None":
            raise ValueError("an output-quantised layer must keep its column order (fake_quant.py:309-316)")
        N = self.out_features
        sigma = torch.as_tensor(sigma).to(device=self.w_codes.device, dtype=torch.long)
        _validate_columns(sigma, N, "output permutation")
        if sigma.numel() != N:
            raise ValueError(f"output permutation must list all {N} output channels")
        self.w_codes = self.w_codes[sigma].contiguous()
        if self.w_sal is not None:
            self.w_sal = self.w_sal[sigma].contiguous()
        sc = torch.zeros_like(self.w_scales)
        sc[:, :N] = self.w_scales[:, :N][:, sigma]
        self.w_scales = sc
        if self.bias is not None:
            self._buffers["bias"] = self.bias.reshape(-1)[sigma].contiguous().view(self.bias.shape)
        self.out_perm = sigma if self.out_perm is None else self.out_perm[sigma]
        return self

    # ---- activation / output fake quantiser of mode="reference" (fake_quant.py:291-304, :309-316)
    def _fake_quant_nonsalient(self, t2d, name, perm=None):
        _, keep = self._device_indices(t2d.device)
        if keep is not None and keep.numel() == 0:
            return t2d.clone()
        if name == "per_group":
            if perm is None:
                perm = keep if self.sort == "none" else _dynamic_perm(t2d, subset=keep, kind=self.sort)
            return ops.fakequant_rows(t2d, perm, self.group_size, self.quant_bits)
        if name == "per_token":
            if keep is None:    # the reference quantises x itself in place (fake_quant.py:63)
                return ops.fakequant_rows(t2d, None, t2d.shape[-1], self.quant_bits, inplace=True)
            return ops.fakequant_rows(t2d, keep, keep.numel(), self.quant_bits)
        if name == "per_tensor":
            if keep is None:
                return ops.fakequant_tensor(t2d, self.quant_bits, inplace=True)
            out = t2d.clone()
            k64 = keep.long()
            out[:, k64] = ops.fakequant_tensor(t2d[:, k64].contiguous(), self.quant_bits, inplace=True)
            return out
        raise ValueError(f"Invalid act_quant: {name}")

    @torch.no_grad()
    def forward(self, x, act_perm=None, out_perm=None):
        if isinstance(x, parallel.BlockedColumns):      # gathered output of a column-parallel producer, [p, M, S]
            return self._forward_blocked_input(x)
        if isinstance(x, NormedRows):                   # the producing norm is fused into the row quantiser
            return self._forward_normed(x, act_perm, out_perm)
        shape = x.shape
        if x.dim() == 3:
            x2 = x.reshape(-1, shape[-1])
        elif x.dim() == 2:
            x2 = x
        else:
            raise ValueError(f"Unsupported input shape: {shape}")       # fake_quant.py:287
        if act_perm is not None or out_perm is not None:               # parity replay of a recorded permutation
            _validate_columns(act_perm, self.in_features, "act_perm")
            _validate_columns(out_perm, self.full_out_features, "out_perm")
        if self._emits_blocked() and x.dim() == 2:
            return self._forward_fused(x, x2, blocked_out=True)
        if self.mode == "fused":
            y = self._forward_fused(x, x2)
        else:
            y = self._forward_reference(x2, act_perm)
        return self._finish(y, x2.dtype, shape if x.dim() == 3 else None, out_perm)

    def _finish(self, y, dtype, shape3, out_perm=None):
        """Everything behind the contraction: gather of a column-parallel layer, output quantiser, dtype, shape."""
        if self.tp_world > 1 and self.tp_gather_output:
            y = parallel.gather_columns(y, self.full_out_features, self.tp_group)
        if self.output_quant_name != "None":                           # fake_quant.py:309-316, on the full width
            y = self._quantize_output(y, out_perm)
        if y.dtype != dtype:
            y = y.to(dtype)
        return y.view(*shape3[:-1], -1) if shape3 is not None else y

    def _norm_fusable(self):
        g = self.group_size
        G = 0 if self.perm is None else (self.perm.numel() + g - 1) // g
        return (self.w_codes is not None or self.w_decode is not None) and G >= 1 and 32 <= g <= 1024 \
            and g & (g - 1) == 0 and (self.int_groups is None or self.int_groups == G) and self.quant_bits <= 7 \
            and 256 <= self.in_features <= 16384 and self.in_features % 8 == 0

    def _forward_normed(self, nx, act_perm=None, out_perm=None):
        x = nx.x
        cdt = self.w_sal.dtype if self.w_sal is not None else (
            self.w_decode.dtype if self.w_decode is not None else x.dtype)
        if self.mode != "fused" or not self._norm_fusable() or x.dtype != cdt or act_perm is not None \
                or x.shape[-1] != self.in_features:
            return self.forward(nx.materialize(), act_perm, out_perm)
        x2 = x.reshape(-1, x.shape[-1])
        w, b = nx.norm.weight, getattr(nx.norm, "bias", None)
        kn = self.perm.numel()

        def producer(perm, sal):
            return ops.norm_quant_rows(x2, w, b, nx.eps, nx.kind, perm, sal, self.group_size, self.quant_bits, Kn=kn)

        if self._emits_blocked() and x.dim() == 2:
            return self._forward_fused(nx, x2, blocked_out=True, producer=producer)
        y = self._forward_fused(nx, x2, producer=producer)
        return self._finish(y, x.dtype, x.shape if x.dim() == 3 else None, out_perm)

    def _emits_blocked(self):
        return (self.tp_world > 1 and self.tp_gather_output and getattr(self, "tp_blocked_output", False)
                and self.mode == "fused" and self.output_quant_name == "None")

    def _forward_blocked_input(self, xb):
        """x arrives as parallel.BlockedColumns: K1 reads the gather layout in place (mq_quant_rows_blocked); layers
        or shapes it does not cover take the dense view."""
        if xb.n_cols != self.in_features:
            raise ValueError(f"blocked input has {xb.n_cols} columns, layer wants {self.in_features}")
        cdt = self.w_sal.dtype if self.w_sal is not None else (
            self.w_decode.dtype if self.w_decode is not None else xb.dtype)
        G = 0 if self.perm is None else (self.perm.numel() + self.group_size - 1) // self.group_size
        g = self.group_size
        ok = (self.mode == "fused" and xb.dtype == cdt and xb.buf.shape[0] > 1 and xb.buf.shape[2] % 8 == 0
              and (self.int_groups is None or self.int_groups == G) and G >= 1 and 32 <= g <= 1024 and g & (g - 1) == 0
              and self.in_features % 8 == 0 and self.quant_bits <= (8 if cdt == torch.float16 else 7))
        if not ok:
            return self.forward(xb.dense())
        if self._emits_blocked():
            return self._forward_fused(xb, xb, blocked_out=True)
        return self._finish(self._forward_fused(xb, xb), xb.dtype, None)

    # ---- column-parallel sharding over out_features; no reference counterpart
    @torch.no_grad()
    def shard_(self, rank, world, group=None, gather_output=True, blocked_output=False):
        """Keep only the output columns that ``rank`` of ``world`` owns (rows of the packed weight / scales / salient
        slice / bias).  With ``gather_output`` the forward all-gathers y back to the full width over ``group``;
        without it the layer returns its [M, N_local] column block (consumer is column-local).  ``blocked_output``
        (mode="fused", 2-D inputs): the gathered output is returned as a parallel.BlockedColumns -- the GEMM writes its
        own block of the [world, M, S] gather buffer, the all-gather runs in place and asynchronously, and the next
        W4A4Linear quantises straight from that layout; call ``.dense()`` on it for any other consumer."""
        if self.tp_world != 1:
            raise RuntimeError("layer is already sharded")
        if self.w_decode is not None:
            raise RuntimeError("shard the layer before prepare_decode (the decode layout is built per shard)")
        if self.output_quant_name != "None" and not gather_output:
            raise ValueError("an output-quantised layer must gather its output: the output quantiser groups columns "
                             "across the whole width (fake_quant.py:309-316)")
        N = self.out_features
        lo, hi = parallel.shard_range(N, world, rank)
        if hi <= lo:
            raise ValueError(f"rank {rank} of {world} owns no output column of a layer with N={N}")
        if self.mode == "fused" and self.w_codes is not None:
            self.w_codes = parallel.shard_rows(self.w_codes, N, world, rank)
            self.w_sal = parallel.shard_rows(self.w_sal, N, world, rank)
            self.w_scales = parallel.shard_scales(self.w_scales, N, world, rank)
        elif self._weight is not None:
            self._buffers["_weight"] = parallel.shard_rows(self._weight, N, world, rank)
            object.__setattr__(self, "_w16_cache", None)
        if self.bias is not None:
            self._buffers["bias"] = parallel.shard_rows(self.bias.reshape(-1), N, world, rank)
        self.full_out_features = N
        self.out_features = hi - lo
        self.tp_rank, self.tp_world, self.tp_group, self.tp_gather_output = rank, world, group, gather_output
        self.tp_blocked_output = bool(blocked_output)
        return self

    def _forward_fused(self, x, x2, blocked_out=False, producer=None):
        cdt = self.w_sal.dtype if self.w_sal is not None else (self.w_decode.dtype if self.w_decode is not None else (
            x2.dtype if x2.dtype in (torch.float16, torch.bfloat16) else torch.float16))
        xin = x2 if x2.dtype == cdt else x2.to(cdt)
        if blocked_out:
            # the GEMM writes this rank's block of the gather buffer; the collective then runs in place and async
            peer = parallel.peer_gather_for(self.tp_group)
            if peer is not None:                # copy engines over peer-mapped memory (parallel.PeerGather)
                slot = peer.buffer(x2.shape[0], self.full_out_features, cdt)
                self._fused_gemm(x, xin, cdt, out=slot["buf"][self.tp_rank], producer=producer)
                return peer.gather(slot, self.full_out_features, parallel.ASYNC_GATHER)
            buf = parallel.gather_buffer(x2.shape[0], self.full_out_features, self.tp_world, cdt, x2.device)
            y = self._fused_gemm(x, xin, cdt, out=buf[self.tp_rank], producer=producer)
            return parallel.gather_columns_blocked(buf[self.tp_rank][:, : y.shape[1]], self.full_out_features,
                                                   self.tp_group, buf=buf)
        return self._fused_gemm(x, xin, cdt, producer=producer)

    def _fused_gemm(self, x, xin, cdt, out=None, producer=None):
        # layers fed by the same tensor with the same K partition (q/k/v, gate/up: same calibration statistics in
        # the reference, fake_quant.py:396-455) share one quantised activation
        a = _shared_act(x, xin, self, producer)
        # weight-bandwidth regime: int4-resident weights.  Up to DECODE_PREFER_TOKENS the decode kernel wins outright;
        # between that and DECODE_MAX_TOKENS it is only taken when the prefill operand is not resident
        if self.w_decode is not None and (a.rows <= ops.DECODE_PREFER_TOKENS or (
                self.w_codes is None and a.rows <= ops.DECODE_MAX_TOKENS)):
            return ops.decode_gemm(a, self.w_decode, self.bias, out=out)
        if self.w_codes is None:
            self._restore_prefill_operand()          # stays resident from now on: the layer serves both regimes
        w = ops.QuantizedRows(codes=self.w_codes, scales=self.w_scales, sal=self.w_sal, rows=self.out_features,
                              Kn=a.Kn, Ks=a.Ks, group_size=self.group_size, int_groups=a.int_groups)
        return ops.mixed_gemm(a, w, self.bias, out_dtype=cdt, out=out)

    # ---- decode regime: INT4-resident weights (BASELINE config 5; no reference counterpart beyond F.linear :306)
    @torch.no_grad()
    def prepare_decode(self, keep_prefill=True):
        """Build the nibble layout the decode kernel (M <= 64 tokens) streams: 4 bits per non-salient weight on the
        device (smoothquant/model_size.py:5-15's accounting made real).  With ``keep_prefill=False`` the int8 codes of
        the prefill kernel are released -- the layer then holds ~0.5 byte per weight; a later forward with more than 64
        tokens rebuilds them from the nibbles (mq_unpack_int4 of a plain nibble copy kept at 0.5 byte per weight would
        double the footprint, so the rebuild re-reads the decode layout instead)."""
        if self.mode != "fused" or self.w_codes is None:
            raise ValueError("prepare_decode needs a packed mode='fused' layer")
        G = (self.perm.numel() + self.group_size - 1) // self.group_size
        if not ops.decode_supported(self.group_size, self.int_groups, G, self.quant_bits):
            raise ValueError("the decode kernel needs integer groups only, quant_bits <= 4 and a group size that is a "
                             f"multiple of 64 (group_size={self.group_size}, int_groups={self.int_groups}/{G}, "
                             f"quant_bits={self.quant_bits})")
        ks = 0 if self.sal_sorted is None else self.sal_sorted.numel()
        wq = ops.QuantizedRows(codes=self.w_codes, scales=self.w_scales, sal=self.w_sal, rows=self.out_features,
                               Kn=self.perm.numel(), Ks=ks, group_size=self.group_size, int_groups=G, n_salient=ks)
        self.w_decode = ops.pack_decode(wq, self.w_sal.dtype if self.w_sal is not None else (
            self.bias.dtype if self.bias is not None and self.bias.dtype in (torch.float16, torch.bfloat16) else torch.float16))
        if not keep_prefill:                # the group scales travel inside the nibble stream: nothing else stays
            self.w_codes = None
            self.w_sal = None
            self.w_scales = None
        return self

    def _restore_prefill_operand(self):
        """Rebuild the prefill kernel's int8 codes / row-major salient slice from the decode layout (a fixed index
        permutation, inverted with tensor views; offline path -- only taken when a layer prepared with
        keep_prefill=False is asked for more than 64 tokens, its dequantised weight or a checkpoint)."""
        d = self.w_decode
        if d is None:
            raise RuntimeError("layer holds neither prefill codes nor a decode layout")
        N, G, g = self.out_features, d.groups, d.group_size
        tiles = (N + 15) // 16
        rec = d.record_bytes                                                # per (tile, block): nibbles, then scales
        nib = d.nib.view(tiles, -1, rec)[:, :, :2048].reshape(tiles, -1, 8, 4, 2, 2, 4)   # [tile, unit, gid, tq, pk, pr, j]
        units = nib.shape[1]
        both = torch.stack(((nib & 15), (nib >> 4)), dim=-1).to(torch.int8)  # [..., j, hl]: low nibble k, high k + 32
        both = ((both ^ 8) - 8)                                              # two's-complement nibble -> int8
        codes = both.permute(0, 5, 2, 1, 7, 4, 3, 6).reshape(tiles * 16, units * 64)[:N]
        ld = max((G * g + 127) // 128 * 128, 128)
        out = torch.zeros((N, ld), dtype=torch.int8, device=d.nib.device)
        out[:, : G * g] = codes[:, : G * g]
        self.w_codes = out
        if d.Ks:
            sal = d.sal.view(d.dtype).view(tiles, -1, 8, 4, 2, 2, 2)        # [tile, step, gid, tq, pk, pr, h]
            steps = sal.shape[1]
            rows = sal.permute(0, 5, 2, 1, 4, 3, 6).reshape(tiles * 16, steps * 16)[:N]
            ws = torch.zeros((N, (d.Ks + 7) // 8 * 8), dtype=d.dtype, device=d.nib.device)
            ws[:, : d.Ks] = rows[:, : d.Ks]
            self.w_sal = ws
        if self.w_scales is None:           # fp32 group-major table from the scale slots behind every record's nibbles
            slots = (rec - 2048) // 32
            sc = d.nib.view(tiles, -1, rec)[:, :, 2048:].contiguous().view(d.dtype).view(tiles, -1, slots, 16)
            if g >= 256:                    # one slot per block, a group spans g / 256 blocks: take its first block
                tab = sc[:, :: g // 256, 0, :].permute(1, 0, 2).reshape(-1, tiles * 16)
            else:                           # block b, slot s holds group b * (256 / g) + s
                tab = sc.permute(1, 2, 0, 3).reshape(-1, tiles * 16)
            full = torch.zeros((max(G, 1), max((N + 3) // 4 * 4, 4)), dtype=torch.float32, device=d.nib.device)
            full[:G, :N] = tab[:G, :N].float()
            self.w_scales = full

    def device_weight_bytes(self) -> int:
        """Bytes this layer's weight operands occupy on the device right now (all resident copies)."""
        n = 0
        for t in (self.w_codes, self.w_scales, self.w_sal, self._weight):
            if t is not None:
                n += t.numel() * t.element_size()
        if self.w_decode is not None:
            n += self.w_decode.nib.numel() + (0 if self.w_decode.sal is None else self.w_decode.sal.numel())
        return n

    def _forward_reference(self, x2, act_perm):
        q_x = self._fake_quant_nonsalient(x2, self.act_quant_name, act_perm)
        w = self._weight
        cdt = w.dtype if w.dtype in (torch.float16, torch.bfloat16) else torch.float16
        # F.linear(q_x, weight, bias) (fake_quant.py:306) as the pure 16-bit tcgen05 GEMM; fp32 modules (OPT, SURVEY
        # F7) are cast to fp16 at the boundary -- stated tolerance in tests/DESIGN.md
        y = ops.mixed_gemm(None, None, self.bias, a_sal=q_x if q_x.dtype == cdt else q_x.to(cdt),
                           w_sal=w if w.dtype == cdt else self._weight16(cdt))
        return y if y.dtype == x2.dtype else y.to(x2.dtype)        # the reference's y lives in x's dtype

    def _weight16(self, cdt):
        cache = getattr(self, "_w16_cache", None)
        if cache is None or cache.dtype != cdt or cache.device != self._weight.device:
            cache = self._weight.to(cdt)
            object.__setattr__(self, "_w16_cache", cache)
        return cache

    def _quantize_output(self, y, out_perm=None):
        # fake_quant.py:309-316: the K-sized input mask indexes y's columns -> only valid when N == K
        if self.salient_indices is not None and self.full_out_features != self.in_features:
            raise IndexError(f"The shape of the mask [{self.in_features}] at index 0 does not match the shape of the "
                             f"indexed tensor [{y.shape[0]}, {self.full_out_features}] at index 1")
        return self._fake_quant_nonsalient(y.contiguous(), self.output_quant_name, out_perm)

    # ---- offline weight path
    @staticmethod
    @torch.no_grad()
    def from_float(module, weight_quant="per_channel", act_quant="per_token", quantize_output=False, importance=None,
                   salient_prop=0, quant_bits=4, group_size=128, mode="reference", sort="max", sort_key=None,
                   weight_perm=None, tp_group=None, tp_gather_output=True, int_fraction=None):
        assert isinstance(module, torch.nn.Linear)                                    # fake_quant.py:335
        _check_choice(weight_quant, _WEIGHT_QUANTS, "weight_quant")                   # ValueError, fake_quant.py:361
        w = module.weight.data
        if not w.is_cuda:
            raise RuntimeError("W4A4Linear.from_float needs the module on a CUDA device (no CPU fallback)")
        new = W4A4Linear(module.in_features, module.out_features, module.bias is not None, act_quant=act_quant,
                         quantize_output=quantize_output, importance=importance, salient_prop=salient_prop,
                         quant_bits=quant_bits, group_size=group_size, mode=mode, sort=sort, device="meta",
                         dtype=w.dtype)
        new._buffers["_weight"] = None
        new.weight_quant_name = weight_quant
        sal, keep = new._device_indices(w.device)
        _validate_columns(weight_perm, module.in_features, "weight_perm")
        if mode == "fused":
            if weight_quant != "per_group" or act_quant != "per_group":
                raise ValueError("mode='fused' needs weight_quant='per_group' and act_quant='per_group'")
            qmax = 2 ** (quant_bits - 1) - 1
            if not 2 <= quant_bits <= 8 or qmax * qmax * group_size >= 2 ** 22:
                # the GEMM turns int32 group sums into floats with the 1.5 * 2^23 bias pattern: exact only below 2^22
                raise ValueError(f"mode='fused' needs 2 <= quant_bits <= 8 and qmax^2 * group_size < 2^22 "
                                 f"(quant_bits={quant_bits}, group_size={group_size})")
            new.int_fraction = int_fraction
            new._pack_fused(w, sort_key if sort_key is not None else importance)
        else:
            if weight_quant == "per_group":       # sorts ALL K columns by the weight's own abs-max (:164-167)
                wq = quantize_weight_per_group_absmax_sort(w, quant_bits, group_size, perm=weight_perm)
            elif weight_quant == "per_channel":   # the reference mutates module.weight here; so do we
                wq = quantize_weight_per_channel_absmax(w if sal is None else w.clone(), quant_bits)
            else:
                wq = quantize_weight_per_tensor_absmax(w if sal is None else w.clone(), quant_bits)
            if sal is not None:                   # FP originals back into the salient columns (:363-365)
                s64 = sal.long()
                wq[:, s64] = module.weight.data[:, s64]
            new._buffers["_weight"] = wq
        new._buffers["bias"] = None if module.bias is None else module.bias.data        # shape [N] (:369-370)
        if tp_group is not None:
            import torch.distributed as dist
            new.shard_(dist.get_rank(tp_group), dist.get_world_size(tp_group), tp_group, tp_gather_output)
        return new

    def _pack_fused(self, w, key):
        """P0 weight packer: static shared permutation (non-salient columns stably sorted by ``key`` ascending),
        codes + group scales of W[:, perm] (== quantize_weight_per_group_absmax, fake_quant.py:30-53) and the
        salient columns in the weight dtype."""
        dev = w.device
        sal, keep = self._device_indices(dev)
        if keep is None:
            keep = torch.arange(self.in_features, dtype=torch.int32, device=dev)
        if self.sort == "none" or key is None:
            perm = keep
        else:
            key = torch.as_tensor(key).to(device=dev, dtype=torch.float32)
            perm = ops.argsort_stable(key, subset=keep)
        _validate_columns(perm, self.in_features, "permutation")      # NaN-free total order by construction; cheap
        _validate_columns(sal, self.in_features, "salient_indices")
        cdt = w.dtype if w.dtype in (torch.float16, torch.bfloat16) else torch.float16
        G = (perm.numel() + self.group_size - 1) // self.group_size
        self.int_groups = _int_groups(G, self.group_size, self.int_fraction)
        wq = ops.quant_rows(w if w.dtype == cdt else w.to(cdt), perm, sal, self.group_size, self.quant_bits,
                            int_groups=self.int_groups)
        self.perm = perm
        self.w_codes, self.w_scales, self.w_sal = wq.codes, wq.scales, wq.sal
        self._set_partition_id()

    @torch.no_grad()
    def dequantized_weight(self):
        """[N, K] weight equivalent to the packed operands: code*scale at the original column of every permuted
        position, the FP salient columns elsewhere (API compatibility; not on the hot path)."""
        released = self.w_codes is None
        if released:
            self._restore_prefill_operand()
        try:
            return self._dequantized_weight()
        finally:
            if released:
                self.w_codes = None
                self.w_sal = None
                self.w_scales = None

    def _dequantized_weight(self):
        N, K, g = self.out_features, self.in_features, self.group_size
        Kn = self.perm.numel()
        G = (Kn + g - 1) // g
        Gi = G if self.int_groups is None else self.int_groups
        dt = self.w_sal.dtype if self.w_sal is not None else torch.float16
        out = torch.zeros((N, K), dtype=dt, device=self.w_codes.device)
        ks = 0 if self.sal_sorted is None else self.sal_sorted.numel()
        if Kn:
            deq = torch.zeros((N, G * g), dtype=dt, device=out.device)
            if Gi:
                sc = self.w_scales[:Gi, :N].t().to(dt)                                 # exact: scales are dt values
                deq[:, : Gi * g] = (self.w_codes[:, : Gi * g].view(N, Gi, g).to(dt) * sc[:, :, None]).view(N, Gi * g)
            if Gi < G:                                                                 # hybrid tail: stored dequantised
                t0 = (ks + 7) // 8 * 8
                deq[:, Gi * g:] = self.w_sal[:, t0: t0 + (G - Gi) * g]
            out[:, self.perm.long()] = deq[:, :Kn]
        if ks:
            out[:, self.sal_sorted.long()] = self.w_sal[:, :ks]
        if self.out_perm is not None:                 # rows are stored in emitted order: present the original order
            orig = torch.empty_like(out)
            orig[self.out_perm] = out
            out = orig
        return out

    def packed_bytes(self) -> int:
        """True storage of the layer (the reference's get_model_size stops counting quantised linears)."""
        if self.mode == "fused" and self.w_codes is not None:
            Kn = self.perm.numel()      # at rest every quantised column is an int4 nibble (a byte when quant_bits > 4)
            cb = (Kn + 1) // 2 if self.quant_bits <= 4 else Kn          # + one 16-bit scale per group
            n = self.out_features * cb + self.out_features * ((Kn + self.group_size - 1) // self.group_size) * 2
            if self.w_sal is not None:
                n += self.out_features * self.sal_sorted.numel() * self.w_sal.element_size()
            n += Kn * 4
        else:
            n = self._weight.numel() * self._weight.element_size()
        if self.bias is not None:
            n += self.bias.numel() * self.bias.element_size()
        return n

    # ---- packed checkpoint.  The reference never serialises its W4A4 layers (it re-quantises on
    # every run, 54 s for OPT-1.3B); the INT8 analogue is examples/export_int8_model.py:45-56.
    @torch.no_grad()
    def packed_state(self) -> Dict[str, object]:
        """Everything needed to rebuild this layer without the float weight: int4 nibbles (mq_pack_int4) + group
        scales in the 16-bit compute dtype + 16-bit salient slice + permutation (fused mode), or the literal
        fake-quantised weight (reference mode).  Tensors are returned on the CPU."""
        if self.tp_world != 1:
            raise RuntimeError("save the unsharded layer and shard after loading (shard_ slices packed rows)")
        cpu = lambda t: None if t is None else t.detach().cpu()
        meta = dict(format=PACKED_FORMAT, in_features=self.in_features, out_features=self.out_features,
                    group_size=self.group_size, quant_bits=self.quant_bits, mode=self.mode, sort=self.sort,
                    act_quant=self.act_quant_name, weight_quant=self.weight_quant_name,
                    output_quant=self.output_quant_name, int_groups=self.int_groups, int_fraction=self.int_fraction,
                    input_folded=self.input_folded)
        st = {"meta": meta, "bias": cpu(self.bias), "salient_indices": cpu(self.salient_indices),
              "out_perm": cpu(self.out_perm)}
        if self.mode == "fused" and self.w_codes is None and self.w_decode is not None:
            self._restore_prefill_operand()          # the checkpoint format is the row-major nibble format
        if self.mode == "fused" and self.w_codes is not None:
            sdt = self.w_sal.dtype if self.w_sal is not None else torch.float32
            sc16 = self.w_scales.to(sdt)
            if not torch.equal(sc16.float(), self.w_scales):        # scales are values of the compute dtype
                sc16 = self.w_scales
            if self.quant_bits <= 4:        # nibbles hold [-8, 7]: wider codes are stored as bytes
                st["w_packed"] = cpu(ops.pack_int4(self.w_codes))
            else:
                st["w_codes8"] = cpu(self.w_codes)
            st.update(w_scales=cpu(sc16), w_sal=cpu(self.w_sal), perm=cpu(self.perm))
        else:
            st["weight"] = cpu(self._weight)
        return st

    @staticmethod
    @torch.no_grad()
    def from_packed(state, device="cuda"):
        """Inverse of :meth:`packed_state`: no quantisation work, only mq_unpack_int4 of the nibbles."""
        meta = state["meta"]
        if meta.get("format") != PACKED_FORMAT:
            raise ValueError(f"unknown packed W4A4 format: {meta.get('format')}")
        dev = torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("W4A4Linear.from_packed needs a CUDA device (no CPU fallback)")
        new = W4A4Linear(meta["in_features"], meta["out_features"], state["bias"] is not None,
                         act_quant=meta["act_quant"], quantize_output=meta["output_quant"] != "None",
                         quant_bits=meta["quant_bits"], group_size=meta["group_size"], mode=meta["mode"],
                         sort=meta["sort"], device="meta")
        new.weight_quant_name = meta["weight_quant"]
        new.salient_indices = state["salient_indices"]
        new.int_groups, new.int_fraction = meta["int_groups"], meta["int_fraction"]
        put = lambda t: None if t is None else t.to(dev)
        new._buffers["bias"] = put(state["bias"])
        new._buffers["_weight"] = None
        K = meta["in_features"]
        _validate_columns(state.get("perm"), K, "checkpoint permutation")
        _validate_columns(state.get("salient_indices"), K, "checkpoint salient_indices")
        if "w_packed" in state or "w_codes8" in state:
            if "w_packed" in state:
                if meta["quant_bits"] > 4:
                    raise ValueError("checkpoint stores int4 nibbles but declares quant_bits > 4")
                new.w_codes = ops.unpack_int4(put(state["w_packed"]).contiguous())
            else:
                new.w_codes = put(state["w_codes8"]).contiguous()
            new.w_scales = put(state["w_scales"]).float().contiguous()
            new.w_sal, new.perm = put(state["w_sal"]), put(state["perm"])
            new.out_perm = put(state.get("out_perm"))
            new._set_partition_id()
            if meta.get("input_folded"):
                new._device_indices(dev)
                new.fold_input_()
        else:
            new._buffers["_weight"] = put(state["weight"])
        new._device_indices(dev)
        return new

    def _set_partition_id(self):
        import hashlib
        sal, _ = self._device_indices(self.perm.device)
        h = hashlib.sha1(self.perm.cpu().numpy().tobytes())
        h.update(b"|" if sal is None else sal.cpu().numpy().tobytes())
        self._partition_id = (h.hexdigest(), self.group_size, self.quant_bits, self.int_groups)

    def __repr__(self):
        return (f"W4A4Linear({self.in_features}, {self.out_features}, bias={self.bias is not None}, "
                f"weight_quant={self.weight_quant_name}, act_quant={self.act_quant_name}, "
                f"output_quant={self.output_quant_name}, salient_indices={self.salient_indices})")


# ------------------------------------------------------------------------------------------- model walkers
def _importance(input_feat, key, guarded):
    """sum over calibration blocks of mean|x| (fake_quant.py:396, :486).  The OPT walker has no guard and raises
    KeyError / TypeError on a missing entry, the Llama walker maps an empty dict to None."""
    if guarded and not input_feat:
        return None
    return sum(input_feat[key]).float()


def _swap(parent, attr, prefix, input_feat, guarded, quantize_output, kw, extra):
    imp = _importance(input_feat, prefix + "." + attr, guarded)
    sk = extra.get("sort_keys")
    setattr(parent, attr, W4A4Linear.from_float(
        getattr(parent, attr), quantize_output=quantize_output, importance=imp,
        sort_key=None if sk is None else sk.get(prefix + "." + attr),
        mode=extra.get("mode", "reference"), sort=extra.get("sort", "max"), tp_group=extra.get("tp_group"),
        int_fraction=extra.get("int_fraction"), **kw))


@torch.no_grad()
def fold_mlp_permutation(gate_proj, up_proj, down_proj):
    """fold down_proj's static input permutation into its producers.  down_proj consumes
    act(gate(x)) * up(x), an elementwise function of the two outputs, so emitting gate / up channels in the order
    [perm_down | salient_down] makes down_proj's activation arrive already partitioned: its quantiser reads contiguous
    16-byte vectors instead of gathering 2-byte elements (K1 contiguous mode, 1.7x faster at K = 11008) and the model
    output is bit-identical (every dot product is unchanged).  Reference anchor: smooth.py:126-141 (which producer
    feeds which linears), fake_quant.py:291-301 (the gather this removes)."""
    sigma = down_proj.fold_input_()
    gate_proj.permute_outputs_(sigma)
    up_proj.permute_outputs_(sigma)
    return sigma


def _mark_shared(parent, attrs):
    """Layers of `attrs` that ended up with the same K partition share one quantised activation per forward."""
    mods = [getattr(parent, a) for a in attrs]
    for m in mods:
        pid = getattr(m, "_partition_id", None)
        m._share_n = 1 if pid is None else sum(getattr(o, "_partition_id", None) == pid for o in mods)


def quantize_opt(model, weight_quant="per_tensor", act_quant="per_tensor", quantize_bmm_input=True, input_feat=None,
                 salient_prop=0, quant_bits=4, group_size=128, **extra):
    """fake_quant.py:377-461 -- OPT: fc1, fc2 per decoder layer; q/k/v (output-quantised when quantize_bmm_input)
    and out_proj per attention block."""
    from transformers.models.opt.modeling_opt import OPTAttention, OPTDecoderLayer
    kw = dict(weight_quant=weight_quant, act_quant=act_quant, salient_prop=salient_prop, quant_bits=quant_bits,
              group_size=group_size)
    for name, m in list(model.model.named_modules()):
        prefix = "model." + name
        if isinstance(m, OPTDecoderLayer):
            for attr in ("fc1", "fc2"):
                _swap(m, attr, prefix, input_feat, False, False, kw, extra)
                _mark_shared(m, (attr,))
        elif isinstance(m, OPTAttention):
            for attr in ("q_proj", "k_proj", "v_proj"):
                _swap(m, attr, prefix, input_feat, False, quantize_bmm_input, kw, extra)
            _swap(m, "out_proj", prefix, input_feat, False, False, kw, extra)
            _mark_shared(m, ("q_proj", "k_proj", "v_proj"))
            _mark_shared(m, ("out_proj",))
    if extra.get("fuse_norm") and extra.get("mode") == "fused":
        # pre-norm OPT: self_attn_layer_norm -> q/k/v, final_layer_norm -> fc1 (smooth.py:144-156)
        for m in model.model.modules():
            if isinstance(m, OPTDecoderLayer) and getattr(m, "do_layer_norm_before", False):
                fuse_norm_(m, "self_attn_layer_norm", [m.self_attn.q_proj, m.self_attn.k_proj, m.self_attn.v_proj],
                           "layer")
                fuse_norm_(m, "final_layer_norm", [m.fc1], "layer")
    return model


def quantize_llama_like(model, weight_quant="per_channel", act_quant="per_token", quantize_bmm_input=False,
                        input_feat=None, salient_prop=0, quant_bits=4, group_size=128, **extra):
    """fake_quant.py:464-561 -- Llama / Mistral: gate, up, down per MLP; q/k/v (+ optional output quant), o_proj."""
    from transformers.models.llama.modeling_llama import LlamaAttention, LlamaMLP
    from transformers.models.mistral.modeling_mistral import MistralAttention, MistralMLP
    kw = dict(weight_quant=weight_quant, act_quant=act_quant, salient_prop=salient_prop, quant_bits=quant_bits,
              group_size=group_size)
    for name, m in list(model.model.named_modules()):
        prefix = "model." + name
        if isinstance(m, (LlamaMLP, MistralMLP)):
            for attr in ("gate_proj", "up_proj", "down_proj"):
                _swap(m, attr, prefix, input_feat, True, False, kw, extra)
            if extra.get("fold") and extra.get("mode") == "fused" and extra.get("tp_group") is None:
                fold_mlp_permutation(m.gate_proj, m.up_proj, m.down_proj)
            _mark_shared(m, ("gate_proj", "up_proj"))
            _mark_shared(m, ("down_proj",))
        elif isinstance(m, (LlamaAttention, MistralAttention)):
            for attr in ("q_proj", "k_proj", "v_proj"):
                _swap(m, attr, prefix, input_feat, True, quantize_bmm_input, kw, extra)
            _swap(m, "o_proj", prefix, input_feat, True, False, kw, extra)
            _mark_shared(m, ("q_proj", "k_proj", "v_proj"))
            _mark_shared(m, ("o_proj",))
    if extra.get("fuse_norm") and extra.get("mode") == "fused":
        # input_layernorm -> q/k/v and post_attention_layernorm -> gate/up (smooth.py:126-141)
        for m in model.model.modules():
            attn, mlp = getattr(m, "self_attn", None), getattr(m, "mlp", None)
            if attn is not None and mlp is not None and hasattr(m, "input_layernorm"):
                fuse_norm_(m, "input_layernorm", [attn.q_proj, attn.k_proj, attn.v_proj], "rms")
                fuse_norm_(m, "post_attention_layernorm", [mlp.gate_proj, mlp.up_proj], "rms")
    return model


def quantize_model(model, weight_quant="per_channel", act_quant="per_token", quantize_bmm_input=False,
                   input_feat=None, salient_prop=None, quant_bits=4, group_size=128, min_prop=0, max_prop=0, **extra):
    """fake_quant.py:734-799 -- dispatch on the model family.  ``min_prop`` / ``max_prop`` are accepted and ignored
    like in the reference.  Mixtral / Falcon walkers are out of scope: ValueError."""
    if input_feat is None:
        input_feat = {}
    from transformers.models.llama.modeling_llama import LlamaPreTrainedModel
    from transformers.models.mistral.modeling_mistral import MistralPreTrainedModel
    from transformers.models.opt.modeling_opt import OPTPreTrainedModel
    kw = dict(weight_quant=weight_quant, act_quant=act_quant, quantize_bmm_input=quantize_bmm_input,
              input_feat=input_feat, salient_prop=salient_prop, quant_bits=quant_bits, group_size=group_size)
    if isinstance(model, OPTPreTrainedModel):
        return quantize_opt(model, **kw, **extra)
    if isinstance(model, (LlamaPreTrainedModel, MistralPreTrainedModel)):
        return quantize_llama_like(model, **kw, **extra)
    raise ValueError(f"Unsupported model type: {type(model)}")


def model_packed_bytes(model) -> Dict[str, int]:
    """Real storage of the swapped linears vs everything else (replaces the analytic get_model_size)."""
    lin = sum(m.packed_bytes() for m in model.modules() if isinstance(m, W4A4Linear))
    other = sum(p.numel() * p.element_size() for p in model.parameters())
    return {"w4a4_linear_bytes": lin, "other_parameter_bytes": other, "total_bytes": lin + other}


# ------------------------------------------------------------------------------------------- packed checkpoints
def save_packed_model(model, path=None) -> Dict[str, object]:
    """Checkpoint of a quantised model: the packed state of every W4A4Linear (by module name) plus the ordinary
    state_dict entries of everything else.  Written with torch.save when ``path`` is given; returned either way."""
    linears = {n: m.packed_state() for n, m in model.named_modules() if isinstance(m, W4A4Linear)}
    rest = {k: v.detach().cpu() for k, v in model.state_dict().items()
            if not any(k == n or k.startswith(n + ".") for n in linears)}
    ckpt = {"format": PACKED_FORMAT, "linears": linears, "rest": rest}
    if path is not None:
        torch.save(ckpt, path)
    return ckpt


def load_packed_model(model, ckpt, device="cuda"):
    """Turn a float skeleton of the same architecture (any weights, any real device) into the quantised
    model of the checkpoint: every named linear becomes ``W4A4Linear.from_packed`` (no re-quantisation), every other
    tensor is loaded from the checkpoint.  ``ckpt`` is a path or the dict returned by save_packed_model."""
    if not isinstance(ckpt, dict):
        ckpt = torch.load(ckpt, map_location="cpu", weights_only=False)
    if ckpt.get("format") != PACKED_FORMAT:
        raise ValueError(f"unknown packed checkpoint format: {ckpt.get('format')}")
    if any(p.is_meta for p in model.parameters()):
        raise ValueError("load_packed_model needs a materialised skeleton: non-persistent buffers (rotary tables) are "
                         "not part of any checkpoint and would stay uninitialised on a meta skeleton")
    model.to(device)
    for name, st in ckpt["linears"].items():
        parent_name, _, attr = name.rpartition(".")
        parent = model.get_submodule(parent_name) if parent_name else model
        old = getattr(parent, attr)                                  # KeyError / AttributeError on a foreign skeleton
        if not isinstance(old, (nn.Linear, W4A4Linear)):
            raise TypeError(f"{name} is a {type(old).__name__}, not a linear layer")
        setattr(parent, attr, W4A4Linear.from_packed(st, device))
    own = {k for n in ckpt["linears"] for k in model.state_dict() if k == n or k.startswith(n + ".")}
    missing, unexpected = model.load_state_dict(ckpt["rest"], strict=False)
    missing = [k for k in missing if k not in own]
    if missing or unexpected:
        raise KeyError(f"packed checkpoint does not match the skeleton: missing {missing}, unexpected {unexpected}")
    return model
