"""GPU calibration collectors: the producers of the hot path's inputs.

The reference gathers its statistics with forward hooks on every ``nn.Linear`` and a torch reduction per call, copied
to the CPU each time:

* ``get_act_scales``  -- running per-channel max |x|   (smoothquant/calibration.py:13-51)  -> ``smooth_lm`` scales
* ``get_calib_feat``  -- per-block per-channel mean |x| (run_experiments/run_experiments.py:55-84) -> ``input_feat``
  (``importance = sum(input_feat[name])``, fake_quant.py:396 / :486)

Here the same hooks call the column-statistic kernel K0 (``mq_colstat``, one pass over x, fp32 accumulation) and keep
the running values on the device; nothing is copied to the host until the end.  Dataset / tokenizer handling is out
of scope (no network datasets on this box): the collectors take an iterable of ``input_ids`` batches instead of a
tokenizer + dataset path, everything else -- hook placement, keys (``named_modules()`` names), reductions, return
types -- follows the reference.  ``collect_sort_keys`` additionally returns the static-permutation keys of
mode="fused" (``max`` / ``mean3std`` / ``argmax``; the last two are parity unpinned).
"""
from __future__ import annotations

import functools
from typing import Dict, Iterable, List

import torch
from torch import nn

from . import ops

__all__ = ["get_act_scales", "get_calib_feat", "collect_sort_keys"]


def _hook_linears(model, fn):
    hooks = []
    for name, m in model.named_modules():
        if isinstance(m, nn.Linear):                                   # calibration.py:33-37 / run_experiments.py:66-70
            hooks.append(m.register_forward_hook(functools.partial(fn, name=name)))
    return hooks


def _first(x):
    return x[0] if isinstance(x, tuple) else x                          # calibration.py:28-30


@torch.no_grad()
def _run(model, batches: Iterable[torch.Tensor]):
    device = next(model.parameters()).device
    if device.type != "cuda":
        raise RuntimeError("the calibration collectors run mq_colstat on a CUDA device (no CPU fallback)")
    model.eval()
    for ids in batches:
        model(ids.to(device))


def get_act_scales(model, batches: Iterable[torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Running per-channel abs-max of every linear's input (calibration.py:13-51): ``{name: float32 [K] on the CPU}``.
    Exact (a maximum has no rounding), so it equals the reference's result on the same batches bit for bit."""
    scales: Dict[str, torch.Tensor] = {}

    def hook(m, x, y, name):
        cur = ops.colstat(_first(x).detach(), "max")                    # view(-1, K).abs().max(dim=0)  (:19-21)
        scales[name] = torch.maximum(scales[name], cur) if name in scales else cur   # :22-25

    hooks = _hook_linears(model, hook)
    try:
        _run(model, batches)
    finally:
        for h in hooks:
            h.remove()
    return {k: v.cpu() for k, v in scales.items()}


def get_calib_feat(model, batches: Iterable[torch.Tensor]) -> Dict[str, List[torch.Tensor]]:
    """Per-block per-channel mean |x| of every linear's input (run_experiments.py:55-84): ``{name: [tensor [K] per
    block]}`` in the activation dtype on the CPU, the ``input_feat`` argument of the walkers.  The reference reduces
    in the activation dtype (torch accumulates in fp32 and rounds once); K0 accumulates in fp32 and the result is
    rounded to the activation dtype here, so values agree up to the summation order (<= 1 ulp of the dtype)."""
    feat: Dict[str, List[torch.Tensor]] = {}

    def hook(m, x, y, name):
        t = _first(x).detach()
        feat.setdefault(name, []).append(ops.colstat(t, "mean").to(t.dtype))           # :60-64

    hooks = _hook_linears(model, hook)
    try:
        _run(model, batches)
    finally:
        for h in hooks:
            h.remove()
    return {k: [t.cpu() for t in v] for k, v in feat.items()}


def collect_sort_keys(model, batches: Iterable[torch.Tensor], kind: str = "max") -> Dict[str, torch.Tensor]:
    """Static-permutation keys of mode="fused" per linear input, keyed like the walkers expect
    (``sort_keys={"model.layers.0.mlp.gate_proj": key, ...}``).  ``max``: running abs-max (== get_act_scales);
    ``mean3std`` / ``argmax``: the statistic of the concatenated batches is not decomposable, so they are taken from
    the LAST batch seen (README.md:36-37 gives no definition over several batches; parity unpinned)."""
    if kind not in ("max", "mean3std", "argmax"):
        raise ValueError(f"Invalid sort: {kind}")
    keys: Dict[str, torch.Tensor] = {}

    def hook(m, x, y, name):
        cur = ops.colstat(_first(x).detach(), kind)
        keys[name] = torch.maximum(keys[name], cur) if (kind == "max" and name in keys) else cur

    hooks = _hook_linears(model, hook)
    try:
        _run(model, batches)
    finally:
        for h in hooks:
            h.remove()
    return {k: v.cpu() for k, v in keys.items()}
