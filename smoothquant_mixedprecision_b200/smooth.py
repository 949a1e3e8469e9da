"""SmoothQuant scale migration (offline weight transform), API-compatible with the reference's
``smoothquant/smooth.py``.

The smoothing scale is folded into the preceding LayerNorm / RMSNorm and the consuming linears once, before
quantisation: there is no run-time component, so this stays a handful of torch tensor ops executed on
whatever device the model lives on.  It is bit-identical to the reference because it evaluates the same expression
in the same dtype:  s = clamp(act^alpha / clamp(max_fcs max_n |W|, 1e-5)^(1-alpha), 1e-5)   (smooth.py:26-39, :56-67).
The optional run-time divisor of the activation kernel (``ops.quant_rows(..., smooth=s)``) exists for producers that
are not a norm layer and is off by default.

Covered model families: OPT (smooth.py:77-90) and Llama / Mistral (smooth.py:126-141) -- the ones BASELINE.json
names.  Bloom / Falcon / Mixtral branches are out of scope.
"""
from __future__ import annotations

import torch
from torch import nn

__all__ = ["smooth_ln_fcs", "smooth_ln_fcs_llama_like", "smooth_lm", "smoothing_scales"]


def _as_list(fcs):
    return fcs if isinstance(fcs, list) else [fcs]


@torch.no_grad()
def smoothing_scales(fcs, act_scales, alpha=0.5):
    """Per-input-channel migration scale shared by every linear in ``fcs`` (they read the same activation)."""
    ref = fcs[0].weight
    a = act_scales.to(device=ref.device, dtype=ref.dtype)
    col_max = torch.stack([fc.weight.abs().amax(dim=0) for fc in fcs]).amax(dim=0).clamp(min=1e-5)
    return (a.pow(alpha) / col_max.pow(1 - alpha)).clamp(min=1e-5).to(ref.device).to(ref.dtype)


def _validate(norm, fcs, act_scales, norm_types):
    assert isinstance(norm, norm_types)
    for fc in fcs:
        assert isinstance(fc, nn.Linear)
        assert norm.weight.numel() == fc.in_features == act_scales.numel()


@torch.no_grad()
def smooth_ln_fcs(ln, fcs, act_scales, alpha=0.5):
    """LayerNorm producer (smooth.py:19-45): weight AND bias of the norm are divided by s."""
    fcs = _as_list(fcs)
    _validate(ln, fcs, act_scales, nn.LayerNorm)
    s = smoothing_scales(fcs, act_scales, alpha)
    ln.weight.div_(s)
    ln.bias.div_(s)
    for fc in fcs:
        fc.weight.mul_(s.view(1, -1))


def _rmsnorm_types():
    from transformers.models.llama.modeling_llama import LlamaRMSNorm
    from transformers.models.mistral.modeling_mistral import MistralRMSNorm
    return (LlamaRMSNorm, MistralRMSNorm)


@torch.no_grad()
def smooth_ln_fcs_llama_like(ln, fcs, act_scales, alpha=0.5):
    """RMSNorm producer (smooth.py:49-71): only the norm weight exists."""
    fcs = _as_list(fcs)
    _validate(ln, fcs, act_scales, _rmsnorm_types())
    s = smoothing_scales(fcs, act_scales, alpha)
    ln.weight.div_(s)
    for fc in fcs:
        fc.weight.mul_(s.view(1, -1))


@torch.no_grad()
def smooth_lm(model, scales, alpha=0.5):
    """smooth.py:75-160 for OPT and Llama / Mistral decoder layers.  ``scales[name]`` is the per-channel max |x| at
    the input of linear ``name`` (calibration.get_act_scales, calibration.py:13-51)."""
    from transformers.models.llama.modeling_llama import LlamaDecoderLayer
    from transformers.models.mistral.modeling_mistral import MistralDecoderLayer
    from transformers.models.opt.modeling_opt import OPTDecoderLayer
    for name, layer in model.named_modules():
        if isinstance(layer, OPTDecoderLayer):
            att = layer.self_attn
            smooth_ln_fcs(layer.self_attn_layer_norm, [att.q_proj, att.k_proj, att.v_proj],
                          scales[name + ".self_attn.q_proj"], alpha)
            smooth_ln_fcs(layer.final_layer_norm, layer.fc1, scales[name + ".fc1"], alpha)
        elif isinstance(layer, (LlamaDecoderLayer, MistralDecoderLayer)):
            att = layer.self_attn
            smooth_ln_fcs_llama_like(layer.input_layernorm, [att.q_proj, att.k_proj, att.v_proj],
                                     scales[name + ".self_attn.q_proj"], alpha)
            smooth_ln_fcs_llama_like(layer.post_attention_layernorm, [layer.mlp.gate_proj, layer.mlp.up_proj],
                                     scales[name + ".mlp.gate_proj"], alpha)
