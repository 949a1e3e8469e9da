"""Generate the golden fixtures in this directory by EXECUTING the real reference.

Run in the build container only (the reference tree does not exist on the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

It imports ``/root/reference/smoothquant`` unmodified, feeds seeded tensors through its quantiser functions,
``W4A4Linear`` and the model walkers, and stores inputs + outputs as small ``.pt`` files.  Because the reference
calls ``torch.argsort`` without ``stable=True`` (fake_quant.py:116, :167, :268) the permutation it actually used is
captured by wrapping ``torch.argsort`` while the reference runs; it is stored next to the outputs so that parity
tests can replay exactly the same partition.
"""
import os
import sys

os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference")

import torch  # noqa: E402
from torch import nn  # noqa: E402

import smoothquant.fake_quant as rfq  # noqa: E402  (the reference)
import smoothquant.smooth as rsm  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


class ArgsortTap:
    """Record every torch.argsort result produced while the reference runs."""

    def __enter__(self):
        self.calls = []
        self._orig = torch.argsort

        def tapped(*a, **k):
            out = self._orig(*a, **k)
            self.calls.append(out.clone())
            return out

        torch.argsort = tapped
        return self

    def __exit__(self, *exc):
        torch.argsort = self._orig


def synth(rows, cols, dtype, seed, outliers=True, zero_block=False, distinct=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(rows, cols, generator=g)
    scale = torch.exp(0.5 * torch.randn(cols, generator=g))
    if outliers:
        idx = torch.randperm(cols, generator=g)[: max(1, cols // 100)]
        scale[idx] *= torch.empty(idx.numel()).uniform_(10, 40, generator=g)
    if distinct:                      # tie-free column maxima even after rounding to fp16
        scale = scale * (1.0 + torch.arange(cols) / cols)
    x = x * scale
    if zero_block:
        x[:, : min(cols, 40)] = 0.0
    return x.to(dtype)


def quantiser_cases():
    out = []
    specs = [
        # rows, cols, g, dtype, seed, kwargs
        (7, 300, 128, torch.float16, 1, {}),
        (16, 256, 64, torch.float16, 2, {}),
        (5, 100, 128, torch.float16, 3, {}),
        (8, 2048, 1024, torch.float16, 4, {}),
        (9, 130, 4, torch.float16, 5, {}),
        (12, 384, 128, torch.bfloat16, 6, {}),
        (6, 200, 64, torch.bfloat16, 7, {"zero_block": True}),
        (10, 300, 128, torch.float32, 8, {}),
        (4, 512, 128, torch.float32, 9, {"zero_block": True}),
        (33, 777, 64, torch.float16, 10, {"distinct": True}),
        (3, 1100, 128, torch.float16, 11, {"zero_block": True}),
    ]
    for rows, cols, g, dtype, seed, kw in specs:
        x = synth(rows, cols, dtype, seed, **kw)
        case = {"x": x, "group_size": g, "n_bits": 4}
        with ArgsortTap() as tap:
            case["act_group_sort"] = rfq.quantize_activation_per_group_absmax_sort(x.clone(), 4, g)
        case["act_group_sort_perm"] = tap.calls[0]
        with ArgsortTap() as tap:
            case["weight_group_sort"] = rfq.quantize_weight_per_group_absmax_sort(x.clone(), 4, g)
        case["weight_group_sort_perm"] = tap.calls[0]
        case["act_group"] = rfq.quantize_activation_per_group_absmax(x.clone(), 4, g).clone()
        case["weight_group"] = rfq.quantize_weight_per_group_absmax(x.clone(), 4, g).clone()
        case["act_token"] = rfq.quantize_activation_per_token_absmax(x.clone(), 4)
        case["act_tensor"] = rfq.quantize_activation_per_tensor_absmax(x.clone(), 4)
        case["weight_channel"] = rfq.quantize_weight_per_channel_absmax(x.clone(), 4)
        case["weight_tensor"] = rfq.quantize_weight_per_tensor_absmax(x.clone(), 4)
        out.append(case)
    # a 3-D activation (shape must be preserved)
    x3 = synth(2 * 5, 192, torch.float16, 12).view(2, 5, 192)
    case = {"x": x3, "group_size": 64, "n_bits": 4}
    with ArgsortTap() as tap:
        case["act_group_sort"] = rfq.quantize_activation_per_group_absmax_sort(x3.clone(), 4, 64)
    case["act_group_sort_perm"] = tap.calls[0]
    case["act_group"] = rfq.quantize_activation_per_group_absmax(x3.clone(), 4, 64).clone()
    out.append(case)
    return out


def layer_cases():
    out = []
    specs = [
        # K, N, M, g, p, dtype, bias, quantize_output, act_quant, weight_quant, three_d, seed
        (256, 192, 24, 64, 0.05, torch.float16, True, False, "per_group", "per_group", False, 21),
        (256, 192, 24, 64, 0.05, torch.float32, True, False, "per_group", "per_group", False, 22),
        (320, 128, 17, 128, 0.10, torch.float16, False, False, "per_group", "per_group", True, 23),
        (128, 128, 16, 64, 0.05, torch.float32, True, True, "per_group", "per_group", False, 24),
        (128, 128, 16, 64, 0.0, torch.float16, True, True, "per_group", "per_group", False, 25),
        (192, 64, 8, 128, 0.0, torch.float16, True, False, "per_token", "per_channel", False, 26),
        (192, 64, 8, 128, 0.05, torch.float32, True, False, "per_tensor", "per_tensor", False, 27),
        (512, 256, 40, 128, 0.05, torch.bfloat16, True, False, "per_group", "per_group", False, 28),
    ]
    for K, N, M, g, p, dtype, bias, qo, aq, wq, three_d, seed in specs:
        gen = torch.Generator().manual_seed(seed)
        lin = nn.Linear(K, N, bias=bias)
        with torch.no_grad():
            lin.weight.copy_(torch.randn(N, K, generator=gen) * 0.05)
            if bias:
                lin.bias.copy_(torch.randn(N, generator=gen) * 0.1)
        lin = lin.to(dtype)
        x = synth(M, K, dtype, seed + 100)
        if three_d:
            x = x[:16].reshape(2, 8, K).contiguous()
        importance = x.reshape(-1, K).abs().float().mean(dim=0)
        w0 = lin.weight.data.clone()
        b0 = lin.bias.data.clone() if bias else None
        with ArgsortTap() as tap:
            mod = rfq.W4A4Linear.from_float(lin, weight_quant=wq, act_quant=aq, quantize_output=qo,
                                            importance=importance if p > 0 else None, salient_prop=p,
                                            quant_bits=4, group_size=g)
        init_sorts = [c for c in tap.calls]
        with ArgsortTap() as tap:
            y = mod(x.clone())
        fwd_sorts = [c for c in tap.calls]
        out.append({
            "K": K, "N": N, "group_size": g, "salient_prop": p, "bias": b0, "weight_fp": w0,
            "weight_q": mod.weight.data.clone(), "x": x, "y": y.clone(), "importance": importance,
            "salient_indices": None if mod.salient_indices is None else mod.salient_indices.clone(),
            "act_quant": aq, "weight_quant": wq, "quantize_output": qo,
            "init_argsorts": init_sorts, "fwd_argsorts": fwd_sorts, "repr": repr(mod),
        })
    return out


def hook_input_feat(model, ids_list):
    """run_experiments.get_calib_feat (run_experiments.py:55-84) with synthetic token blocks."""
    feat = {}
    hooks = []

    def hook(m, x, y, name):
        x = x[0] if isinstance(x, tuple) else x
        feat.setdefault(name, []).append(x.view(-1, x.shape[-1]).abs().mean(dim=0).cpu().detach())

    for name, m in model.named_modules():
        if isinstance(m, nn.Linear):
            hooks.append(m.register_forward_hook(lambda m, x, y, name=name: hook(m, x, y, name)))
    with torch.no_grad():
        for ids in ids_list:
            model(ids)
    for h in hooks:
        h.remove()
    return feat


def hook_act_scales(model, ids_list):
    """calibration.get_act_scales statistic (calibration.py:18-25) with synthetic token blocks."""
    scales = {}
    hooks = []

    def hook(m, x, y, name):
        x = x[0] if isinstance(x, tuple) else x
        cur = x.view(-1, x.shape[-1]).abs().detach().max(dim=0)[0].float().cpu()
        scales[name] = torch.max(scales[name], cur) if name in scales else cur

    for name, m in model.named_modules():
        if isinstance(m, nn.Linear):
            hooks.append(m.register_forward_hook(lambda m, x, y, name=name: hook(m, x, y, name)))
    with torch.no_grad():
        for ids in ids_list:
            model(ids)
    for h in hooks:
        h.remove()
    return scales


def model_cases():
    from transformers import LlamaConfig, LlamaForCausalLM, OPTConfig, OPTForCausalLM
    out = {}
    # --- tiny Llama: smooth_lm(alpha=.85) then quantize_model(per_group, per_group, 5% salient) ---
    torch.manual_seed(1234)
    cfg = LlamaConfig(vocab_size=128, hidden_size=64, intermediate_size=176, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=4, max_position_embeddings=64)
    model = LlamaForCausalLM(cfg).eval()
    gen = torch.Generator().manual_seed(7)
    calib = [torch.randint(0, 128, (1, 32), generator=gen) for _ in range(2)]
    ids = torch.randint(0, 128, (2, 24), generator=gen)
    state0 = {k: v.clone() for k, v in model.state_dict().items()}
    act_scales = hook_act_scales(model, calib)
    rsm.smooth_lm(model, act_scales, 0.85)
    state_smoothed = {k: v.clone() for k, v in model.state_dict().items()}
    input_feat = hook_input_feat(model, calib)
    with torch.no_grad():
        logits_fp = model(ids).logits.clone()
    with ArgsortTap() as tap:
        rfq.quantize_model(model, weight_quant="per_group", act_quant="per_group", input_feat=input_feat,
                           salient_prop=0.05, group_size=32)
    with torch.no_grad():
        logits_q = model(ids).logits.clone()
    out["llama"] = {"config": cfg.to_dict(), "state0": state0, "state_smoothed": state_smoothed,
                    "act_scales": act_scales, "alpha": 0.85, "input_feat": input_feat, "ids": ids,
                    "logits_fp_smoothed": logits_fp, "logits_w4a4": logits_q, "group_size": 32,
                    "salient_prop": 0.05,
                    "quant_weights": {n: m.weight.data.clone() for n, m in model.named_modules()
                                      if isinstance(m, rfq.W4A4Linear)},
                    "salient": {n: m.salient_indices.clone() for n, m in model.named_modules()
                                if isinstance(m, rfq.W4A4Linear)}}
    # --- tiny OPT: quantize_opt defaults quantize_bmm_input=True (fake_quant.py:381) ---
    torch.manual_seed(4321)
    ocfg = OPTConfig(vocab_size=128, hidden_size=64, ffn_dim=192, num_hidden_layers=2, num_attention_heads=4,
                     max_position_embeddings=64, word_embed_proj_dim=64)
    omodel = OPTForCausalLM(ocfg).eval()
    ostate0 = {k: v.clone() for k, v in omodel.state_dict().items()}
    oscales = hook_act_scales(omodel, calib)
    rsm.smooth_lm(omodel, oscales, 0.5)
    ostate_s = {k: v.clone() for k, v in omodel.state_dict().items()}
    ofeat = hook_input_feat(omodel, calib)
    rfq.quantize_opt(omodel, weight_quant="per_group", act_quant="per_group", input_feat=ofeat,
                     salient_prop=0.05, group_size=32)
    with torch.no_grad():
        ologits = omodel(ids).logits.clone()
    out["opt"] = {"config": ocfg.to_dict(), "state0": ostate0, "state_smoothed": ostate_s,
                  "act_scales": oscales, "alpha": 0.5, "input_feat": ofeat, "ids": ids,
                  "logits_w4a4": ologits, "group_size": 32, "salient_prop": 0.05,
                  "quant_weights": {n: m.weight.data.clone() for n, m in omodel.named_modules()
                                    if isinstance(m, rfq.W4A4Linear)},
                  "salient": {n: m.salient_indices.clone() for n, m in omodel.named_modules()
                              if isinstance(m, rfq.W4A4Linear)}}
    return out


def main():
    torch.set_num_threads(4)
    torch.save(quantiser_cases(), os.path.join(HERE, "quantisers.pt"))
    torch.save(layer_cases(), os.path.join(HERE, "layers.pt"))
    torch.save(model_cases(), os.path.join(HERE, "models.pt"))
    for f in ("quantisers.pt", "layers.pt", "models.pt"):
        print(f, os.path.getsize(os.path.join(HERE, f)) // 1024, "KiB")
    print("torch", torch.__version__)


if __name__ == "__main__":
    main()
