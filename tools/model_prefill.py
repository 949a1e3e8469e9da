"""Whole-model prefill through the drop-in API (BASELINE.json config 3): a random-init HF Llama-2-7B, smoothed with
``smooth_lm(alpha=0.85)`` and swapped to W4A4 MixedQuant linears with ``quantize_llama_like(mode="fused")``, runs a
full forward (embeddings, RMSNorm, rotary attention via stock PyTorch SDPA, SiLU MLP, fp16 lm_head) over
batch x seq tokens.  Only the linear layers are this repository's kernels; everything else is stock HF / PyTorch and
is reported as the non-W4A4 share.

    python tools/model_prefill.py [--layers 32] [--batch 16] [--seq 2048] [--steps 3] [--fp16-baseline]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def build_model(layers, dev, seed=0):
    import torch
    from transformers import LlamaConfig, LlamaForCausalLM
    cfg = LlamaConfig(vocab_size=32000, hidden_size=4096, intermediate_size=11008, num_hidden_layers=layers,
                      num_attention_heads=32, num_key_value_heads=32, max_position_embeddings=4096,
                      attn_implementation="sdpa")
    torch.manual_seed(seed)
    with torch.device(dev):
        model = LlamaForCausalLM(cfg).to(torch.float16)
    model.eval()
    return model


def synthetic_calibration(model, dev, seed=1):
    """Per-linear statistics shaped like the reference's hooks produce them: input_feat[name] = [mean|x| vector]
    (run_experiments.py:57-64) and act_scales[name] = max|x| vector (calibration.py:18-25).  Linears that read the
    same tensor share one vector, as they do in the reference."""
    import torch
    g = torch.Generator(device=dev).manual_seed(seed)
    feat, scales = {}, {}
    for name, m in model.model.named_modules():
        if isinstance(m, torch.nn.Linear):
            key = "model." + name
            parent, leaf = key.rsplit(".", 1)
            src = {"q_proj": "x", "k_proj": "x", "v_proj": "x", "gate_proj": "h", "up_proj": "h"}.get(leaf, leaf)
            tag = parent + ":" + src
            if tag not in feat:
                K = m.in_features
                c = torch.exp(0.5 * torch.randn(K, device=dev, generator=g))
                hot = torch.rand(K, device=dev, generator=g) < 0.01
                c = torch.where(hot, c * 20.0, c)
                feat[tag] = (c * 0.8, c * 4.0)
            mean_abs, max_abs = feat[tag]
            feat[key] = [mean_abs.cpu()]
            scales[key] = max_abs
    input_feat = {k: v for k, v in feat.items() if ":" not in k}
    return input_feat, scales


def run(layers=32, batch=16, seq=2048, steps=3, warmup=2, quantize=True, dev="cuda:0", fold=False, fuse_norm=False):
    import torch
    from smoothquant_mixedprecision_b200 import ops
    from smoothquant_mixedprecision_b200.fake_quant import quantize_llama_like, W4A4Linear
    from smoothquant_mixedprecision_b200.smooth import smooth_lm
    t0 = time.perf_counter()
    model = build_model(layers, dev)
    n_w4a4 = 0
    if quantize:
        input_feat, act_scales = synthetic_calibration(model, dev)
        smooth_lm(model, act_scales, 0.85)
        quantize_llama_like(model, "per_group", "per_group", input_feat=input_feat, salient_prop=0.05, group_size=128,
                            mode="fused", sort="max", sort_keys=act_scales, fold=fold, fuse_norm=fuse_norm)
        n_w4a4 = sum(isinstance(m, W4A4Linear) for m in model.modules())
        torch.cuda.empty_cache()
    setup_s = time.perf_counter() - t0
    ids = torch.randint(0, 32000, (batch, seq), device=dev, generator=torch.Generator(device=dev).manual_seed(3))

    @torch.no_grad()
    def step():
        out = model.model(input_ids=ids, use_cache=False).last_hidden_state   # decoder stack
        return model.lm_head(out[:, -1:, :])                                  # next-token logits of every sequence

    for _ in range(warmup):
        step()
    ops.GEMM_EVENTS, ops.QUANT_EVENTS = [], []
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        logits = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    gev, ops.GEMM_EVENTS = ops.GEMM_EVENTS, None
    qev, ops.QUANT_EVENTS = ops.QUANT_EVENTS, None
    gemm_ms = sum(a.elapsed_time(b) for a, b, _, _ in gev) / steps
    quant_ms = sum(a.elapsed_time(b) for a, b, _ in qev) / steps
    return {"workload": f"HF Llama-2-7B random-init ({layers} layers), smooth_lm alpha=0.85 + quantize_llama_like fused "
                        f"g=128 5% salient (fold={fold}, fuse_norm={fuse_norm}), full forward {batch}x{seq} tokens" if quantize else
                        f"HF Llama-2-7B random-init ({layers} layers) fp16 nn.Linear (stock PyTorch), {batch}x{seq}",
            "tokens_per_s": batch * seq / (ms * 1e-3), "ms_per_step": ms, "w4a4_layers": n_w4a4,
            "w4a4_gemm_ms": gemm_ms, "w4a4_quant_ms": quant_ms,
            "w4a4_share": (gemm_ms + quant_ms) / ms if ms else None,
            "logits_finite": bool(torch.isfinite(logits.float()).all()), "setup_s": setup_s}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--seq", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--fp16-baseline", action="store_true", help="also time the un-quantised fp16 model")
    ap.add_argument("--serving-options", action="store_true", help="also time fold=True, fuse_norm=True")
    args = ap.parse_args()
    print(json.dumps(run(args.layers, args.batch, args.seq, args.steps)), flush=True)
    if args.serving_options:
        import torch
        torch.cuda.empty_cache()
        print(json.dumps(run(args.layers, args.batch, args.seq, args.steps, fold=True, fuse_norm=True)), flush=True)
    if args.fp16_baseline:
        import torch
        torch.cuda.empty_cache()
        print(json.dumps(run(args.layers, args.batch, args.seq, args.steps, quantize=False)), flush=True)


if __name__ == "__main__":
    main()
