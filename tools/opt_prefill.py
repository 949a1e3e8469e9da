"""BASELINE.json configs 1 and 2 through the drop-in API on one B200: a random-init HF OPT model, calibrated with
the GPU collectors (calibration.get_calib_feat / collect_sort_keys on synthetic token blocks, run_experiments.py:55-84),
swapped by ``quantize_opt`` (q/k/v outputs fake-quantised as in the reference, fake_quant.py:381, :423) in
``mode="reference"`` (literal semantics) and ``mode="fused"`` (static shared permutation), full forward over
batch x seq tokens.  Reports tokens/s of both modes and of the stock fp16 model, and the logit gap between the modes
(a quantisation-noise-level quantity, SURVEY 8(a')).

    python tools/opt_prefill.py --config 1     # OPT-125m, g=128, 5 % salient, max sort, seq 512
    python tools/opt_prefill.py --config 2     # OPT-1.3B, g=64, 10 % salient, mean+3STD sort, prefill 2048
"""
import argparse
import copy
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CONFIGS = {
    1: dict(name="OPT-125m", hidden=768, ffn=3072, layers=12, heads=12, g=128, p=0.05, sort="max", batch=1, seq=512),
    2: dict(name="OPT-1.3B", hidden=2048, ffn=8192, layers=24, heads=32, g=64, p=0.10, sort="mean3std", batch=1, seq=2048),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=2, choices=(1, 2))
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--batch", type=int, default=0)
    args = ap.parse_args()
    import torch
    from transformers import OPTConfig, OPTForCausalLM
    from smoothquant_mixedprecision_b200 import calibration, ops
    from smoothquant_mixedprecision_b200.fake_quant import quantize_opt, W4A4Linear, clear_act_cache
    c = CONFIGS[args.config]
    batch = args.batch or c["batch"]
    dev = "cuda:0"
    torch.manual_seed(0)
    cfg = OPTConfig(vocab_size=50272, hidden_size=c["hidden"], ffn_dim=c["ffn"], num_hidden_layers=c["layers"],
                    num_attention_heads=c["heads"], max_position_embeddings=2048, word_embed_proj_dim=c["hidden"])
    base = OPTForCausalLM(cfg).to(torch.float16).to(dev).eval()
    gen = torch.Generator().manual_seed(1)
    calib = [torch.randint(0, 50272, (1, 512), generator=gen) for _ in range(2)]       # two 512-token blocks
    feat = calibration.get_calib_feat(base, calib)
    keys = calibration.collect_sort_keys(base, calib, c["sort"])
    ids = torch.randint(0, 50272, (batch, c["seq"]), generator=gen).to(dev)

    def timed(model):
        @torch.no_grad()
        def step():
            return model(ids).logits
        for _ in range(2):
            out = step()
        ops.GEMM_EVENTS, ops.QUANT_EVENTS = [], []
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            out = step()
        e1.record()
        torch.cuda.synchronize()
        gev, ops.GEMM_EVENTS = ops.GEMM_EVENTS, None
        qev, ops.QUANT_EVENTS = ops.QUANT_EVENTS, None
        ms = e0.elapsed_time(e1) / args.steps
        return out.float(), ms, sum(a.elapsed_time(b) for a, b, *_ in gev) / args.steps, \
            sum(a.elapsed_time(b) for a, b, *_ in qev) / args.steps

    res = {"config": args.config, "workload": f"{c['name']} random-init fp16, W4A4 g={c['g']} {int(c['p'] * 100)}% salient, "
           f"{c['sort']} sort, quantize_bmm_input=True, {batch}x{c['seq']} tokens, calibration: 2 synthetic 512-token blocks"}
    ref_logits, ms, _, _ = timed(base)
    res["fp16_stock"] = {"tokens_per_s": batch * c["seq"] / (ms * 1e-3), "ms": ms}
    outs = {}
    for mode in ("reference", "fused"):
        m = copy.deepcopy(base)
        extra = dict(mode=mode, sort=c["sort"])
        if mode == "fused":
            extra["sort_keys"] = keys        # collector names == walker names ("model.decoder.layers.N...")
        quantize_opt(m, "per_group", "per_group", input_feat=feat, salient_prop=c["p"], group_size=c["g"], **extra)
        n = sum(isinstance(x, W4A4Linear) for x in m.modules())
        clear_act_cache()
        logits, ms, gemm_ms, quant_ms = timed(m)
        outs[mode] = logits
        res[mode] = {"tokens_per_s": batch * c["seq"] / (ms * 1e-3), "ms": ms, "w4a4_layers": n, "gemm_ms": gemm_ms,
                     "quant_rows_ms": quant_ms, "finite": bool(torch.isfinite(logits).all()),
                     "rel_l2_vs_fp16": float((logits - ref_logits).norm() / ref_logits.norm())}
        del m
        torch.cuda.empty_cache()
    d = outs["fused"] - outs["reference"]
    res["mode_gap_rel_l2"] = float(d.norm() / outs["reference"].norm())
    res["top1_agreement_modes"] = float((outs["fused"].argmax(-1) == outs["reference"].argmax(-1)).float().mean())
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
