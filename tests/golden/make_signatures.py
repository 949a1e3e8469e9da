"""Record the reference's public signatures (names, parameter order, defaults) -> signatures.json.
Run in the build container only:  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_signatures.py"""
import inspect
import json
import os
import sys

sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference")
import smoothquant.fake_quant as rfq  # noqa: E402
import smoothquant.smooth as rsm  # noqa: E402

NAMES_FQ = ["quantize_weight_per_channel_absmax", "quantize_weight_per_tensor_absmax",
            "quantize_weight_per_group_absmax", "quantize_activation_per_token_absmax",
            "quantize_activation_per_tensor_absmax", "quantize_activation_per_group_absmax",
            "quantize_activation_per_group_absmax_sort", "quantize_weight_per_group_absmax_sort",
            "quantize_opt", "quantize_llama_like", "quantize_model"]


def sig(fn):
    return [[p.name, None if p.default is inspect._empty else repr(p.default)]
            for p in inspect.signature(fn).parameters.values()]


out = {"fake_quant": {n: sig(getattr(rfq, n)) for n in NAMES_FQ},
       "W4A4Linear.__init__": sig(rfq.W4A4Linear.__init__), "W4A4Linear.from_float": sig(rfq.W4A4Linear.from_float),
       "smooth": {n: sig(getattr(rsm, n)) for n in ["smooth_ln_fcs", "smooth_ln_fcs_llama_like", "smooth_lm"]}}
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "signatures.json"), "w"), indent=1)
