"""GPU tests of the two components either side of the path: the calibration collectors that produce
its inputs (calibration.py:13-51, run_experiments.py:55-84) and the packed checkpoint that stores its weights."""
import functools
import io

import pytest
import torch
from torch import nn

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def fq():
    from smoothquant_mixedprecision_b200 import fake_quant
    return fake_quant


def _tiny_llama(seed, dtype=torch.float16):
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(seed)
    cfg = LlamaConfig(vocab_size=128, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=4, max_position_embeddings=128)
    return LlamaForCausalLM(cfg).to(dtype).to(DEV).eval()


def _batches(n=3, seq=64):
    g = torch.Generator().manual_seed(5)
    return [torch.randint(0, 128, (2, seq), generator=g) for _ in range(n)]


def _reference_hooks(model, batches):
    """The reference's two hook recipes, verbatim in behaviour (calibration.py:18-30, run_experiments.py:57-64), with
    stock torch reductions."""
    scales, feat = {}, {}

    def hook(m, x, y, name):
        x = x[0] if isinstance(x, tuple) else x
        t = x.view(-1, x.shape[-1]).abs().detach()
        mx = torch.max(t, dim=0)[0].float().cpu()
        scales[name] = torch.max(scales[name], mx) if name in scales else mx
        feat.setdefault(name, []).append(x.view(-1, x.shape[-1]).abs().mean(dim=0).cpu().detach())

    hooks = [m.register_forward_hook(functools.partial(hook, name=n)) for n, m in model.named_modules()
             if isinstance(m, nn.Linear)]
    with torch.no_grad():
        for ids in batches:
            model(ids.to(DEV))
    for h in hooks:
        h.remove()
    return scales, feat


def test_collectors_match_the_reference_hook_recipes():
    from smoothquant_mixedprecision_b200 import calibration as C
    model, batches = _tiny_llama(0), _batches()
    ref_scales, ref_feat = _reference_hooks(model, batches)
    scales = C.get_act_scales(model, batches)
    feat = C.get_calib_feat(model, batches)
    assert set(scales) == set(ref_scales) == set(feat) and "lm_head" in scales
    for n in ref_scales:
        assert scales[n].dtype == torch.float32 and torch.equal(scales[n], ref_scales[n]), n     # a max is exact
        assert len(feat[n]) == len(batches)
        for a, b in zip(feat[n], ref_feat[n]):
            assert a.dtype == b.dtype == torch.float16
            # same fp32-accumulated mean rounded once to fp16; summation order may move it by one fp16 ulp (2^-10 rel)
            assert torch.allclose(a.float(), b.float(), rtol=2 ** -10, atol=1e-7), n
    keys = C.collect_sort_keys(model, batches, "max")
    assert all(torch.equal(keys[n], ref_scales[n]) for n in ref_scales)
    with pytest.raises(ValueError):
        C.collect_sort_keys(model, batches, "median")
    assert not any(m._forward_hooks for m in model.modules())                                   # hooks removed


@pytest.mark.parametrize("mode,int_fraction,bias", [("fused", None, True), ("fused", 0.5, False), ("fused", 0.0, True),
                                                    ("reference", None, True)])
def test_packed_layer_round_trip_is_bit_exact(fq, mode, int_fraction, bias):
    torch.manual_seed(3)
    M, K, N, g = 80, 640, 384, 128
    x = (torch.randn(M, K, device=DEV) * torch.exp(0.5 * torch.randn(K, device=DEV))).half()
    imp = x.abs().float().mean(dim=0).cpu()
    lin = nn.Linear(K, N, bias=bias).half().to(DEV)
    kw = dict(mode="fused", sort_key=imp, int_fraction=int_fraction) if mode == "fused" else {}
    mod = fq.W4A4Linear.from_float(lin, "per_group", "per_group", importance=imp, salient_prop=0.05, group_size=g, **kw)
    buf = io.BytesIO()
    torch.save(mod.packed_state(), buf)
    buf.seek(0)
    state = torch.load(buf, weights_only=False)
    back = fq.W4A4Linear.from_packed(state, DEV)
    assert repr(back) == repr(mod) and back.mode == mod.mode and back.int_groups == mod.int_groups
    if mode == "fused":
        assert state["w_packed"].dtype == torch.uint8 and state["w_packed"].shape == (N, mod.w_codes.shape[1] // 2)
        assert state["w_scales"].dtype == torch.float16                   # scales are fp16 values: stored as such
        for name in ("w_codes", "w_scales", "w_sal", "perm"):
            assert torch.equal(getattr(back, name), getattr(mod, name)), name
        assert back._partition_id == mod._partition_id
    else:
        assert torch.equal(back.weight, mod.weight)
    fq.clear_act_cache()
    y0 = mod(x).clone()
    fq.clear_act_cache()
    assert torch.equal(back(x), y0)
    with pytest.raises(ValueError):
        fq.W4A4Linear.from_packed({**state, "meta": {**state["meta"], "format": 99}}, DEV)
    with pytest.raises(RuntimeError):
        fq.W4A4Linear.from_packed(state, "cpu")                           # no CPU fallback


def test_packed_model_round_trip(fq, tmp_path):
    from smoothquant_mixedprecision_b200 import calibration as C
    model, batches = _tiny_llama(1), _batches()
    feat = C.get_calib_feat(model, batches)
    keys = {n: v for n, v in C.collect_sort_keys(model, batches, "max").items()}
    fq.quantize_llama_like(model, "per_group", "per_group", input_feat=feat, salient_prop=0.05, group_size=128,
                           mode="fused", sort_keys=keys)
    ids = batches[0].to(DEV)
    with torch.no_grad():
        ref = model(ids).logits.clone()
    path = tmp_path / "packed.pt"
    ckpt = fq.save_packed_model(model, path)
    n_lin = sum(isinstance(m, fq.W4A4Linear) for m in model.modules())
    assert len(ckpt["linears"]) == n_lin == 14 and "lm_head.weight" in ckpt["rest"]
    assert not any(".w_codes" in k or "._weight" in k for k in ckpt["rest"])
    skeleton = _tiny_llama(2)                                               # different weights everywhere
    fq.load_packed_model(skeleton, path, DEV)
    assert sum(isinstance(m, fq.W4A4Linear) for m in skeleton.modules()) == n_lin
    fq.clear_act_cache()
    with torch.no_grad():
        out = skeleton(ids).logits
    assert torch.equal(out, ref)                                           # same packed operands, same kernels
    with pytest.raises(ValueError):
        fq.load_packed_model(_tiny_llama(2), {**ckpt, "format": 0}, DEV)
    with pytest.raises(KeyError):
        fq.load_packed_model(_tiny_llama(2), {**ckpt, "rest": {k: v for k, v in ckpt["rest"].items()
                                                                if k != "lm_head.weight"}}, DEV)
