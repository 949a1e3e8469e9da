"""CPU-side checks: the C-ABI library loads and exports every symbol include/mq.h declares, the Python surface keeps
the reference's signatures, errors are raised before the ABI is crossed, and there is no CPU fallback."""
import inspect
import json
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from smoothquant_mixedprecision_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "mq.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(mq_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(_lib.lib, name), name
    m = re.search(r"#define MQ_ABI_VERSION (\d+)", open(os.path.join(ROOT, "include", "mq.h")).read())
    assert _lib.lib.mq_version() == int(m.group(1)) == 2
    assert _lib.lib.mq_set_option(b"no_such_option", 1) == -1
    assert b"unknown option" in _lib.lib.mq_last_error()


def test_signatures_match_reference():
    from smoothquant_mixedprecision_b200 import fake_quant as fq, smooth as sm
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "signatures.json")))

    def check(fn, want, skip_self=False):
        got = [[p.name, None if p.default is inspect._empty else repr(p.default)]
               for p in inspect.signature(fn).parameters.values() if p.kind != p.VAR_KEYWORD]
        assert got[: len(want)] == want, (fn, got, want)
        assert all(d is not None for _, d in got[len(want):]), "extra parameters must be defaulted"

    for n, want in gold["fake_quant"].items():
        check(getattr(fq, n), want)
    check(fq.W4A4Linear.__init__, gold["W4A4Linear.__init__"])
    check(fq.W4A4Linear.from_float, gold["W4A4Linear.from_float"])
    for n, want in gold["smooth"].items():
        check(getattr(sm, n), want)


def test_drop_in_import_paths():
    import smoothquant
    from smoothquant import quantize_model, smooth_lm                    # smoothquant/__init__.py:1-4
    from smoothquant.fake_quant import W4A4Linear, quantize_llama_like, quantize_opt  # run_experiments.py:21
    from smoothquant.smooth import smooth_lm as s2                       # ppl_eval.py:4
    assert smooth_lm is s2 and callable(quantize_model) and smoothquant.__all__ == ["smooth_lm", "quantize_model"]
    assert W4A4Linear.__name__ == "W4A4Linear" and callable(quantize_llama_like) and callable(quantize_opt)
    from smoothquant.model_size import get_model_size                    # run_experiments.py:22
    import inspect
    assert [p.name for p in inspect.signature(get_model_size).parameters.values()] == \
        ["model", "data_width", "salient_prop", "group_size"]            # smoothquant/model_size.py:5


def test_model_size_matches_reference_formula_on_float_models():
    """smoothquant/model_size.py:5-15 on a model without swapped linears (bits, not bytes)."""
    from smoothquant.model_size import get_model_size
    m = torch.nn.Sequential(torch.nn.Linear(64, 32), torch.nn.Linear(32, 8, bias=False))
    n = 64 * 32 + 32 + 32 * 8
    assert get_model_size(m) == n * 16
    assert get_model_size(m, data_width=4, group_size=128) == n * (4 + 20 / 128)
    assert get_model_size(m, data_width=4, salient_prop=0.05, group_size=64) == \
        n * ((4 + 20 / 64) * 0.95 + (16 + 20 / 64) * 0.05)


def test_column_list_validation():
    """a permutation / salient list with out-of-range or duplicate ids must be rejected before it reaches
    the kernels (they turn it into unchecked 16-bit shared-memory offsets)."""
    from smoothquant_mixedprecision_b200.fake_quant import _validate_columns
    _validate_columns(torch.tensor([3, 0, 2]), 4, "ok")
    _validate_columns(None, 4, "none")
    for bad in ([0, 4], [-1, 2], [1, 1, 2]):
        with pytest.raises(ValueError):
            _validate_columns(torch.tensor(bad), 4, "perm")


def test_no_cpu_fallback_and_python_side_errors():
    from smoothquant_mixedprecision_b200 import fake_quant as fq, ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.colstat(torch.zeros(4, 8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        fq.W4A4Linear.from_float(torch.nn.Linear(8, 8))
    with pytest.raises(ValueError):
        fq.W4A4Linear(8, 8, act_quant="per_row")
    with pytest.raises(ValueError):
        fq.W4A4Linear.from_float(torch.nn.Linear(8, 8), weight_quant="nope")
    with pytest.raises(ValueError):
        fq.quantize_model(torch.nn.Linear(2, 2))
    m = fq.W4A4Linear(16, 8, importance=torch.arange(16.0), salient_prop=0.2)
    assert m.salient_indices.tolist() == [15, 14, 13]                    # max(1, int(.2*16)) most important channels
    assert "W4A4Linear(16, 8, bias=True" in repr(m)


def test_product_code_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "smoothquant_mixedprecision_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src, os.path.join(dirpath, f)


def test_packed_checkpoint_and_calibration_host_errors():
    """Host-side behaviour of the components either side of the path (no GPU work): format / device checks of the
    packed checkpoint, collectors refuse a CPU model, import paths of the shims."""
    from smoothquant_mixedprecision_b200 import calibration, fake_quant as fq
    from smoothquant.fake_quant import save_packed_model, load_packed_model          # exported through the shim
    from smoothquant.calibration import get_act_scales as ref_sig
    import inspect
    assert save_packed_model is fq.save_packed_model and load_packed_model is fq.load_packed_model
    assert list(inspect.signature(ref_sig).parameters) == ["model", "tokenizer", "dataset_path", "num_samples",
                                                           "seq_len"]                   # calibration.py:13
    meta = dict(format=fq.PACKED_FORMAT, in_features=8, out_features=4, group_size=4, quant_bits=4, mode="reference",
                sort="max", act_quant="per_group", weight_quant="per_group", output_quant="None", int_groups=None,
                int_fraction=None)
    state = {"meta": meta, "bias": None, "salient_indices": None, "weight": torch.zeros(4, 8, dtype=torch.float16)}
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        fq.W4A4Linear.from_packed(state, "cpu")
    with pytest.raises(ValueError, match="format"):
        fq.W4A4Linear.from_packed({**state, "meta": {**meta, "format": 0}}, "cuda")
    with pytest.raises(ValueError, match="format"):
        fq.load_packed_model(torch.nn.Linear(2, 2), {"format": 0, "linears": {}, "rest": {}})
    with torch.device("meta"):
        skeleton = torch.nn.Sequential(torch.nn.Linear(2, 2))
    with pytest.raises(ValueError, match="materialised"):
        fq.load_packed_model(skeleton, {"format": fq.PACKED_FORMAT, "linears": {}, "rest": {}})
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        calibration.get_act_scales(torch.nn.Sequential(torch.nn.Linear(4, 4)), [torch.zeros(1, 4)])
    with pytest.raises(ValueError):
        calibration.collect_sort_keys(torch.nn.Linear(2, 2), [], "median")


def test_fused_norm_falls_back_to_the_plain_norm_off_the_gpu_path():
    """FusedNorm only hands out a lazy NormedRows for 16-bit CUDA rows; everything else gets the wrapped norm's output.
    NormedRows carries the tensor attributes the attention / MLP blocks read and materialises the plain norm."""
    import torch
    from smoothquant_mixedprecision_b200 import fake_quant as fq
    ln = torch.nn.LayerNorm(32)
    x = torch.randn(2, 5, 32)
    assert torch.equal(fq.FusedNorm(ln, "layer")(x), ln(x))
    with pytest.raises(ValueError):
        fq.FusedNorm(ln, "batch")
    nr = fq.NormedRows(x, ln, "layer")
    assert nr.shape == x.shape and nr.dim() == 3 and nr.size(-1) == 32 and nr.dtype == x.dtype and nr.device == x.device
    assert nr.eps == ln.eps and torch.equal(nr.materialize(), ln(x)) and nr.materialize() is nr.materialize()
    # a parent whose consumers are not packed fused layers keeps its norm
    parent = torch.nn.Module()
    parent.norm = ln
    assert not fq.fuse_norm_(parent, "norm", [torch.nn.Linear(32, 8)], "layer") and parent.norm is ln


def test_integration_stub_matches_the_abi():
    """The ctypes stub INTEGRATION.md shows a maintainer binds the same number of arguments as include/mq.h declares
    (the stub went stale once when the workspace arguments were dropped)."""
    from smoothquant_mixedprecision_b200 import _lib
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    found = re.findall(r"_lib\.(mq_\w+)\.argtypes = \[(.*?)\]", text, flags=re.S)
    assert found, "no argtypes in the INTEGRATION.md stub"
    for name, body in found:
        n = len([a for a in body.replace("\n", " ").split(",") if a.strip()])
        assert name in _lib.SIGNATURES and n == len(_lib.SIGNATURES[name][1]), (name, n, len(_lib.SIGNATURES[name][1]))
    assert "mq_version() == 2" in text
