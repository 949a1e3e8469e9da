"""Pin the CPU oracle (oracle/mq_oracle.py) against fixtures produced by the real reference
(tests/golden/make_golden.py).  Everything here is bit-exact: torch.equal, no tolerance."""
import torch

from oracle import mq_oracle as O


def test_group_quantisers_match_reference(golden_quantisers):
    for c in golden_quantisers:
        x, g, nb = c["x"], c["group_size"], c["n_bits"]
        x2 = x.reshape(-1, x.shape[-1])
        gq = O.group_quant(x2, nb, g)
        assert torch.equal(gq.deq.reshape(x.shape), c["act_group"])          # fake_quant.py:78-101
        if "weight_group" in c:
            assert torch.equal(gq.deq, c["weight_group"])                    # fake_quant.py:30-53
        # the decomposition is consistent: codes * scales == dequantised, |codes| <= 7
        G = gq.scales.shape[1]
        rec = (gq.codes.view(x2.shape[0], G, g).to(x.dtype) * gq.scales[:, :, None]).view(x2.shape[0], -1)
        assert torch.equal(rec[:, : x2.shape[1]], gq.deq)
        assert int(gq.codes.abs().max()) <= 7
        assert torch.equal(gq.codes[:, x2.shape[1]:], torch.zeros_like(gq.codes[:, x2.shape[1]:]))


def test_sorted_quantisers_match_reference_with_recorded_perm(golden_quantisers):
    for c in golden_quantisers:
        x, g, nb = c["x"], c["group_size"], c["n_bits"]
        out, gq, perm = O.act_quant_group_sort(x, nb, g, perm=c["act_group_sort_perm"])
        assert torch.equal(out, c["act_group_sort"])                         # fake_quant.py:105-154
        if "weight_group_sort" in c:
            outw, _, _ = O.weight_quant_group_sort(x, nb, g, perm=c["weight_group_sort_perm"])
            assert torch.equal(outw, c["weight_group_sort"])                 # fake_quant.py:157-207


def test_canonical_stable_perm_agrees_on_tie_free_keys(golden_quantisers):
    n = 0
    for c in golden_quantisers:
        x = c["x"].reshape(-1, c["x"].shape[-1])
        key = O.sort_key(x, "max")
        if key.unique().numel() != key.numel():
            continue
        n += 1
        assert torch.equal(O.stable_argsort(key), c["act_group_sort_perm"])
        out, _, _ = O.act_quant_group_sort(c["x"], c["n_bits"], c["group_size"])
        assert torch.equal(out, c["act_group_sort"])
    assert n >= 1


def test_degenerate_quantisers(golden_quantisers):
    for c in golden_quantisers:
        if "act_token" not in c:
            continue
        x = c["x"]
        assert torch.equal(O.per_row_quant(x, 4).deq, c["act_token"])         # fake_quant.py:57-64
        assert torch.equal(O.per_row_quant(x, 4).deq, c["weight_channel"])    # fake_quant.py:10-16
        assert torch.equal(O.per_tensor_quant(x, 4), c["act_tensor"])         # fake_quant.py:68-75
        assert torch.equal(O.per_tensor_quant(x, 4), c["weight_tensor"])      # fake_quant.py:20-26


def test_layer_weight_and_forward(golden_layers):
    for c in golden_layers:
        sal = c["salient_indices"]
        K = c["K"]
        if sal is not None:
            # canonical salient selection == the reference's whenever importance is tie-free
            if c["importance"].unique().numel() == K:
                assert torch.equal(O.select_salient(c["importance"], c["salient_prop"]), sal)
        init = c["init_argsorts"]
        wperm = None
        if c["weight_quant"] == "per_group":
            wperm = init[1] if sal is not None else init[0]
        wq = O.w4a4_weight_reference(c["weight_fp"], sal, 4, c["group_size"], c["weight_quant"], perm=wperm)
        assert torch.equal(wq, c["weight_q"])                                 # fake_quant.py:347-365
        fwd = c["fwd_argsorts"]
        aperm = fwd[0] if c["act_quant"] == "per_group" else None
        operm = fwd[2] if (c["quantize_output"] and c["act_quant"] == "per_group") else None
        bias = c["bias"]
        y = O.w4a4_forward_reference(c["x"], wq, bias, sal, 4, c["group_size"], act_quant=c["act_quant"],
                                     quantize_output=c["quantize_output"], act_perm=aperm, out_perm=operm)
        assert y.shape == c["y"].shape
        if c["x"].dtype == torch.float32:
            assert torch.equal(y, c["y"])                                     # fake_quant.py:279-322
        else:
            # half-precision F.linear on CPU is not guaranteed bitwise reproducible across thread counts
            assert torch.allclose(y.float(), c["y"].float(), rtol=2e-3, atol=2e-3)


def test_fused_oracle_int_exact_equals_dequantised_product(golden_layers):
    c = golden_layers[0]
    x = c["x"].reshape(-1, c["K"])
    part = O.fused_partition(c["K"], c["salient_indices"], O.sort_key(x, "max").float(), c["group_size"])
    assert part.perm.numel() + part.salient.numel() == c["K"]
    assert torch.equal(torch.sort(torch.cat([part.perm, part.salient])).values, torch.arange(c["K"]))
    y_int, aq, wq = O.fused_layer_oracle(x, c["weight_fp"], c["bias"], part, exact_int=True)
    y_deq, _, _ = O.fused_layer_oracle(x, c["weight_fp"], c["bias"], part, exact_int=False)
    rel = (y_int - y_deq).norm() / y_int.norm()
    assert rel < 2e-3          # int-exact vs fp16-rounded dequantised operands (~2.4e-4)


def test_smooth_scales_match_reference(golden_models):
    m = golden_models["llama"]
    s0, s1 = m["state0"], m["state_smoothed"]
    pre = "model.layers.0."
    ws = [s0[pre + f"self_attn.{p}_proj.weight"] for p in "qkv"]
    s = O.smooth_scales(m["act_scales"][pre + "self_attn.q_proj"], ws, m["alpha"])      # smooth.py:56-67
    assert torch.equal(s0[pre + "input_layernorm.weight"] / s, s1[pre + "input_layernorm.weight"])
    assert torch.equal(ws[0] * s.view(1, -1), s1[pre + "self_attn.q_proj.weight"])


def test_int4_pack_roundtrip():
    codes = torch.randint(-7, 8, (5, 64), dtype=torch.int8)
    assert torch.equal(O.int4_unpack(O.int4_pack(codes)), codes)
