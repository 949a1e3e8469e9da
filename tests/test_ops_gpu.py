"""GPU parity tests of the C-ABI kernels (through smoothquant_mixedprecision_b200.ops -> ctypes -> libmq.so) against
the CPU oracle and the reference-generated golden fixtures.  Integer / code / scale / permutation results are
compared bit for bit; the GEMM within a stated tolerance."""
import pytest
import torch

from oracle import mq_oracle as O

pytestmark = pytest.mark.gpu

DEV = "cuda"


@pytest.fixture(scope="module")
def ops():
    from smoothquant_mixedprecision_b200 import ops as _ops
    return _ops


@pytest.fixture(scope="module")
def mqlib():
    from smoothquant_mixedprecision_b200 import _lib
    return _lib


def synth(rows, cols, dtype, seed, outliers=True):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(rows, cols, generator=g)
    scale = torch.exp(0.5 * torch.randn(cols, generator=g))
    if outliers:
        idx = torch.randperm(cols, generator=g)[: max(1, cols // 100)]
        scale[idx] *= torch.empty(idx.numel()).uniform_(10, 40, generator=g)
    return (x * scale).to(dtype)


# ------------------------------------------------------------------------------------------------ K0 / K0b
@pytest.mark.parametrize("shape,dtype", [((64, 256), torch.float16), ((33, 777), torch.float16),
                                         ((1000, 4096), torch.bfloat16), ((17, 130), torch.float32),
                                         ((4096, 1024), torch.float16)])
def test_colstat(ops, shape, dtype):
    x = synth(*shape, dtype, seed=shape[0])
    xg = x.to(DEV)
    assert torch.equal(ops.colstat(xg, "max").cpu(), O.sort_key(x, "max").float())          # fake_quant.py:113
    mean = x.abs().double().mean(dim=0).float()
    assert torch.allclose(ops.colstat(xg, "mean").cpu(), mean, rtol=1e-6, atol=0)            # run_experiments.py:60
    assert torch.allclose(ops.colstat(xg, "mean3std").cpu(), O.sort_key(x, "mean3std"), rtol=2e-4, atol=1e-6)
    assert torch.equal(ops.colstat(xg, "argmax").cpu(), O.sort_key(x, "argmax"))


@pytest.mark.parametrize("n", [1, 5, 255, 256, 4096, 11008])
def test_argsort_stable(ops, n):
    g = torch.Generator().manual_seed(n)
    key = torch.randint(0, max(2, n // 3), (n,), generator=g).float()       # many ties
    kg = key.to(DEV)
    assert torch.equal(ops.argsort_stable(kg).cpu().long(), O.stable_argsort(key))
    assert torch.equal(ops.argsort_stable(kg, descending=True).cpu().long(), O.stable_argsort(key, descending=True))
    if n >= 5:
        subset = torch.sort(torch.randperm(n, generator=g)[: n - n // 4]).values
        want = subset[O.stable_argsort(key[subset])]
        got = ops.argsort_stable(kg, subset=subset.to(DEV).int()).cpu().long()
        assert torch.equal(got, want)


def test_argsort_orders_nan_and_inf_keys_like_torch(ops):
    """NaN keys compare false with everything; the rank must still be a total order (NaN above +inf,
    ties by index) so that the output is a permutation."""
    key = torch.tensor([1.0, float("nan"), -2.0, float("inf"), float("nan"), -float("inf"), 1.0, 0.0])
    for desc in (False, True):
        got = ops.argsort_stable(key.to(DEV), descending=desc).cpu().long()
        assert sorted(got.tolist()) == list(range(8))
        assert torch.equal(got, torch.argsort(key, stable=True, descending=desc))


# ------------------------------------------------------------------------------------------------ K1 / P0
def _check_quant_rows(ops, x, g, n_bits, perm, sal, smooth=None):
    xs = x if smooth is None else (x / smooth)
    want = O.group_quant(xs[:, perm], n_bits, g)
    q = ops.quant_rows(x.to(DEV), perm.to(DEV), sal.to(DEV) if sal is not None else None, g, n_bits,
                       smooth=None if smooth is None else smooth.to(DEV))
    G = want.scales.shape[1]
    codes = q.codes.cpu()
    assert torch.equal(codes[:, : G * g], want.codes)
    assert int(codes[:, G * g:].abs().max()) == 0 if codes.shape[1] > G * g else True
    assert torch.equal(q.scales.cpu()[:, : x.shape[0]].t(), want.scales.float())
    if sal is not None:
        s = q.sal.cpu()
        assert torch.equal(s[:, : sal.numel()], xs[:, sal])
        assert float(s[:, sal.numel():].abs().sum()) == 0.0


def test_quant_rows_matches_oracle_on_golden_inputs(ops, golden_quantisers, mqlib):
    for force in (0, 1):
        mqlib.set_option("quant_force_generic", force)
        try:
            for i, c in enumerate(golden_quantisers):
                x = c["x"].reshape(-1, c["x"].shape[-1])
                K = x.shape[1]
                gen = torch.Generator().manual_seed(100 + i)
                order = torch.randperm(K, generator=gen)
                ks = max(1, K // 20)
                sal = torch.sort(order[:ks]).values
                perm = order[ks:]
                _check_quant_rows(ops, x, c["group_size"], c["n_bits"], perm, sal)
        finally:
            mqlib.set_option("quant_force_generic", 0)


@pytest.mark.parametrize("R,K,g,dtype", [(64, 1024, 64, torch.float16), (37, 4096, 128, torch.float16),
                                         (16, 2048, 1024, torch.float16), (50, 512, 32, torch.bfloat16),
                                         (9, 768, 128, torch.float32), (130, 11008, 128, torch.float16),
                                         (3, 256, 16, torch.float16)])
def test_quant_rows_fast_path(ops, R, K, g, dtype):
    x = synth(R, K, dtype, seed=R + K)
    gen = torch.Generator().manual_seed(R)
    order = torch.randperm(K, generator=gen)
    ks = max(1, int(0.05 * K))
    _check_quant_rows(ops, x, g, 4, order[ks:], torch.sort(order[:ks]).values)
    _check_quant_rows(ops, x, g, 4, torch.arange(K), None)                     # no salient, identity order
    if dtype != torch.float32:
        sm = (torch.rand(K, generator=gen) + 0.5).to(dtype)
        _check_quant_rows(ops, x, g, 4, order[ks:], torch.sort(order[:ks]).values, smooth=sm)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("R,K,g,ks", [(130, 4096, 128, 204), (37, 11008, 128, 550), (64, 2048, 64, 204), (9, 1024, 32, 0),
                                      (33, 2048, 1024, 100), (16, 8192, 256, 819), (5, 512, 128, 0)])
def test_quant_rows_h2_contiguous_layout_and_kernel_agreement(ops, mqlib, R, K, g, ks, dtype):
    """The packed-pair kernel (quant_rows_h2) in both of its modes against the oracle and against the r01 kernels:
    gather mode (explicit permutation) and contiguous mode (perm=None: quantised columns first, salient tail behind,
    the layout a folded producer emits), including magnitudes that drive scales into the fp16 denormal range (exact
    fp32 fallback inside the kernel) and groups straddling Kn (zero padding, fake_quant.py:84-86)."""
    x = synth(R, K, dtype, seed=R * 3 + K)
    x[: R // 3] *= 1e-4                                     # tiny rows: denormal scales -> slow path
    x[R // 3, : K // 2] = 0.0
    Kn = K - ks
    sal = torch.arange(Kn, K) if ks else None
    want = O.group_quant(x[:, :Kn], 4, g)
    G = want.scales.shape[1]
    outs = {}
    for h2 in (1, 2):
        mqlib.set_option("quant_h2", h2)
        try:
            q = ops.quant_rows(x.to(DEV), None, None if sal is None else sal.to(DEV), g, 4, Kn=Kn)
            qg = ops.quant_rows(x.to(DEV), torch.arange(Kn).to(DEV), None if sal is None else sal.to(DEV), g, 4)
        finally:
            mqlib.set_option("quant_h2", 0)
        for name, r in (("contig", q), ("gather", qg)):
            codes = r.codes.cpu()
            assert torch.equal(codes[:, : G * g], want.codes), (name, h2)
            assert int(codes[:, G * g:].abs().max()) == 0 if codes.shape[1] > G * g else True
            assert torch.equal(r.scales.cpu()[:, :R].t(), want.scales.float()), (name, h2)
            if ks:
                assert torch.equal(r.sal.cpu()[:, :ks], x[:, Kn:]), (name, h2)
                assert float(r.sal.cpu()[:, ks:].abs().sum()) == 0.0
        outs[h2] = q
    assert torch.equal(outs[1].codes, outs[2].codes) and torch.equal(outs[1].scales[:, :R], outs[2].scales[:, :R])


def test_quant_rows_zero_rows_and_nbits8(ops):
    x = torch.zeros(8, 256, dtype=torch.float16)
    _check_quant_rows(ops, x, 128, 4, torch.arange(256), None)                 # all-zero groups: scale = fp16(1e-5)/7
    x = synth(8, 256, torch.float16, 3)
    _check_quant_rows(ops, x, 64, 8, torch.arange(256), None)


def test_fakequant_rows_matches_reference_outputs(ops, golden_quantisers):
    for c in golden_quantisers:
        x = c["x"]
        xg = x.to(DEV)
        perm = c["act_group_sort_perm"]
        out = ops.fakequant_rows(xg, perm.to(DEV), c["group_size"], c["n_bits"])
        assert torch.equal(out.cpu(), c["act_group_sort"])                       # fake_quant.py:105-154
        assert torch.equal(xg.cpu(), x)                                          # input untouched
        out2 = ops.fakequant_rows(xg, None, c["group_size"], c["n_bits"])
        assert torch.equal(out2.cpu(), c["act_group"])                           # fake_quant.py:78-101
        if "weight_group_sort" in c:
            outw = ops.fakequant_rows(xg, c["weight_group_sort_perm"].to(DEV), c["group_size"], c["n_bits"])
            assert torch.equal(outw.cpu(), c["weight_group_sort"])               # fake_quant.py:157-207
        if "act_token" in c:
            x2 = x.reshape(-1, x.shape[-1])
            tok = ops.fakequant_rows(xg, None, x2.shape[1], 4)
            assert torch.equal(tok.cpu(), c["act_token"])                        # fake_quant.py:57-64
            ten = ops.fakequant_tensor(xg, 4)
            assert torch.equal(ten.cpu(), c["act_tensor"])                       # fake_quant.py:68-75
            z = xg.clone()
            r = ops.fakequant_rows(z, None, x2.shape[1], 4, inplace=True)
            assert r.data_ptr() == z.data_ptr() and torch.equal(z.cpu(), c["act_token"])


def test_dynamic_sort_pipeline_equals_reference_on_tie_free_keys(ops, golden_quantisers):
    n = 0
    for c in golden_quantisers:
        x = c["x"].reshape(-1, c["x"].shape[-1])
        key = O.sort_key(x, "max")
        if key.unique().numel() != key.numel():
            continue
        n += 1
        xg = c["x"].to(DEV)
        perm = ops.argsort_stable(ops.colstat(xg, "max"))
        assert torch.equal(perm.cpu().long(), c["act_group_sort_perm"])
        out = ops.fakequant_rows(xg, perm, c["group_size"], c["n_bits"])
        assert torch.equal(out.cpu(), c["act_group_sort"])
    assert n >= 1


def test_fakequant_passthrough_of_unlisted_columns(ops):
    x = synth(20, 512, torch.float16, 5)
    sal = torch.tensor([3, 77, 300, 511])
    keep = O.nonsalient_columns(512, sal)
    want = x.clone()
    want[:, keep] = O.group_quant(x[:, keep], 4, 128).deq
    got = ops.fakequant_rows(x.to(DEV), keep.to(DEV), 128, 4)
    assert torch.equal(got.cpu(), want)


def test_fast_and_generic_kernels_agree_at_full_size(ops, mqlib):
    x = synth(2048, 4096, torch.float16, 11).to(DEV)
    order = torch.randperm(4096, generator=torch.Generator().manual_seed(1))
    perm, sal = order[204:].to(DEV), torch.sort(order[:204]).values.to(DEV)
    a = ops.quant_rows(x, perm, sal, 128)
    d = ops.fakequant_rows(x, perm, 128)
    mqlib.set_option("quant_force_generic", 1)
    try:
        b = ops.quant_rows(x, perm, sal, 128)
        e = ops.fakequant_rows(x, perm, 128)
    finally:
        mqlib.set_option("quant_force_generic", 0)
    assert torch.equal(a.codes, b.codes) and torch.equal(a.scales, b.scales) and torch.equal(a.sal, b.sal)
    assert torch.equal(d, e)
    # properties: |codes| <= 7, each group attains 7 somewhere unless it is all zero, dequantisation is idempotent
    assert int(a.codes.abs().max()) == 7
    assert torch.equal(ops.fakequant_rows(d, perm, 128), d)
    assert mqlib.quant_fault()[0] == 0


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_division_free_codes_on_rounding_boundaries(ops, mqlib, dtype):
    """The fast kernel computes round_half_even(rnd_T(x / scale)) without the IEEE divide (remainder-corrected
    reciprocal multiply).  Adversarial rows: every group holds its abs-max A (so scale = rnd_T(rnd_T(A) / 7)) and
    values within a few T-ulps of scale * (k + 0.5), i.e. on and around the rounding boundaries where the quotient's
    rounding to T decides the code.  Must agree bit for bit with the generic kernel (true divide) and the oracle."""
    g, G, R = 128, 8, 512
    gen = torch.Generator().manual_seed(5)
    A = (torch.rand(R, G, generator=gen) * 30 + 0.01).to(dtype)
    lo = torch.tensor(1e-5, dtype=dtype)
    scale = (torch.maximum(A, lo) / 7).to(dtype)                      # dtype-local ops like the reference
    k = torch.randint(-7, 7, (R, G, g), generator=gen).to(torch.float32)
    x = (scale.float()[:, :, None] * (k + 0.5)).to(dtype)
    steps = torch.randint(-3, 4, (R, G, g), generator=gen)
    for _ in range(3):                                                # walk up to 3 ulps either way
        # step through T's own grid (an fp32 nextafter would round straight back to the same T value)
        ulp = (x.abs().float() * 2.0 ** (-10 if dtype == torch.float16 else -7)).clamp_min(1e-8)
        up, dn = (x.float() + ulp).to(dtype), (x.float() - ulp).to(dtype)
        x = torch.where(steps > 0, up, torch.where(steps < 0, dn, x))
        steps = steps - steps.sign()
    x = torch.minimum(torch.maximum(x, -A[:, :, None]), A[:, :, None])
    x[:, :, 0] = A
    x2 = x.reshape(R, G * g).to(DEV)
    a = ops.quant_rows(x2, None, None, g)
    d = ops.fakequant_rows(x2, None, g)
    mqlib.set_option("quant_force_generic", 1)
    try:
        b = ops.quant_rows(x2, None, None, g)
        e = ops.fakequant_rows(x2, None, g)
    finally:
        mqlib.set_option("quant_force_generic", 0)
    assert torch.equal(a.codes, b.codes) and torch.equal(a.scales, b.scales)
    assert torch.equal(d, e)
    q = O.group_quant(x2.cpu(), 4, g)
    assert torch.equal(a.codes.cpu()[:, : G * g], q.codes) and torch.equal(d.cpu(), q.deq)
    # the case is adversarial only if boundary quotients actually occur
    quot = (x2.cpu().float().view(R, G, g) / q.scales.float().view(R, G, 1)).to(dtype).float()
    assert int(((quot - quot.floor()) == 0.5).sum()) > 100


def test_int4_pack_roundtrip(ops):
    codes = torch.randint(-7, 8, (33, 256), dtype=torch.int8)
    packed = ops.pack_int4(codes.to(DEV))
    assert torch.equal(packed.cpu(), O.int4_pack(codes))
    assert torch.equal(ops.unpack_int4(packed).cpu(), codes)


# ------------------------------------------------------------------------------------------------ K2
def _gemm_reference(a, w, bias, Ks, M, N):
    """fp32 composition on the GPU from integer codes and scales (exact partial sums: |sum| < 2^24)."""
    ref = torch.zeros(M, N, dtype=torch.float32, device=DEV)
    g = a.group_size
    for gi in range(a.groups):
        p = a.codes[:, gi * g:(gi + 1) * g].float() @ w.codes[:, gi * g:(gi + 1) * g].float().t()
        ref += p * (a.scales[gi, :M, None] * w.scales[gi, None, :N])
    if Ks:
        ref += a.sal[:, :Ks].float() @ w.sal[:, :Ks].float().t()
    if bias is not None:
        ref += bias.float()
    return ref


@pytest.mark.parametrize("M,N,K,g,p,dtype,bn", [
    (128, 128, 512, 128, 0.05, torch.float16, 128),
    (24, 192, 256, 64, 0.05, torch.float16, 128),
    (300, 328, 1024, 64, 0.10, torch.float16, 128),
    (512, 768, 2048, 128, 0.05, torch.bfloat16, 256),
    (1000, 1024, 4096, 128, 0.05, torch.float16, 256),
    (256, 512, 4096, 1024, 0.0, torch.float16, 256),
    (2048, 4096, 4096, 128, 0.05, torch.float16, 256),
])
def test_mixed_gemm_matches_composed_reference(ops, mqlib, M, N, K, g, p, dtype, bn):
    x = synth(M, K, dtype, seed=M + N).to(DEV)
    w = (torch.randn(N, K, generator=torch.Generator().manual_seed(N)) * 0.02).to(dtype).to(DEV)
    bias = (torch.randn(N, generator=torch.Generator().manual_seed(7)) * 0.1).to(dtype).to(DEV)
    order = torch.randperm(K, generator=torch.Generator().manual_seed(K))
    ks = max(1, int(p * K)) if p > 0 else 0
    perm = order[ks:].to(DEV)
    sal = torch.sort(order[:ks]).values.to(DEV) if ks else None
    a = ops.quant_rows(x, perm, sal, g)
    wq = ops.quant_rows(w, perm, sal, g)
    mqlib.set_option("gemm_bn", bn)
    try:
        y = ops.mixed_gemm(a, wq, bias)
        torch.cuda.synchronize()
    finally:
        mqlib.set_option("gemm_bn", 0)
    assert mqlib.gemm_fault()[0] == 0
    ref = _gemm_reference(a, wq, bias, ks, M, N)
    # tolerance: output rounding to fp16 (2^-11) / bf16 (2^-8) relative + fp32 accumulation-order noise
    rtol = 1e-3 if dtype == torch.float16 else 8e-3
    err = (y.float() - ref).abs()
    assert float((err - rtol * ref.abs()).max()) <= 2e-3 * float(ref.abs().max())
    assert float(err.norm() / ref.norm()) < (5e-4 if dtype == torch.float16 else 4e-3)


def test_mixed_gemm_against_cpu_oracle_on_golden_layer(ops, golden_layers):
    """Full fused path (K1 on both operands + K2) against oracle.fused_layer_oracle evaluated in float64."""
    c = golden_layers[0]
    K, g = c["K"], c["group_size"]
    x = c["x"].reshape(-1, K)
    part = O.fused_partition(K, c["salient_indices"], O.sort_key(x, "max").float(), g)
    y64, aq, wq = O.fused_layer_oracle(x, c["weight_fp"], c["bias"], part)
    perm, sal = part.perm.to(DEV), part.salient.to(DEV)
    a = ops.quant_rows(x.to(DEV), perm, sal, g)
    w = ops.quant_rows(c["weight_fp"].to(DEV), perm, sal, g)
    G = aq.scales.shape[1]
    assert torch.equal(a.codes.cpu()[:, : G * g], aq.codes) and torch.equal(w.codes.cpu()[:, : G * g], wq.codes)
    assert torch.equal(a.scales.cpu()[:, : x.shape[0]].t(), aq.scales.float())
    assert torch.equal(w.scales.cpu()[:, : c["N"]].t(), wq.scales.float())
    bias = c["bias"].reshape(-1).to(DEV) if c["bias"] is not None else None
    y = ops.mixed_gemm(a, w, bias).cpu().double()
    err = (y - y64).abs()
    assert float(err.max()) <= 1e-3 * float(y64.abs().max()) + 1e-3       # fp16 output rounding of |y| <~ 10


def test_pure_f16_gemm_reference_mode(ops):
    M, N, K = 200, 328, 520                                                   # K not a multiple of 64: TMA zero fill
    a = torch.randn(M, K, generator=torch.Generator().manual_seed(1)).half().to(DEV)
    w = (torch.randn(N, K, generator=torch.Generator().manual_seed(2)) * 0.05).half().to(DEV)
    b = torch.randn(N, generator=torch.Generator().manual_seed(3)).half().to(DEV)
    y = ops.mixed_gemm(None, None, b, a_sal=a, w_sal=w)
    ref = a.float() @ w.float().t() + b.float()
    assert float((y.float() - ref).abs().max()) <= 2e-3 * float(ref.abs().max())


@pytest.mark.parametrize("g,Gi", [(128, 0), (128, 3), (128, 7), (64, 4), (64, 3), (32, 8), (32, 5)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_hybrid_split_tail_is_the_reference_dequantised_value(ops, mqlib, g, Gi, dtype):
    """mq_quant_rows_split: the first Gi groups are the integer codes / scales of the plain quantiser, the remaining
    groups are bit for bit the fake-quantised (dequantised) values the reference quantiser returns, laid out behind
    the salient columns; fast and generic kernels agree; the mixed GEMM on the split operands matches the composed
    float64 reference."""
    M, K, N, ks = 160, 1024, 192, 51
    x = synth(M, K, dtype, 21).to(DEV)
    w = (torch.randn(N, K, generator=torch.Generator().manual_seed(4)) * 0.05).to(dtype).to(DEV)
    order = torch.randperm(K, generator=torch.Generator().manual_seed(2))
    perm, sal = order[ks:].to(DEV), torch.sort(order[:ks]).values.to(DEV)
    Kn = K - ks
    G = (Kn + g - 1) // g
    full = ops.quant_rows(x, perm, sal, g)
    deq = ops.fakequant_rows(x, perm, g)                      # == quantize_activation_per_group_absmax on x[:, perm]
    a = ops.quant_rows(x, perm, sal, g, int_groups=Gi)
    mqlib.set_option("quant_force_generic", 1)
    try:
        a_gen = ops.quant_rows(x, perm, sal, g, int_groups=Gi)
    finally:
        mqlib.set_option("quant_force_generic", 0)
    t0, tail = (ks + 7) // 8 * 8, (G - Gi) * g
    assert a.Ks == t0 + tail and a.groups == Gi and a.n_salient == ks
    if Gi:
        assert torch.equal(a.codes[:, : Gi * g], full.codes[:, : Gi * g]) and torch.equal(a.scales, full.scales[:Gi])
        assert torch.equal(a.codes[:, : Gi * g], a_gen.codes[:, : Gi * g]) and torch.equal(a.scales, a_gen.scales)
    assert torch.equal(a.sal[:, :ks], full.sal[:, :ks]) and not a.sal[:, ks:t0].any()
    want_tail = torch.zeros(M, tail, dtype=dtype, device=DEV)
    cols = perm[Gi * g:].long()
    want_tail[:, : cols.numel()] = deq[:, cols]
    assert torch.equal(a.sal[:, t0: t0 + tail], want_tail)
    assert torch.equal(a.sal[:, : t0 + tail], a_gen.sal[:, : t0 + tail])
    wq = ops.quant_rows(w, perm, sal, g, int_groups=Gi)
    y = ops.mixed_gemm(a, wq, None)
    wfull = ops.quant_rows(w, perm, sal, g)
    Gg = G * g
    ad = (full.codes[:, :Gg].double().view(M, G, g) * full.scales[:G, :M].t().double()[:, :, None]).view(M, Gg)
    wd = (wfull.codes[:, :Gg].double().view(N, G, g) * wfull.scales[:G, :N].t().double()[:, :, None]).view(N, Gg)
    ref = ad @ wd.t() + full.sal[:, :ks].double() @ wfull.sal[:, :ks].double().t()
    err = float((y.double() - ref).abs().max())
    # fp16 / bf16 rounding of the output and of the tail operands (2^-11 / 2^-8 relative), written as a bound
    tol = (2e-3 if dtype == torch.float16 else 1.6e-2) * float(ref.abs().max()) + 1e-3
    assert err <= tol, (err, tol)
    assert mqlib.gemm_fault()[0] == 0 and mqlib.quant_fault()[0] == 0


@pytest.mark.parametrize("M,N,K,g,p,frac", [(256, 256, 512, 128, 0.0, 1.0), (1000, 1000, 2048, 128, 0.05, 1.0),
                                            (300, 512, 1024, 64, 0.05, 1.0), (777, 768, 1024, 32, 0.1, 1.0),
                                            (2048, 512, 4096, 1024, 0.05, 1.0), (1500, 1280, 2048, 128, 0.05, 0.5),
                                            (640, 256, 1024, 128, 0.05, 0.0)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_cta_pair_kernel_is_bit_identical_to_single_cta_kernel(ops, mqlib, M, N, K, g, p, frac, dtype):
    """The cta_group::2 kernel (256 x 256 tiles over two SMs) issues the same MMAs and accumulates every output
    element in the same order as the single-CTA kernel: outputs must be equal bit for bit, ragged edges included."""
    x = synth(M, K, dtype, 5).to(DEV)
    w = (torch.randn(N, K, generator=torch.Generator().manual_seed(6)) * 0.05).to(dtype).to(DEV)
    order = torch.randperm(K, generator=torch.Generator().manual_seed(7))
    ks = int(p * K)
    perm = order[ks:].int().to(DEV)
    sal = torch.sort(order[:ks]).values.int().to(DEV) if ks else None
    G = (K - ks + g - 1) // g
    step = max(1, 128 // g)
    gi = int(round(G * frac)) // step * step
    a = ops.quant_rows(x, perm, sal, g, int_groups=gi)
    wq = ops.quant_rows(w, perm, sal, g, int_groups=gi)
    bias = torch.randn(N, generator=torch.Generator().manual_seed(8)).to(dtype).to(DEV)
    try:
        mqlib.set_option("gemm_pair", 2)
        y0 = ops.mixed_gemm(a, wq, bias)
        mqlib.set_option("gemm_pair", 1)
        y1 = ops.mixed_gemm(a, wq, bias)
    finally:
        mqlib.set_option("gemm_pair", 0)
    torch.cuda.synchronize()
    assert mqlib.gemm_fault()[0] == 0
    assert torch.equal(y0, y1)


# ------------------------------------------------------------------------------------------------ K2d (decode, int4)
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("M,N,K,g,p", [(1, 256, 512, 128, 0.05), (5, 1000, 1024, 64, 0.1), (8, 4096, 4096, 128, 0.05),
                                       (16, 520, 2048, 256, 0.0), (24, 11008, 4096, 128, 0.05), (33, 4096, 11008, 128, 0.05),
                                       (64, 1024, 8192, 1024, 0.1), (64, 8192, 8192, 128, 0.1), (3, 72, 320, 64, 0.3)])
def test_decode_gemm_matches_prefill_kernel_and_composed_reference(ops, mqlib, M, N, K, g, p, dtype):
    """mq_decode_gemm (INT4-resident weights, mma.sync swap-AB) against the prefill kernel on the same operands and
    against the fp64 composition of the codes / scales.  The integer part is the same arithmetic operation for
    operation; the salient slice differs only in fp32 summation order -> 2 ulp of the output type."""
    x = synth(M, K, dtype, seed=M + N)
    gen = torch.Generator().manual_seed(N + K)
    w = (torch.randn(N, K, generator=gen) * 0.05).to(dtype)
    bias = (torch.randn(N, generator=gen) * 0.1).to(dtype)
    order = torch.randperm(K, generator=gen)
    ks = max(1, int(p * K)) if p > 0 else 0
    perm = order[ks:].to(DEV)
    sal = torch.sort(order[:ks]).values.to(DEV) if ks else None
    a = ops.quant_rows(x.to(DEV), perm, sal, g)
    wq = ops.quant_rows(w.to(DEV), perm, sal, g)
    dw = ops.pack_decode(wq, dtype)
    Kn, G = K - ks, (K - ks + g - 1) // g
    # 0.5 byte per (padded) weight + one 16-bit scale per (row, group)
    assert dw.nib.numel() == ((N + 15) // 16) * ((G * g + 255) // 256) * (2048 + 32 * max(1, 256 // g))
    y_dec = ops.decode_gemm(a, dw, bias.to(DEV))
    y_pre = ops.mixed_gemm(a, wq, bias.to(DEV))
    torch.cuda.synchronize()
    assert y_dec.shape == (M, N) and y_dec.dtype == dtype
    ref = _gemm_reference(a, wq, bias.to(DEV), ks, M, N).double()
    scale = float(ref.abs().max())
    tol = (8e-3 if dtype == torch.bfloat16 else 1e-3) * scale + 1e-3
    assert float((y_dec.double() - ref).abs().max()) <= tol
    ulp = 2.0 ** (-7 if dtype == torch.bfloat16 else -10)
    assert float((y_dec.float() - y_pre.float()).abs().max()) <= 2 * ulp * scale
    if ks == 0:                                   # integer part without split-K: bit-identical to the prefill kernel
        mqlib.set_option("decode_split", 1)
        try:
            assert torch.equal(ops.decode_gemm(a, dw, bias.to(DEV)), y_pre)
        finally:
            mqlib.set_option("decode_split", 0)


@pytest.mark.parametrize("split,tpc", [(2, 0), (4, 0), (8, 0), (1, 3), (4, 5)])
@pytest.mark.parametrize("M,N,K,g,p", [(16, 1024, 8192, 128, 0.1), (7, 520, 4096, 1024, 0.05), (40, 2048, 2048, 64, 0.0)])
def test_decode_gemm_split_k_is_deterministic_and_exact(ops, mqlib, M, N, K, g, p, split, tpc):
    """Cluster split-K (decode_split CTAs share the row tiles and cut K on group boundaries, partial tiles summed in rank
    order through the leader's shared memory) against the fp64 composition; two runs are bit-identical."""
    x = synth(M, K, torch.float16, seed=M)
    gen = torch.Generator().manual_seed(K)
    w = (torch.randn(N, K, generator=gen) * 0.05).half()
    order = torch.randperm(K, generator=gen)
    ks = int(p * K)
    perm = order[ks:].to(DEV)
    sal = torch.sort(order[:ks]).values.to(DEV) if ks else None
    a = ops.quant_rows(x.to(DEV), perm, sal, g)
    wq = ops.quant_rows(w.to(DEV), perm, sal, g)
    dw = ops.pack_decode(wq, torch.float16)
    mqlib.set_option("decode_split", split)
    mqlib.set_option("decode_tpc", tpc)
    try:
        y1 = ops.decode_gemm(a, dw, None).clone()
        y2 = ops.decode_gemm(a, dw, None)
        torch.cuda.synchronize()
    finally:
        mqlib.set_option("decode_split", 0)
        mqlib.set_option("decode_tpc", 0)
    assert torch.equal(y1, y2)
    ref = _gemm_reference(a, wq, None, ks, M, N).double()
    assert float((y1.double() - ref).abs().max()) <= 1e-3 * float(ref.abs().max()) + 1e-3


@pytest.mark.parametrize("g", [64, 128, 512])
def test_decode_layer_releases_prefill_codes_and_rebuilds_them(ops, g):
    """W4A4Linear.prepare_decode(keep_prefill=False): the layer keeps 4 bits per non-salient weight on the device, runs
    M <= 64 through the decode kernel, and rebuilds the prefill operand bit for bit when a longer batch arrives."""
    from smoothquant_mixedprecision_b200 import fake_quant as fq
    torch.manual_seed(3)
    K, N = 2048, 1536
    lin = torch.nn.Linear(K, N).half().to(DEV)
    imp = torch.rand(K)
    mod = fq.W4A4Linear.from_float(lin, "per_group", "per_group", importance=imp, salient_prop=0.05, group_size=g,
                                   mode="fused", sort_key=imp, int_fraction=1.0)
    x16 = torch.randn(16, K, device=DEV).half()
    x300 = torch.randn(300, K, device=DEV).half()
    y16, y300, wd = mod(x16).clone(), mod(x300).clone(), mod.weight.clone()
    codes, wsal, full = mod.w_codes.clone(), mod.w_sal.clone(), mod.device_weight_bytes()
    wsc = mod.w_scales.clone()
    mod.prepare_decode(keep_prefill=False)
    assert mod.w_codes is None and mod.w_sal is None and mod.w_scales is None    # scales live in the nibble stream
    assert mod.device_weight_bytes() < 0.58 * full                  # 4-bit codes + 16-bit scales + the salient slice
    assert torch.equal(mod.weight, wd)                         # dequantised view from the nibble layout
    assert mod.w_codes is None                                 # ... without keeping the rebuilt operand
    yd = mod(x16)
    assert float((yd.float() - y16.float()).abs().max()) <= 2 * 2.0 ** -10 * float(y16.abs().max())
    assert torch.equal(mod(x300), y300)                        # > 64 tokens: prefill operand rebuilt, bit-identical
    assert torch.equal(mod.w_codes, codes) and torch.equal(mod.w_sal, wsal)
    assert torch.equal(mod.w_scales[:, :N], wsc[:, :N])
    st = mod.packed_state()
    back = fq.W4A4Linear.from_packed(st, DEV)
    assert torch.equal(back(x300), y300)
    with pytest.raises(ValueError):
        fq.W4A4Linear.from_float(lin, "per_group", "per_group", group_size=32, mode="fused").prepare_decode()
    fq.clear_act_cache()


def _torch_norm(x, w, b, eps, kind):
    """What the model runs in front of the linears: transformers' LlamaRMSNorm.forward / F.layer_norm."""
    if kind == "rms":
        h = x.float()
        h = h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + eps)
        return w * h.to(x.dtype)
    return torch.nn.functional.layer_norm(x, (x.shape[-1],), w, b, eps)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("kind", ["rms", "layer"])
@pytest.mark.parametrize("M,K,g,p", [(300, 4096, 128, 0.05), (77, 768, 64, 0.05), (1024, 5120, 128, 0.10),
                                     (33, 8192, 1024, 0.0), (4096, 2048, 32, 0.05)])
def test_norm_quant_rows_is_norm_then_row_quantiser(ops, M, K, g, p, kind, dtype):
    """mq_norm_quant_rows.  Floating-point part: the normalised rows equal the torch norm on all but a
    sliver of the elements and differ there by last-place units of the 16-bit type -- the row statistic is an fp32 sum
    in another order.  Integer part: codes, scales and the salient copy are those of the row quantiser / the oracle's group_quant
    on the kernel's own normalised rows, bit for bit, with a permutation and in contiguous mode."""
    from oracle import mq_oracle as O
    torch.manual_seed(M + K)
    x = (torch.randn(M, K, device=DEV) * torch.exp(0.5 * torch.randn(K, device=DEV)) + 0.3).to(dtype)
    w = (1 + 0.2 * torch.randn(K, device=DEV)).to(dtype)
    b = (0.1 * torch.randn(K, device=DEV)).to(dtype) if kind == "layer" else None
    eps = 1e-6 if kind == "rms" else 1e-5
    ks = int(round(K * p))
    order = torch.randperm(K, device=DEV)
    sal = torch.sort(order[:ks]).values.int() if ks else None
    ref = _torch_norm(x, w, b, eps, kind)
    for perm in (order[ks:].int(), None):
        if perm is None:
            Kn = K - ks
            sal_c = torch.arange(Kn, K, device=DEV).int() if ks else None
            a, y = ops.norm_quant_rows(x, w, b, eps, kind, None, sal_c, g, Kn=Kn, want_y=True)
            want = ops.quant_rows(y, None, sal_c, g, 4, Kn=Kn)
            cols, scol = torch.arange(Kn), sal_c
        else:
            a, y = ops.norm_quant_rows(x, w, b, eps, kind, perm, sal, g, want_y=True)
            want = ops.quant_rows(y, perm, sal, g, 4)
            cols, scol = perm.long().cpu(), sal
        # floating point: a last-place difference of the inner rounding rn(x * rstd) times the norm weight is at most
        # ~2 units in the last place of y; and the rows agree exactly on all but a sliver of the elements
        ulp = torch.finfo(dtype).eps * torch.exp2(torch.floor(torch.log2(
            ref.float().abs().clamp_min(torch.finfo(dtype).tiny))))
        d = (y.float() - ref.float()).abs()
        # (+ 2e-6 absolute: where w * n + b cancels to |y| << 1 the fp32 rounding noise of the O(1) terms is many ulps
        # of the tiny result -- bf16 keeps 8 bits at any magnitude)
        assert bool((d <= 3 * ulp + 2e-6).all()), float(((d - 2e-6) / ulp).max())
        assert float((d > 0).float().mean()) < 5e-3, float((d > 0).float().mean())
        # integer: the fused kernel quantises exactly like K1 on the same rows ...
        G = want.groups
        assert torch.equal(a.codes[:, : G * g], want.codes[:, : G * g])
        assert torch.equal(a.scales[:, :M], want.scales[:, :M])
        if ks:
            assert torch.equal(a.sal[:, :ks], want.sal[:, :ks]) and torch.equal(a.sal[:, :ks], y[:, scol.long()])
        # ... and like the reference quantiser (oracle) on sampled rows
        rows = torch.arange(0, M, max(1, M // 16))[:16]
        oq = O.group_quant(y[rows.to(DEV)].cpu()[:, cols], 4, g)
        assert torch.equal(a.codes[rows.to(DEV)].cpu()[:, : G * g], oq.codes[:, : G * g])
        assert torch.equal(a.scales[:G].cpu()[:, rows].t(), oq.scales.float())
    # without want_y the result is the same
    a2 = ops.norm_quant_rows(x, w, b, eps, kind, order[ks:].int(), sal, g)
    a1, _ = ops.norm_quant_rows(x, w, b, eps, kind, order[ks:].int(), sal, g, want_y=True)
    assert torch.equal(a2.codes, a1.codes) and torch.equal(a2.scales[:, :M], a1.scales[:, :M])
