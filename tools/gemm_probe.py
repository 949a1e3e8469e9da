"""GPU bring-up probe for the mixed GEMM (K2) and the row quantiser (K1): runs a ladder of cases, each in its own
subprocess (a faulting kernel must not poison the next case), and prints one line per case.

    python tools/gemm_probe.py            # whole ladder
    python tools/gemm_probe.py CASE_JSON  # one case (internal)
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = [
    # name, M, N, G, g, Ks, dtype, bn
    dict(name="f16_1tile_k64", M=128, N=128, G=0, g=128, Ks=64, dt="f16", bn=128),
    dict(name="f16_1tile_k256", M=128, N=128, G=0, g=128, Ks=256, dt="f16", bn=128),
    dict(name="f16_bn256", M=128, N=256, G=0, g=128, Ks=128, dt="f16", bn=256),
    dict(name="i8_1group", M=128, N=128, G=1, g=128, Ks=0, dt="f16", bn=128),
    dict(name="i8_2groups", M=128, N=128, G=2, g=128, Ks=0, dt="f16", bn=128),
    dict(name="i8_8groups_bn256", M=128, N=256, G=8, g=128, Ks=0, dt="f16", bn=256),
    dict(name="i8_g64", M=128, N=128, G=5, g=64, Ks=0, dt="f16", bn=128),
    dict(name="i8_g256", M=128, N=128, G=3, g=256, Ks=0, dt="f16", bn=128),
    dict(name="mixed_small", M=128, N=128, G=3, g=128, Ks=40, dt="f16", bn=128),
    dict(name="mixed_g64_odd", M=200, N=328, G=5, g=64, Ks=24, dt="f16", bn=128),
    dict(name="mixed_bf16", M=300, N=512, G=4, g=128, Ks=56, dt="bf16", bn=256),
    dict(name="mixed_multi_tile", M=1000, N=1024, G=31, g=128, Ks=204, dt="f16", bn=256),
    dict(name="mixed_7b_2k_bn128", M=2048, N=4096, G=31, g=128, Ks=204, dt="f16", bn=128),
    dict(name="mixed_7b_2k_bn256", M=2048, N=4096, G=31, g=128, Ks=204, dt="f16", bn=256),
    dict(name="mixed_g1024", M=512, N=512, G=4, g=1024, Ks=64, dt="f16", bn=256),
]


def run_case(c):
    import torch
    from smoothquant_mixedprecision_b200 import ops, _lib
    torch.manual_seed(0)
    dev = "cuda"
    dt = torch.float16 if c["dt"] == "f16" else torch.bfloat16
    M, N, G, g, Ks = c["M"], c["N"], c["G"], c["g"], c["Ks"]
    _lib.set_option("gemm_bn", c["bn"])
    a = w = None
    ref = torch.zeros(M, N, dtype=torch.float32, device=dev)
    if G:
        ldc = (G * g + 127) // 128 * 128
        ac = torch.zeros(M, ldc, dtype=torch.int8, device=dev)
        wc = torch.zeros(N, ldc, dtype=torch.int8, device=dev)
        ac[:, :G * g] = torch.randint(-7, 8, (M, G * g), device=dev, dtype=torch.int8)
        wc[:, :G * g] = torch.randint(-7, 8, (N, G * g), device=dev, dtype=torch.int8)
        lda_s, ldw_s = (M + 3) // 4 * 4, (N + 3) // 4 * 4
        asc = torch.rand(G, lda_s, device=dev) * 0.1 + 0.01
        wsc = torch.rand(G, ldw_s, device=dev) * 0.01 + 0.001
        for gi in range(G):
            p = ac[:, gi * g:(gi + 1) * g].float() @ wc[:, gi * g:(gi + 1) * g].float().t()
            ref += p * (asc[gi, :M, None] * wsc[gi, None, :N])
    asal = wsal = None
    if Ks:
        lds = (Ks + 7) // 8 * 8
        asal = torch.zeros(M, lds, dtype=dt, device=dev)
        wsal = torch.zeros(N, lds, dtype=dt, device=dev)
        asal[:, :Ks] = torch.randn(M, Ks, device=dev).to(dt)
        wsal[:, :Ks] = (torch.randn(N, Ks, device=dev) * 0.05).to(dt)
        ref += asal.float() @ wsal.float().t()
    bias = (torch.randn(N, device=dev) * 0.1).to(dt)
    ref += bias.float()
    if G:
        a = ops.QuantizedRows(codes=ac, scales=asc, sal=asal, rows=M, Kn=G * g, Ks=Ks, group_size=g)
        w = ops.QuantizedRows(codes=wc, scales=wsc, sal=wsal, rows=N, Kn=G * g, Ks=Ks, group_size=g)
        y = ops.mixed_gemm(a, w, bias)
    else:
        y = ops.mixed_gemm(None, None, bias, a_sal=asal, w_sal=wsal, Ks=Ks)
    torch.cuda.synchronize()
    fault = _lib.gemm_fault()
    err = (y.float() - ref).abs()
    denom = ref.abs().max().item() + 1e-9
    bad = (err > 2e-2 * denom).nonzero()
    out = dict(name=c["name"], max_abs=err.max().item(), ref_max=denom, rel=err.max().item() / denom,
               fault=fault[:5], nbad=int(bad.shape[0]))
    if bad.shape[0]:
        out["first_bad"] = bad[:6].tolist()
        out["rows_bad"] = sorted(set(bad[:, 0].tolist()))[:12]
        out["cols_bad"] = sorted(set(bad[:, 1].tolist()))[:12]
        i, j = bad[0].tolist()
        out["y_vs_ref"] = [y[i, j].item(), ref[i, j].item()]
    # timing for the big ones
    if M * N >= 2048 * 4096:
        for _ in range(3):
            ops.mixed_gemm(a, w, bias)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            ops.mixed_gemm(a, w, bias)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        out["ms"] = ms
        out["tops"] = 2.0 * M * N * (G * g + Ks) / ms / 1e9
    print("CASE " + json.dumps(out), flush=True)


def main():
    if len(sys.argv) > 1:
        run_case(json.loads(sys.argv[1]))
        return
    for c in CASES:
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), json.dumps(c)], capture_output=True,
                               text=True, timeout=120)
            lines = [l for l in r.stdout.splitlines() if l.startswith("CASE ")]
            if lines:
                print(lines[-1], flush=True)
            else:
                print(f"CASE-FAIL {c['name']} rc={r.returncode} :: {(r.stdout + r.stderr)[-600:]}", flush=True)
        except subprocess.TimeoutExpired:
            print(f"CASE-TIMEOUT {c['name']}", flush=True)


if __name__ == "__main__":
    main()
