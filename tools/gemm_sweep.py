"""BASELINE.json config 4 / 5: isolated mixed-GEMM sweep over M x Llama-2 7B / 13B / 70B linear shapes x group x salient.

Each point quantises synthetic activations (LLM-like outlier channels) and weights with the fused row quantiser,
runs the mixed GEMM, checks it against a float64-accumulated composed reference (dequantised codes x codes + salient)
on a row sample, and times the GEMM with CUDA events.  One JSON line per point.

    python tools/gemm_sweep.py [--quick] [--family 7b,13b,70b] [--out FILE]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from smoothquant_mixedprecision_b200 import ops, _lib  # noqa: E402

SHAPES = {  # (N, K)
    "7b": [(4096, 4096), (11008, 4096), (4096, 11008)],
    "13b": [(5120, 5120), (13824, 5120), (5120, 13824)],
    "70b": [(8192, 8192), (1024, 8192), (28672, 8192), (8192, 28672)],
}


def point(M, N, K, g, p, iters, dev="cuda"):
    gen = torch.Generator(device=dev).manual_seed(M * 7919 + N + K + g)
    c = torch.exp(0.5 * torch.randn(K, device=dev, generator=gen))
    hot = torch.rand(K, device=dev, generator=gen) < 0.01
    c = torch.where(hot, c * (10 + 30 * torch.rand(K, device=dev, generator=gen)), c)
    x = (torch.randn(M, K, device=dev, generator=gen) * c).half()
    w = (torch.randn(N, K, device=dev, generator=gen) * 0.02).half()
    imp = x.abs().float().mean(dim=0)
    ks = max(1, int(p * K)) if p > 0 else 0
    order = torch.argsort(imp, descending=True, stable=True)
    sal = torch.sort(order[:ks]).values.int() if ks else None
    keep = torch.sort(order[ks:]).values.int()
    perm = ops.argsort_stable(x.abs().float().amax(dim=0), subset=keep)
    a = ops.quant_rows(x, perm, sal, g)
    wq = ops.quant_rows(w, perm, sal, g)
    y = ops.mixed_gemm(a, wq, None)
    # composed reference on a sample of rows (float64 accumulate)
    rows = torch.randperm(M, device=dev, generator=gen)[: min(M, 16)]
    G = a.scales.shape[0]
    ad = (a.codes[rows, : G * g].double().view(-1, G, g) * a.scales[:G, rows].t().double()[:, :, None]).view(-1, G * g)
    wd = (wq.codes[:, : G * g].double().view(N, G, g) * wq.scales[:G, :N].t().double()[:, :, None]).view(N, G * g)
    ref = ad @ wd.t()
    if ks:
        ref += a.sal[rows].double() @ wq.sal.double().t()
    err = float((y[rows].double() - ref).abs().max())
    tol = 1e-3 * float(ref.abs().max()) + 1e-3          # fp16 output rounding (2^-11 relative) + slack
    for _ in range(2):
        ops.mixed_gemm(a, wq, None)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.mixed_gemm(a, wq, None)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    wbytes = wq.codes.numel() + wq.scales.numel() * 4 + (wq.sal.numel() * 2 if ks else 0)
    return dict(M=M, N=N, K=K, g=g, p=p, ms=round(ms, 5), tops=round(2.0 * M * N * K / ms / 1e9, 1),
                weight_gbs=round(wbytes / ms / 1e6, 1), max_err=err, tol=tol, ok=bool(err <= tol))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--family", default="7b,13b")
    ap.add_argument("--out", default=None)
    ap.add_argument("--g", default=None, help="comma list overriding the group sizes")
    ap.add_argument("--p", default=None, help="comma list overriding the salient proportions")
    args = ap.parse_args()
    Ms = [16, 64, 256, 1024, 4096, 8192] if args.quick else [16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192]
    gs = [128] if args.quick else [64, 128, 1024]
    ps = [0.05] if args.quick else [0.0, 0.05, 0.10]
    if args.g:
        gs = [int(v) for v in args.g.split(",")]
    if args.p:
        ps = [float(v) for v in args.p.split(",")]
    out = open(args.out, "w") if args.out else None
    bad = 0
    for fam in args.family.split(","):
        for (N, K) in SHAPES[fam]:
            for g in gs:
                for p in ps:
                    for M in Ms:
                        r = point(M, N, K, g, p, iters=5 if M >= 2048 else 20)
                        r["family"] = fam
                        bad += 0 if r["ok"] else 1
                        line = json.dumps(r)
                        print(line, flush=True)
                        if out:
                            out.write(line + "\n")
    fault = _lib.gemm_fault()[0] or _lib.quant_fault()[0]
    print(json.dumps({"points_failed": bad, "fault": fault}), flush=True)
    if out:
        out.write(json.dumps({"points_failed": bad, "fault": fault}) + "\n")
    sys.exit(1 if bad or fault else 0)


if __name__ == "__main__":
    main()
