"""Bandwidth probe of the row quantiser (K1) and column statistics (K0) at Llama-2-7B prefill shapes."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from smoothquant_mixedprecision_b200 import ops, _lib  # noqa: E402


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    dev = "cuda"
    for (M, K, g, p) in [(32768, 4096, 128, 0.05), (8192, 11008, 128, 0.05), (32768, 4096, 64, 0.10),
                         (16384, 4096, 1024, 0.05)]:
        x = torch.randn(M, K, device=dev, dtype=torch.float16)
        order = torch.randperm(K, device=dev)
        ks = max(1, int(p * K))
        perm, sal = order[ks:].int(), torch.sort(order[:ks]).values.int()
        Kn = K - ks
        G = (Kn + g - 1) // g
        for opt in [dict(), dict(quant_rows_per_stage=2), dict(quant_rows_per_stage=8), dict(quant_warps=16)]:
            for k, v in opt.items():
                _lib.set_option(k, v)
            ms = timeit(lambda: ops.quant_rows(x, perm, sal, g))
            for k in opt:
                _lib.set_option(k, 0)
            ldc = (G * g + 127) // 128 * 128
            nbytes = M * K * 2 + M * ldc + M * G * 4 + M * ((ks + 7) // 8 * 8) * 2 + K * 4
            print("K1 " + json.dumps(dict(M=M, K=K, g=g, p=p, opt=opt, ms=ms, gbs=nbytes / ms / 1e6)), flush=True)
        ms = timeit(lambda: ops.fakequant_rows(x, perm, g))
        print("K1deq " + json.dumps(dict(M=M, K=K, g=g, ms=ms, gbs=2 * M * K * 2 / ms / 1e6)), flush=True)
        ms = timeit(lambda: ops.colstat(x, "max"))
        print("K0 " + json.dumps(dict(M=M, K=K, ms=ms, gbs=M * K * 2 / ms / 1e6)), flush=True)
        key = ops.colstat(x, "max")
        ms = timeit(lambda: ops.argsort_stable(key))
        print("K0b " + json.dumps(dict(n=K, ms=ms)), flush=True)
        y = torch.empty_like(x)
        ms = timeit(lambda: y.copy_(x))
        print("copy " + json.dumps(dict(M=M, K=K, ms=ms, gbs=2 * M * K * 2 / ms / 1e6)), flush=True)
    print("fault", _lib.quant_fault())


if __name__ == "__main__":
    main()
