"""Time the mixed GEMM alone at one shape: python tools/gemm_time.py M N K g p [iters] [debug]"""
import sys, os, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from smoothquant_mixedprecision_b200 import ops, _lib

M, N, K, g = (int(v) for v in sys.argv[1:5])
p = float(sys.argv[5])
iters = int(sys.argv[6]) if len(sys.argv) > 6 else 10
debug = int(sys.argv[7]) if len(sys.argv) > 7 else 0
dev = "cuda"
torch.manual_seed(0)
x = (torch.randn(M, K, device=dev) * torch.exp(0.5 * torch.randn(K, device=dev))).half()
w = (torch.randn(N, K, device=dev) / K ** 0.5).half()
order = torch.randperm(K, device=dev)
ks = max(1, int(p * K)) if p > 0 else 0
perm = order[ks:].int()
sal = torch.sort(order[:ks]).values.int() if ks else None
frac = float(os.environ.get("MQ_INT_FRACTION", "1.0"))       # hybrid engine: share of the groups kept as integer codes
G = (perm.numel() + g - 1) // g
step = max(1, 128 // g)
gi = max(0, min(G, int(round(G * frac)) // step * step))
a = ops.quant_rows(x, perm, sal, g, int_groups=gi)
wq = ops.quant_rows(w, perm, sal, g, int_groups=gi)
_lib.set_option("gemm_debug", debug)
if os.environ.get("MQ_PAIR"):
    _lib.set_option("gemm_pair", int(os.environ["MQ_PAIR"]))      # 1: CTA-pair kernel, 2: single-CTA kernel
if debug:
    _lib.set_option("gemm_profile", 1)      # the debug bits live in the profiling variant of the kernel
for _ in range(3):
    y = ops.mixed_gemm(a, wq, None)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    y = ops.mixed_gemm(a, wq, None)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
if os.environ.get("MQ_PROFILE"):
    _lib.set_option("gemm_profile", 1)
    y = ops.mixed_gemm(a, wq, None)
    torch.cuda.synchronize()
    pr = _lib.gemm_profile(148)
    names = ["cycles", "ns", "epi_wait_ld", "mma_wait_full", "mma_wait_tmem", "epi_wait_acc", "epi_wait_scale", "epi_busy"]
    import statistics
    med = {n: statistics.median(r[i] for r in pr) for i, n in enumerate(names)}
    med["clock_ghz"] = med["cycles"] / max(med["ns"], 1)
    print("PROF " + json.dumps(med), flush=True)
    _lib.set_option("gemm_profile", 0)
print("GEMM " + json.dumps(dict(M=M, N=N, K=K, g=g, p=p, int_groups=gi, groups=G, pair=os.environ.get('MQ_PAIR', ''), debug=debug, ms=ms, tops=2.0 * M * N * K / ms / 1e9,
                                fault=_lib.gemm_fault()[:5])), flush=True)
