"""Time the activation row quantiser alone: python tools/k1_time.py M K g p [contig]
contig = 1: identity permutation over the first Kn columns, salient = the contiguous tail (the folded-producer layout).
MQ_H2=2 selects the r01 kernels (quant_rows_fast)."""
import sys, os, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from smoothquant_mixedprecision_b200 import ops, _lib
if os.environ.get('MQ_RPS'): _lib.set_option('quant_rows_per_stage', int(os.environ['MQ_RPS']))
if os.environ.get('MQ_QWARPS'): _lib.set_option('quant_warps', int(os.environ['MQ_QWARPS']))
if os.environ.get('MQ_H2'): _lib.set_option('quant_h2', int(os.environ['MQ_H2']))
M, K, g = (int(v) for v in sys.argv[1:4]); p = float(sys.argv[4])
contig = len(sys.argv) > 5 and sys.argv[5] == "1"
torch.manual_seed(0)
x = (torch.randn(M, K, device="cuda") * torch.exp(0.5 * torch.randn(K, device="cuda"))).half()
order = torch.randperm(K, device="cuda"); ks = max(1, int(p * K)) if p > 0 else 0
if contig:
    perm, Kn = None, K - ks
    sal = torch.arange(K - ks, K, device="cuda").int() if ks else None
else:
    perm = order[ks:].int(); Kn = None
    sal = torch.sort(order[:ks]).values.int() if ks else None
norm = os.environ.get("MQ_NORM")          # rms | layer: the fused norm + quantiser; "unfused": torch RMSNorm, then K1
w = (1 + 0.1 * torch.randn(K, device="cuda")).half()
b = torch.zeros(K, device="cuda").half() if norm == "layer" else None
def rms(t):
    h = t.float(); h = h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + 1e-6); return w * h.to(t.dtype)
def run():
    if norm in ("rms", "layer"): return ops.norm_quant_rows(x, w, b, 1e-6, norm, perm, sal, g, Kn=Kn)
    if norm == "unfused": return ops.quant_rows(rms(x), perm, sal, g, Kn=Kn)
    if norm == "unfused_layer": return ops.quant_rows(torch.nn.functional.layer_norm(x, (K,), w, w), perm, sal, g, Kn=Kn)
    return ops.quant_rows(x, perm, sal, g, Kn=Kn)
for _ in range(3): a = run()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): a = run()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
G = a.scales.shape[0]
nbytes = M * K * 2 + a.codes.numel() + G * M * 4 + (a.sal.numel() * 2 if a.sal is not None else 0)
print("K1T " + json.dumps(dict(M=M, K=K, g=g, p=p, contig=contig, h2=os.environ.get('MQ_H2', '1'), norm=norm, ms=ms, gbs=nbytes / ms / 1e6)))
