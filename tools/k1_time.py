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
for _ in range(3): a = ops.quant_rows(x, perm, sal, g, Kn=Kn)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): a = ops.quant_rows(x, perm, sal, g, Kn=Kn)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
G = a.scales.shape[0]
nbytes = M * K * 2 + a.codes.numel() + G * M * 4 + (a.sal.numel() * 2 if a.sal is not None else 0)
print("K1T " + json.dumps(dict(M=M, K=K, g=g, p=p, contig=contig, h2=os.environ.get('MQ_H2', '1'), ms=ms, gbs=nbytes / ms / 1e6)))
