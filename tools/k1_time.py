"""Time the activation row quantiser alone: python tools/k1_time.py M K g p"""
import sys, os, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from smoothquant_mixedprecision_b200 import ops, _lib
if os.environ.get('MQ_RPS'): _lib.set_option('quant_rows_per_stage', int(os.environ['MQ_RPS']))
if os.environ.get('MQ_QWARPS'): _lib.set_option('quant_warps', int(os.environ['MQ_QWARPS']))
M, K, g = (int(v) for v in sys.argv[1:4]); p = float(sys.argv[4])
torch.manual_seed(0)
x = (torch.randn(M, K, device="cuda") * torch.exp(0.5 * torch.randn(K, device="cuda"))).half()
order = torch.randperm(K, device="cuda"); ks = max(1, int(p * K)) if p > 0 else 0
perm = order[ks:].int(); sal = torch.sort(order[:ks]).values.int() if ks else None
for _ in range(3): a = ops.quant_rows(x, perm, sal, g)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): a = ops.quant_rows(x, perm, sal, g)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
G = a.scales.shape[0]
nbytes = M * K * 2 + a.codes.numel() + G * M * 4 + (a.sal.numel() * 2 if a.sal is not None else 0)
print("K1T " + json.dumps(dict(M=M, K=K, g=g, p=p, ms=ms, gbs=nbytes / ms / 1e6)))
