"""One decode-regime linear, eager launches (for ncu): python tools/decode_one.py M N K [iters]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from smoothquant_mixedprecision_b200 import ops
M, N, K = (int(v) for v in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 5
torch.manual_seed(0)
order = torch.randperm(K, device="cuda")
ks = int(0.10 * K)
perm, sal = order[ks:].int(), torch.sort(order[:ks]).values.int()
w = (torch.randn(N, K, device="cuda") / K ** 0.5).half()
wq = ops.quant_rows(w, perm, sal, 128)
dws = [ops.pack_decode(wq) for _ in range(3)]           # 3 copies > L2: every launch streams from HBM
x = (torch.randn(M, K, device="cuda") * torch.exp(0.5 * torch.randn(K, device="cuda"))).half()
a = ops.quant_rows(x, perm, sal, 128)
for i in range(iters):
    y = ops.decode_gemm(a, dws[i % 3], None)
torch.cuda.synchronize()
print("ok", tuple(y.shape), dws[0].nbytes())
