"""Decode regime (BASELINE config 5): time mq_decode_gemm (int4-resident weights) and the prefill kernel on the
Llama-2-70B linear shapes at M in {1, 16, 64}:  python tools/decode_time.py [M ...]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from smoothquant_mixedprecision_b200 import ops, _lib
if os.environ.get("MQ_PDL"):
    _lib.set_option("pdl", int(os.environ["MQ_PDL"]))        # 2 = no programmatic dependent launch

SHAPES = [("70b_q_o", 8192, 8192), ("70b_k_v", 1024, 8192), ("70b_gate_up", 28672, 8192), ("70b_down", 8192, 28672),
          ("7b_qkvo", 4096, 4096)]
G_SIZE, P = 128, 0.10


def time_rotating(fns, iters=40):
    """Average GPU time of fns[i % len(fns)], replayed from a CUDA graph so that the host cost of the Python / ctypes
    call (~35 us, more than the kernel at these sizes) is not what is measured.  The operand sets together exceed the
    126 MB L2, so every call streams its weights from HBM (no explicit flush: a memset would leave the L2 full of dirty
    lines whose write-back the timed kernel then pays for)."""
    for f in fns:
        f()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            for i in range(iters):
                fns[i % len(fns)]()
    torch.cuda.synchronize()
    graph.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (3 * iters)


def measure(Ms, shapes=SHAPES, dev="cuda", l2_bytes=300 << 20):
    out = []
    for name, N, K in shapes:
        torch.manual_seed(N + K)
        order = torch.randperm(K, device=dev)
        ks = int(P * K)
        perm, sal = order[ks:].int(), torch.sort(order[:ks]).values.int()
        w = (torch.randn(N, K, device=dev) / K ** 0.5).half()
        wq = ops.quant_rows(w, perm, sal, G_SIZE)
        dw = ops.pack_decode(wq)
        del w
        copies = max(2, min(48, -(-l2_bytes // dw.nbytes())))
        dws = [dw] + [ops.DecodeWeights(nib=dw.nib.clone(), sal=dw.sal.clone(), rows=dw.rows, groups=dw.groups,
                                        group_size=dw.group_size, Ks=dw.Ks, dtype=dw.dtype) for _ in range(copies - 1)]
        pcopies = max(2, min(24, -(-l2_bytes // (wq.codes.numel() + wq.sal.numel() * 2))))
        wqs = [wq] + [ops.QuantizedRows(codes=wq.codes.clone(), scales=wq.scales.clone(), sal=wq.sal.clone(), rows=wq.rows,
                                        Kn=wq.Kn, Ks=wq.Ks, group_size=wq.group_size, int_groups=wq.int_groups,
                                        n_salient=wq.n_salient) for _ in range(pcopies - 1)]
        for M in Ms:
            x = (torch.randn(M, K, device=dev) * torch.exp(0.5 * torch.randn(K, device=dev))).half()
            a = ops.quant_rows(x, perm, sal, G_SIZE)
            ms_d = time_rotating([lambda d=d: ops.decode_gemm(a, d, None) for d in dws])
            ms_p = time_rotating([lambda q=q: ops.mixed_gemm(a, q, None) for q in wqs])
            wb = dw.nbytes() + a.codes.numel() + M * N * 2
            out.append(dict(shape=name, N=N, K=K, M=M, us_decode=ms_d * 1e3, us_prefill_kernel=ms_p * 1e3,
                            int4_weight_bytes=dw.nbytes(), gbs=wb / ms_d / 1e6, weight_copies=copies))
        del wq, dw, dws, wqs
    return out


if __name__ == "__main__":
    Ms = [int(v) for v in sys.argv[1:]] or [1, 16, 64]
    for r in measure(Ms):
        print("DEC " + json.dumps(r), flush=True)
