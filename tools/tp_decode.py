"""Config 5 under column parallelism: one W4A4 linear of Llama-2-70B at decode batch sizes, sharded over the GPUs of the
box, int4-resident weights (mq_decode_gemm), output all-gathered in place.

    python -m torch.distributed.run --nproc-per-node N tools/tp_decode.py [M ...]

Per shape and M: microseconds per layer call (row quantiser + decode GEMM on N / world rows + gather), replayed from a
CUDA graph over enough weight copies that every call streams its weights from HBM (a rank's shard of one layer would
otherwise sit in the 126 MB L2), max over ranks.  `us_local` is the same without the gather (gather_output=False)."""
import copy
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from smoothquant_mixedprecision_b200 import parallel as par
from smoothquant_mixedprecision_b200.fake_quant import W4A4Linear, clear_act_cache

SHAPES = [("70b_q_o", 8192, 8192), ("70b_gate_up", 28672, 8192), ("70b_down", 8192, 28672)]
G_SIZE, P = 128, 0.10


def graph_time(fn, calls, dev):
    """us per call of fn(i), i = 0..calls-1, captured once and replayed 3 times"""
    for i in range(min(calls, 3)):
        fn(i)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for i in range(calls):
                fn(i)
    torch.cuda.synchronize()
    g.replay()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / (3 * calls) * 1e3], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    Ms = [int(v) for v in sys.argv[1:]] or [1, 16]
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    par.ASYNC_GATHER = False                 # a decode layer's consumer follows immediately: nothing to overlap
    out = []
    for name, N, K in SHAPES:
        torch.manual_seed(N + K)
        lin = torch.nn.Linear(K, N, bias=False, device=dev, dtype=torch.float16)
        with torch.no_grad():
            lin.weight.copy_(torch.randn(N, K, device=dev) / K ** 0.5)
        imp = torch.rand(K)
        base = W4A4Linear.from_float(lin, "per_group", "per_group", importance=imp, salient_prop=P, group_size=G_SIZE,
                                     mode="fused", sort_key=imp, int_fraction=1.0)
        del lin
        variants = {}
        for tag, gather in (("us", True), ("us_local", False)):
            m = copy.deepcopy(base)
            if world > 1:
                m.shard_(rank, world, gather_output=gather, blocked_output=True)
            m.prepare_decode(keep_prefill=False)
            per = m.device_weight_bytes()
            copies = max(2, min(64, -(-(300 << 20) // per)))
            variants[tag] = [m] + [copy.deepcopy(m) for _ in range(copies - 1)]
        del base
        for M in Ms:
            x = (torch.randn(M, K, device=dev) * torch.exp(0.5 * torch.randn(K, device=dev))).half()
            rec = dict(shape=name, N=N, K=K, M=M, world=world, int4_weight_bytes_per_rank=variants["us"][0].device_weight_bytes(),
                       weight_copies=len(variants["us"]))
            for tag, mods in variants.items():
                def call(i, mods=mods):
                    clear_act_cache()
                    y = mods[i % len(mods)](x)
                    if hasattr(y, "wait"):
                        y.wait()
                rec[tag] = round(graph_time(call, 2 * len(mods), dev), 2)
            out.append(rec)
            if rank == 0:
                print("TPDEC " + json.dumps(rec), flush=True)
        del variants
        torch.cuda.empty_cache()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
