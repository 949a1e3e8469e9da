"""Benchmark of the W4A4 MixedQuant linear hot path (BASELINE.json config 3).

Workload (one "step"): every W4A4 linear of Llama-2-7B (32 decoder layers x {q,k,v,o,gate,up,down} = 224 layers,
group 128, 5 % salient FP16 channels, static max-sort permutation, mode="fused") applied to one batch of 16 x 2048 =
32 768 tokens per GPU, chained through their real shapes (x -> q,k,v ; v -> o ; o -> gate,up ; up -> down -> next
layer).  Attention / norms / lm_head are not part of the path and are not executed.  Synthetic fp16 activations and
random-init weights.  N > 1: the HEADLINE is column-parallel (ONE 32 768-token batch, the 224 linears sharded over
out_features, "strong"); the data-parallel figure (one replica per GPU, no collective, "weak") is the `data_parallel`
sub-record.  `--parallel dp|tp` measures one of them alone.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

`value`   tokens/s with inputs resident in HBM (CUDA events, max over ranks)
`e2e`     the same through the module API with HOST buffers: pinned host activations -> device, the 224 layers,
          last output -> host, all inside the timed region
`roofline` of the dominant kernel (mixed GEMM), CUDA events around every launch inside the timed region
`cpu_baseline` the CPU oracle port of the reference layer on a bounded sample (rank 0, N = 1 only)
`gpu_reference` / `fp16_cublas`  (N = 1) the reference's fake-quant forward arithmetic (fake_quant.py:280-322) written
          with stock PyTorch CUDA ops, and plain fp16 F.linear, over one decoder layer's 7 linears at the full 32 768
          tokens on the same GPU (x 32 layers): the like-for-like GPU baselines of SURVEY 8(d)
`gemm_tops` (N = 1) the isolated mixed GEMM at the three Llama-2-7B linear shapes, M = 32 768
`column_parallel` (N > 1) comm / compute split of the headline: NCCL rank count, bytes received per rank, the step
          with every gather serialised (CUDA events around each collective) vs overlapped, the limiting collective
          with its bytes and GB/s
`decode`  (N = 1) config 5: mq_decode_gemm from INT4-resident weights at the Llama-2-70B / 7B linear shapes, M = 1, 16
`--impl reference` times that CPU path alone (all host threads) and prints the same line with "impl": "reference".
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

HIDDEN, INTER, LAYERS = 4096, 11008, 32
GROUP, SALIENT = 128, 0.05
SHAPES = {"q_proj": (HIDDEN, HIDDEN), "k_proj": (HIDDEN, HIDDEN), "v_proj": (HIDDEN, HIDDEN),
          "o_proj": (HIDDEN, HIDDEN), "gate_proj": (INTER, HIDDEN), "up_proj": (INTER, HIDDEN),
          "down_proj": (HIDDEN, INTER)}           # name -> (N, K)
METRIC = "llama2_7b_w4a4_linear_prefill_tokens_per_s"
# column-parallel plan (smoothquant_mixedprecision_b200/parallel.py: layer_plan_llama)
PLAN = [("q_proj", False), ("k_proj", False), ("v_proj", True), ("o_proj", True), ("gate_proj", False),
        ("up_proj", True), ("down_proj", True)]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tokens", type=int, default=16 * 2048, help="tokens per GPU per step")
    ap.add_argument("--layers", type=int, default=LAYERS)
    ap.add_argument("--cpu-tokens", type=int, default=4096, help="tokens of the bounded CPU sample")
    ap.add_argument("--cpu-layers", type=int, default=2, help="decoder layers of the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baselines", action="store_true")
    ap.add_argument("--parallel", default="auto", choices=["auto", "dp", "tp"],
                    help="N > 1: tp = column-parallel over out_features with NCCL all-gathers (strong scaling of one "
                         "batch); dp = one replica per GPU over different sequences (weak scaling, no collective); "
                         "auto = tp as the headline with the dp figure as the `data_parallel` sub-record")
    ap.add_argument("--tp-chunks", type=int, default=2,
                    help="tp: row chunks of the batch pipelined through the layers so that one chunk's all-gather "
                         "overlaps the other chunk's GEMM")
    ap.add_argument("--transport", default="auto", choices=["auto", "peer", "nccl"],
                    help="tp: how gathered outputs travel -- peer = copy engines over peer-mapped memory, ordered by "
                         "stream memory operations (parallel.PeerGather, no SMs); nccl = all_gather_into_tensor; "
                         "auto = peer up to 4 GPUs, nccl beyond (measured: profiles/r02_tp_sweep.txt)")
    ap.add_argument("--reserve-sms", type=int, default=None,
                    help="tp over NCCL: SMs the persistent GEMM / row quantiser leave free so that the collective's "
                         "kernel runs concurrently instead of queueing behind them (default 0: measured no gain, "
                         "profiles/r02_tp_sweep.txt)")
    ap.add_argument("--nccl-max-ctas", type=int, default=None, help="tp over NCCL: NCCL_MAX_CTAS (default: NCCL's own)")
    ap.add_argument("--peer-mode", default="pull", choices=["pull", "push", "mcast"],
                    help="peer transport: pull peers' blocks, push the own block to every peer, or push once to the "
                         "multicast address of the gather buffers")
    ap.add_argument("--peer-lanes", type=int, default=None, help="peer transport: copy streams pulling in parallel")
    ap.add_argument("--no-decode", action="store_true")
    return ap.parse_args()


def ops_per_token(layers):
    return 2.0 * layers * sum(n * k for n, k in SHAPES.values())


# ---------------------------------------------------------------------------------------------- CPU arm (oracle)
def cpu_layer_sample(tokens, threads, n_layers):
    """`n_layers` decoder layers' 7 linears each through the oracle's literal reference layer (fp32, like the
    reference's own CPU runs, examples/group_size_analysis.py:62-66).  Returns a closure running one pass and the setup
    seconds."""
    import torch
    from oracle import mq_oracle as O
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(0)
    t0 = time.perf_counter()
    xs = {K: torch.randn(tokens, K, generator=g) * torch.exp(0.5 * torch.randn(K, generator=g))
          for K in (HIDDEN, INTER)}
    stack = []
    for _ in range(n_layers):
        layers = {}
        for name, (N, K) in SHAPES.items():
            w = torch.randn(N, K, generator=g) / K ** 0.5
            imp = xs[K].abs().mean(dim=0)
            sal = O.select_salient(imp, SALIENT)
            layers[name] = (O.w4a4_weight_reference(w, sal, 4, GROUP), sal, K)
        stack.append(layers)
    setup = time.perf_counter() - t0

    def run():
        out = None
        for layers in stack:
            for name, (wq, sal, K) in layers.items():
                out = O.w4a4_forward_reference(xs[K], wq, None, sal, 4, GROUP)
        return out
    return run, setup


def time_cpu(tokens, steps, warmup, n_layers=2):
    import torch
    threads = os.cpu_count() or 1
    run, setup = cpu_layer_sample(tokens, threads, n_layers)
    for _ in range(warmup):
        run()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        run()
        ts.append(time.perf_counter() - t0)
    per_layer = sum(ts) / len(ts) / n_layers
    return {"tok_s": tokens / (per_layer * LAYERS), "s_per_layer": per_layer, "threads": threads, "setup_s": setup,
            "sample": f"{n_layers} of {LAYERS} decoder layers (7 linears each), {tokens} tokens, fp32, oracle port of "
                      f"W4A4Linear.forward; tokens/s = tokens / (t_layer * {LAYERS}) -- every layer has the same "
                      f"shapes and the layer is row-independent apart from the column abs-max, so time is linear in "
                      f"layers and tokens",
            "torch_threads": torch.get_num_threads()}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # bounded sample: the largest of (layers, tokens) whose (steps + warmup) passes fit ~150 s on this host, sized from
    # a probe pass of 1 layer x 1024 tokens (time is linear in both, see `sample`)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    probe = time_cpu(1024, 1, 1, 1)
    t_probe = probe["s_per_layer"]
    tokens, n_layers = 1024, 1
    for cand_layers, cand_tokens in ((args.cpu_layers, args.cpu_tokens), (1, args.cpu_tokens), (1, 2048)):
        if t_probe * (cand_tokens / 1024.0) * cand_layers * (steps + warmup) <= 150.0:
            tokens, n_layers = cand_tokens, cand_layers
            break
    r = time_cpu(tokens, steps, warmup, n_layers)
    line = {"impl": "reference", "metric": METRIC, "value": r["tok_s"], "unit": "tokens/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["s_per_layer"] * LAYERS * 1e3 *
            (args.tokens / tokens), "higher_is_better": True,
            # same job as the GPU arm at this N: one 32768-token batch (column-parallel headline) for N > 1
            "scaling": "strong" if args.gpus > 1 and args.parallel != "dp" else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "Llama-2-7B W4A4 linears g=128 5% salient, CPU fake-quant path", "tokens_per_step":
                       args.tokens, "group_size": GROUP, "salient_prop": SALIENT},
            "cpu_baseline": {"value": r["tok_s"], "unit": "tokens/s", "cores": r["threads"], "kind": "port",
                             "sample": r["sample"]},
            "e2e": {"value": r["tok_s"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------- GPU arm
class ClockSampler:
    """nvidia-smi sampling of SM clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                      stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out = self.p.communicate(timeout=5)[0]
        except subprocess.TimeoutExpired:
            self.p.kill()
            out = ""
        clocks, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in out.splitlines():
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6 or not f[0].isdigit():
                continue
            clocks.append(int(f[0]))
            mx = int(f[1])
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        busy = [c for c in clocks if mx and c > 0.3 * mx] or clocks
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(clocks)}


# ---------------------------------------------------------------------------------------------- GPU baselines
def torch_reference_forward(x, wq, keep_mask, g):
    """The reference layer's forward arithmetic (fake_quant.py:291-306 -> :105-154) with stock PyTorch CUDA ops:
    boolean-mask compaction of the non-salient columns, per-call column abs-max, argsort, gather, zero pad, per-group
    abs-max scale, div / round / mul in place, inverse permutation, scatter back, F.linear on the fake-quantised fp16
    operands.  Baseline only -- nothing of this repo's kernels runs here."""
    import torch
    import torch.nn.functional as F
    t = x[:, keep_mask]
    perm = torch.argsort(t.abs().amax(dim=0))
    tp = t[:, perm]
    C = tp.shape[1]
    G = -(-C // g)
    if G * g != C:
        tp = F.pad(tp, (0, G * g - C))
    tg = tp.view(tp.shape[0], G, g)
    sc = tg.abs().amax(dim=-1, keepdim=True).clamp_(min=1e-5).div_(7)
    tg.div_(sc).round_().mul_(sc)
    q = x.clone()
    q[:, keep_mask] = tp[:, :C][:, torch.argsort(perm)]
    return F.linear(q, wq)


def gpu_baselines(mods, x0, steps=3):
    """One decoder layer's 7 linears at the full token count: (a) the reference's fake-quant arithmetic under stock
    PyTorch, (b) plain fp16 F.linear (cuBLAS).  Returns seconds per decoder layer for each."""
    import torch
    import torch.nn.functional as F
    dev = x0.device
    ws, masks = {}, {}
    for name, m in mods.items():
        ws[name] = m.weight.contiguous()              # fake-quantised fp16 weight [N, K] (dequantised view)
        mask = torch.ones(m.in_features, dtype=torch.bool, device=dev)
        mask[m.sal_sorted.long()] = False
        masks[name] = mask

    def chain(fn):
        x = x0
        fn("q_proj", x)
        fn("k_proj", x)
        v = fn("v_proj", x)
        o = fn("o_proj", v)
        fn("gate_proj", o)
        u = fn("up_proj", o)
        return fn("down_proj", u)

    def timeit(fn):
        chain(fn)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            chain(fn)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / steps

    t_ref = timeit(lambda n, x: torch_reference_forward(x, ws[n], masks[n], GROUP))
    t_fp16 = timeit(lambda n, x: F.linear(x, ws[n]))
    return t_ref, t_fp16


def isolated_gemm_tops(mods, x_by_k, iters=10):
    """The mixed GEMM alone (activation already quantised) at the three Llama-2-7B linear shapes, M = tokens."""
    import torch
    from smoothquant_mixedprecision_b200 import ops
    out = {}
    for name in ("q_proj", "gate_proj", "down_proj"):
        m = mods[name]
        x = x_by_k[m.in_features]
        a = ops.quant_rows(x, m.perm, m.sal_sorted, m.group_size, m.quant_bits, Kn=m.perm.numel(),
                           int_groups=m.int_groups)
        w = ops.QuantizedRows(codes=m.w_codes, scales=m.w_scales, sal=m.w_sal, rows=m.out_features, Kn=a.Kn, Ks=a.Ks,
                              group_size=m.group_size, int_groups=a.int_groups)
        for _ in range(3):
            ops.mixed_gemm(a, w, None)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            ops.mixed_gemm(a, w, None)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        M, N, K = x.shape[0], m.out_features, m.in_features
        out[f"{M}x{N}x{K}"] = {"ms": ms, "tops": 2.0 * M * N * K / (ms * 1e-3) / 1e12}
    return out


# tcgen05 issue-rate ceilings measured with back-to-back MMAs on all SMs (tools/ubench.cu, profiles/r01_ubench.log)
UBENCH = {"i8_tops": 4260.0, "f16_tflops": 1666.0, "at_mhz": 1831.0}
NOMINAL = {"i8_tops": 4500.0, "f16_tflops": 2250.0}


def build_stack(layers, tokens, dev, seed, tp=None):
    """Random-init Llama-2-7B linears packed for mode="fused" (P0 on the GPU) + synthetic calibration statistics."""
    import torch
    from smoothquant_mixedprecision_b200.fake_quant import W4A4Linear
    g = torch.Generator(device=dev).manual_seed(seed)
    chan = {K: torch.exp(0.5 * torch.randn(K, device=dev, generator=g)) for K in (HIDDEN, INTER)}
    stack = []
    for _ in range(layers):
        mods = {}
        # one calibration vector per distinct INPUT tensor: q/k/v see the same hidden state and gate/up the same
        # attention output, so the reference's hooks give them identical statistics (fake_quant.py:396-455)
        feat = {src: chan[K] * (1 + 0.05 * torch.rand(K, device=dev, generator=g))
                for src, K in (("x", HIDDEN), ("attn", HIDDEN), ("o", HIDDEN), ("mlp", INTER))}
        src_of = {"q_proj": "x", "k_proj": "x", "v_proj": "x", "o_proj": "attn", "gate_proj": "o", "up_proj": "o",
                  "down_proj": "mlp"}
        for name, (N, K) in SHAPES.items():
            lin = torch.nn.Linear(K, N, bias=False, device=dev, dtype=torch.float16)
            with torch.no_grad():
                lin.weight.copy_(torch.randn(N, K, device=dev, generator=g) / K ** 0.5)
            # synthetic calibration: importance = mean|x| (run_experiments.py:60), sort key = max|x| (calibration.py:18-25)
            imp = feat[src_of[name]]
            mods[name] = W4A4Linear.from_float(lin, "per_group", "per_group", importance=imp.cpu(),
                                               salient_prop=SALIENT, group_size=GROUP, mode="fused",
                                               sort_key=imp * 4.0)
            del lin
        # down_proj's static permutation folded into gate / up (SiLU * mul commutes with a channel permutation):
        # its quantiser then reads contiguous vectors (fake_quant.fold_mlp_permutation)
        from smoothquant_mixedprecision_b200.fake_quant import fold_mlp_permutation
        fold_mlp_permutation(mods["gate_proj"], mods["up_proj"], mods["down_proj"])
        if tp is not None:            # column-parallel: gather only where the consumer needs the full width
            for name in mods:
                mods[name].shard_(tp[0], tp[1], gather_output=dict(PLAN)[name], blocked_output=True)
        stack.append(mods)
    x0 = (torch.randn(tokens, HIDDEN, device=dev, generator=g) * chan[HIDDEN]).half()
    return stack, x0


def run_stack_tp(stack, xs):
    """Column-parallel pass over row chunks `xs` of the batch.  Gathered layers (v, o, up, down) return
    parallel.BlockedColumns whose all-gather is still in flight; the collective of one chunk overlaps the GEMMs issued
    after it (the column-local q / k / gate of the same chunk and everything of the other chunk), and the consumer's
    row quantiser reads the gathered [p, M, S] buffer in place."""
    for m in stack:
        vs = []
        for x in xs:
            vs.append(m["v_proj"](x))
            m["q_proj"](x)
            m["k_proj"](x)
        os_ = [m["o_proj"](v) for v in vs]
        us = []
        for o in os_:
            us.append(m["up_proj"](o))
            m["gate_proj"](o)
        xs = [m["down_proj"](u) for u in us]
    return xs


def run_stack(stack, x):
    for m in stack:
        m["q_proj"](x)
        m["k_proj"](x)
        v = m["v_proj"](x)
        o = m["o_proj"](v)
        m["gate_proj"](o)
        u = m["up_proj"](o)
        x = m["down_proj"](u)
    return x


def gpu_arm(args):
    import torch
    import torch.distributed as dist
    from smoothquant_mixedprecision_b200 import ops, _lib
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if args.nccl_max_ctas is not None:
            os.environ["NCCL_MAX_CTAS"] = str(args.nccl_max_ctas)
        dist.init_process_group("nccl", device_id=dev)
    M = args.tokens
    mode = args.parallel
    if mode == "auto":
        mode = "tp" if world > 1 else "dp"
    tp = mode == "tp" and world > 1

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        sync_all()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps

    def measure(stack, x0, use_tp, clocks_on):
        """resident + e2e timing of one stack; returns the numbers of the JSON line"""
        hx = torch.empty(x0.shape, dtype=x0.dtype, pin_memory=True).copy_(x0)
        hy = torch.empty(x0.shape, dtype=x0.dtype, pin_memory=True)
        xd = torch.empty_like(x0)
        nch = max(1, args.tp_chunks) if use_tp else 1
        rows = [(i * M // nch, (i + 1) * M // nch) for i in range(nch)]

        def run(x):
            if not use_tp:
                return run_stack(stack, x)
            outs = run_stack_tp(stack, [x[a:b] for a, b in rows])
            return outs

        def step_resident():
            run(x0)

        def step_e2e():
            xd.copy_(hx, non_blocking=True)
            y = run(xd)
            if use_tp:
                for (a, b), yc in zip(rows, y):
                    hy[a:b].copy_(yc.dense() if hasattr(yc, "dense") else yc, non_blocking=True)
            else:
                hy.copy_(y, non_blocking=True)
            torch.cuda.current_stream().synchronize()      # the caller holds the result on the host

        for _ in range(max(3, args.warmup)):
            step_resident()
        c0 = ops.COUNTERS["launch_calls"]
        ops.GEMM_EVENTS, ops.QUANT_EVENTS = [], []
        sampler = ClockSampler(local) if (rank == 0 and clocks_on) else None
        ms = timed(step_resident, args.steps)
        clocks = sampler.stop() if sampler else None
        r = {"ms": ms, "clocks": clocks, "launches": (ops.COUNTERS["launch_calls"] - c0) // args.steps}
        ev, ops.GEMM_EVENTS = ops.GEMM_EVENTS, None
        qev, ops.QUANT_EVENTS = ops.QUANT_EVENTS, None
        r["quant_ms"] = sum(a.elapsed_time(b) for a, b, _ in qev) / args.steps
        r["quant_bytes"] = sum(e[2] for e in qev) / args.steps
        r["quant_launches"] = len(qev) // args.steps
        r["gemm_ms"] = sum(a.elapsed_time(b) for a, b, _, _ in ev) / args.steps
        r["ops_i8"] = sum(e[2] for e in ev) / args.steps
        r["ops_f16"] = sum(e[3] for e in ev) / args.steps
        r["gemm_launches"] = len(ev) // args.steps
        step_e2e()
        r["ms_e2e"] = timed(step_e2e, args.steps)
        r["h2d"], r["d2h"] = hx.numel() * 2, hy.numel() * 2
        if use_tp:
            # comm / compute split: the same step with every gather serialised behind its GEMM and timed on the
            # stream (CUDA events around the blocking collective), max over ranks of the per-step totals
            from smoothquant_mixedprecision_b200 import parallel as par
            par.ASYNC_GATHER, par.COMM_EVENTS = False, []
            ops.GEMM_EVENTS, ops.QUANT_EVENTS = [], []
            step_resident()
            par.COMM_EVENTS, ops.GEMM_EVENTS, ops.QUANT_EVENTS = [], [], []
            ms_serial = timed(step_resident, args.steps)
            cev, par.COMM_EVENTS, par.ASYNC_GATHER = par.COMM_EVENTS, None, True
            gev, ops.GEMM_EVENTS = ops.GEMM_EVENTS, None
            kev, ops.QUANT_EVENTS = ops.QUANT_EVENTS, None
            by_bytes = {}
            for a, b, nbytes in cev:
                d = by_bytes.setdefault(nbytes, [0.0, 0])
                d[0] += a.elapsed_time(b)
                d[1] += 1
            comm_ms = sum(v[0] for v in by_bytes.values()) / args.steps
            t = torch.tensor([comm_ms, sum(a.elapsed_time(b) for a, b, _, _ in gev) / args.steps,
                              sum(a.elapsed_time(b) for a, b, _ in kev) / args.steps], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            worst = max(by_bytes.items(), key=lambda kv: kv[1][0])
            per_call_ms = worst[1][0] / worst[1][1]
            r["comm"] = {
                "nccl_ranks": world, "nccl_version": ".".join(str(v) for v in torch.cuda.nccl.version()),
                "collectives_per_step": len(cev) // args.steps, "row_chunks": nch,
                "bytes_received_per_rank_per_step": sum(k * v[1] for k, v in by_bytes.items()) // args.steps,
                "serialised": {"ms_per_step": ms_serial, "comm_ms": float(t[0]), "gemm_ms": float(t[1]),
                               "quantiser_ms": float(t[2])},
                "overlapped": {"ms_per_step": ms, "exposed_comm_ms": max(0.0, ms - (ms_serial - float(t[0])))},
                "limiting_collective": {
                    "what": "all-gather of one row chunk of a gathered layer output "
                            f"([{world}, {rows[0][1] - rows[0][0]}, S] fp16, in place)",
                    "bytes_received_per_rank": worst[0], "calls_per_step": worst[1][1] // args.steps,
                    "ms_per_call": per_call_ms, "algbw_GBs": worst[0] / per_call_ms / 1e6,
                    "busbw_note": "bytes received per rank / time: (p-1)/p of the gathered tensor over NVLink"}}
        return r

    transport = None
    if tp:
        from smoothquant_mixedprecision_b200 import parallel as par
        transport = "nccl"
        if args.transport == "peer" or (args.transport == "auto" and world <= 4):
            try:
                par.enable_peer_gather(lanes=args.peer_lanes, mode=args.peer_mode)
                transport = "peer"
            except Exception as e:      # no peer-mapped memory on this box: NCCL carries the gathers
                print(f"bench.py: peer transport unavailable ({type(e).__name__}: {e}); using NCCL", file=sys.stderr)
        ok = torch.tensor([1 if transport == "peer" else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            par.TRANSPORT, transport = "nccl", "nccl"
        reserve = args.reserve_sms or 0
        if reserve:
            _lib.set_option("sm_reserve", reserve)
    stack, x0 = build_stack(args.layers, M, dev, seed=1234 + (0 if tp else rank), tp=(rank, world) if tp else None)
    job_tokens = M if tp else world * M
    main = measure(stack, x0, tp, True)
    other = None
    if world > 1 and args.parallel == "auto":           # the data-parallel figure beside the column-parallel headline
        _lib.set_option("sm_reserve", 0)
        del stack
        torch.cuda.empty_cache()
        stack, x0 = build_stack(args.layers, M, dev, seed=1234 + rank, tp=None)
        other = measure(stack, x0, False, False)
    ms, ms_e2e, clocks, launches = main["ms"], main["ms_e2e"], main["clocks"], main["launches"]
    quant_ms, quant_bytes, gemm_ms = main["quant_ms"], main["quant_bytes"], main["gemm_ms"]
    ops_i8, ops_f16 = main["ops_i8"], main["ops_f16"]
    fault = _lib.gemm_fault()[0] or _lib.quant_fault()[0]
    if fault:
        raise SystemExit(f"bench.py: kernel fault record {fault:#x}")
    if rank != 0:
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    bf16 = peaks.get("bf16_tflops_sustained") or 1400.0          # fallback: B200_PROFILING.md sustained figure
    peak_src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained; int8 taken as 2x)" if peaks else \
        "fallback 1.4 PF bf16 sustained; int8 taken as 2x"
    traffic = {}                                                  # per-launch DRAM bytes from the committed ncu pass
    try:
        import glob
        tfile = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")))[-1]
        traffic = {k: v["dram_bytes_per_launch"] for k, v in json.load(open(tfile))["kernels"].items()}
    except (IndexError, OSError, KeyError, ValueError):
        pass
    t_roof_ms = (ops_i8 / (2 * bf16 * 1e12) + ops_f16 / (bf16 * 1e12)) * 1e3
    achieved = (ops_i8 + ops_f16) / (gemm_ms * 1e-3) / 1e12
    peak_mix = (ops_i8 + ops_f16) / (t_roof_ms * 1e-3) / 1e12

    def frac_against(p_i8, p_f16):          # same serial-pipe roofline with other peak pairs (TOP/s)
        return ((ops_i8 / (p_i8 * 1e12) + ops_f16 / (p_f16 * 1e12)) * 1e3) / gemm_ms

    clk = (clocks or {}).get("sm_mhz") or UBENCH["at_mhz"]
    ub_scale = clk / UBENCH["at_mhz"]
    fractions = {
        "vs_2x_bf16_sustained_measured": frac_against(2 * bf16, bf16),
        "vs_tcgen05_ubench_at_bench_clock": frac_against(UBENCH["i8_tops"] * ub_scale, UBENCH["f16_tflops"] * ub_scale),
        "vs_nominal_4500_2250": frac_against(NOMINAL["i8_tops"], NOMINAL["f16_tflops"]),
        "ubench_source": "tools/ubench.cu back-to-back tcgen05.mma on 148 SMs: kind::i8 4260 TOP/s, kind::f16 1666 "
                         f"TFLOP/s at 1831 MHz, scaled to the median SM clock of this run ({clk} MHz)"}
    line = {"metric": METRIC, "value": job_tokens / (ms * 1e-3), "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong" if tp else "weak", "vs_baseline": None,
            "dtype": "s8 codes (int32 accumulate) + f16 salient",
            "data": "synthetic",
            "config": {"workload": "Llama-2-7B W4A4 linears (224 layers), g=128, 5% salient FP16, static max-sort, "
                                   "prefill 16x2048 tokens per GPU (dp) or per job (tp)", "tokens_per_gpu": M, "layers": args.layers,
                       "group_size": GROUP, "salient_prop": SALIENT,
                       "parallelism": (f"tp{world}: column-parallel over out_features, ONE {M}-token batch per job; "
                                       "v/o/up/down outputs all-gathered in place into [p, M, S] (" + str(transport) + ", async, "
                                       f"{max(1, args.tp_chunks)} row chunks) and quantised from that layout") if tp
                       else f"dp{world}",
                       "l2": "inputs (268 MB activations, 6.9 GB weights) exceed the 126 MB L2; no flush"},
            "tops": job_tokens * ops_per_token(args.layers) / (ms * 1e-3) / 1e12,
            "e2e": {"value": job_tokens / (ms_e2e * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": main["h2d"],
                    "d2h_bytes_per_step": main["d2h"], "ms_per_step": ms_e2e},
            "gpu_launches": launches,
            "roofline": {"kernel": "mixed_gemm_kernel", "bound": "tensor", "achieved": achieved, "peak": peak_mix,
                         "unit": "TOP/s", "frac": achieved / peak_mix, "traffic": traffic.get("mixed_gemm_kernel"),
                         "traffic_unit": "DRAM bytes per launch (ncu, profiles/*_traffic.json)", "peak_source": peak_src,
                         "share_of_step": gemm_ms / ms, "launches_per_step": main["gemm_launches"],
                         "fractions": fractions,
                         "ops_note": "ops = 2*M*N*Kn on the int8 pipe (true quantised columns, no padding) + 2*M*N*Ks "
                                     "on the f16 pipe"},
            "roofline_quantiser": {"kernel": "quant_rows_h2", "bound": "hbm",
                                   "achieved": quant_bytes / (quant_ms * 1e-3) / 1e9,
                                   "peak": peaks.get("hbm_gbs") or 6500.0, "unit": "GB/s",
                                   "frac": quant_bytes / (quant_ms * 1e-3) / 1e9 / (peaks.get("hbm_gbs") or 6500.0),
                                   "traffic": traffic.get("quant_rows_h2", traffic.get("quant_rows_fast")),
                                   "algorithmic_bytes_per_launch": quant_bytes / max(1, main["quant_launches"]),
                                   "share_of_step": quant_ms / ms,
                                   "launches_per_step": main["quant_launches"]},
            "clocks": clocks}
    if tp:
        line["column_parallel"] = main["comm"]
        line["column_parallel"]["transport"] = (
            f"peer-mapped memory ({args.peer_mode}): copy-engine transfers over NVLink ordered by stream memory operations "
            "(mq_peer_gather / mq_peer_push, no SMs); process group (NCCL) used for set-up and the timing barriers only"
            if transport == "peer"
            else f"NCCL all_gather_into_tensor, in place; NCCL_MAX_CTAS={os.environ.get('NCCL_MAX_CTAS', 'default')}, "
                 f"{reserve} SMs left free by the persistent GEMM / row quantiser")
    if other is not None:
        line["data_parallel"] = {
            "value": world * M / (other["ms"] * 1e-3), "unit": "tokens/s", "scaling": "weak",
            "ms_per_step": other["ms"], "tokens_per_gpu": M,
            "e2e": {"value": world * M / (other["ms_e2e"] * 1e-3), "unit": "tokens/s"},
            "what": f"one replica per GPU over different sequences, no data-path collective (dp{world})"}
    if world == 1 and not args.no_gpu_baselines:
        xk = {HIDDEN: x0, INTER: torch.randn(M, INTER, device=dev, dtype=torch.float16)}
        line["gemm_tops"] = isolated_gemm_tops(stack[0], xk)
        del xk
        t_ref, t_fp16 = gpu_baselines(stack[0], x0)
        line["gpu_reference"] = {
            "value": M / (t_ref * LAYERS), "unit": "tokens/s", "ms_per_decoder_layer": t_ref * 1e3,
            "what": "the reference's fake-quant W4A4Linear.forward arithmetic (fake_quant.py:280-322) under stock "
                    "PyTorch CUDA ops, fp16, 1 decoder layer (7 linears) x 32768 tokens, x 32"}
        line["fp16_cublas"] = {
            "value": M / (t_fp16 * LAYERS), "unit": "tokens/s", "ms_per_decoder_layer": t_fp16 * 1e3,
            "what": "plain fp16 F.linear (cuBLAS) on the same 7 shapes x 32768 tokens, x 32"}
    if world == 1 and not args.no_decode:
        # config 5 (decode regime): mq_decode_gemm from INT4-resident weights, CUDA-graph replay over operand copies
        # larger than L2 (tools/decode_time.py); bytes = int4 weights + scales + salient slice + activations + y
        del stack
        torch.cuda.empty_cache()
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import decode_time
        hbm = peaks.get("hbm_gbs") or 6500.0
        line["decode"] = {
            "what": "one linear, M tokens, g=128, 10% salient, weights INT4-resident (4 bits per non-salient weight); "
                    "us per call, achieved GB/s on the bytes the call must move, fraction of the measured HBM peak; "
                    "us_prefill_kernel = the tcgen05 prefill kernel on the same shape (int8-resident codes)",
            "hbm_peak_GBs": hbm,
            "points": [dict(shape=r["shape"], N=r["N"], K=r["K"], M=r["M"], us=round(r["us_decode"], 2),
                            us_prefill_kernel=round(r["us_prefill_kernel"], 2), GBs=round(r["gbs"], 1),
                            frac_hbm=round(r["gbs"] / hbm, 3)) for r in decode_time.measure([1, 16])]}
    if world == 1 and not args.no_cpu_baseline:
        r = time_cpu(args.cpu_tokens, 2, 1, args.cpu_layers)
        line["cpu_baseline"] = {"value": r["tok_s"], "unit": "tokens/s", "cores": r["threads"], "kind": "port",
                                "sample": r["sample"]}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
