"""Column-parallel / data-parallel host logic (smoothquant_mixedprecision_b200/parallel.py).

CPU part: world_size-2 gloo processes exercise the shard bookkeeping and the gather / all-reduce plumbing on plain
tensors (no compute kernels).  GPU part: a fused layer split into column shards on one device reproduces the
unsharded output bit for bit (the shards run the same kernels on row slices of the packed weight)."""
import copy
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from smoothquant_mixedprecision_b200 import parallel as P


def test_shard_ranges_cover_and_align():
    for N in (8, 64, 768, 1000, 4096, 11008, 28672):
        for world in (1, 2, 3, 4, 8):
            S = P.shard_size(N, world)
            assert S % 8 == 0 and S * world >= N
            spans = [P.shard_range(N, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == N
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and a <= b and c <= d
    with pytest.raises(ValueError):
        P.shard_range(64, 2, 2)
    # 7B shapes at 8 GPUs: 512 and 1376 columns per rank
    assert P.shard_range(4096, 8, 3) == (1536, 2048) and P.shard_range(11008, 8, 7) == (9632, 11008)


def test_split_sequences():
    for n in (1, 7, 16, 33):
        for world in (1, 2, 4, 8):
            spans = [P.split_sequences(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1 and all(b == c for (_, b), (c, _) in zip(spans, spans[1:]))


def test_shard_scales_layout():
    sc = torch.arange(3 * 24, dtype=torch.float32).view(3, 24)
    out = P.shard_scales(sc, 20, 2, 1)             # S = 16 -> rank 1 owns columns 16..19
    assert out.shape == (3, 4) and torch.equal(out, sc[:, 16:20])
    out0 = P.shard_scales(sc, 20, 2, 0)
    assert out0.shape == (3, 16) and torch.equal(out0, sc[:, :16])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, N, M):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(7)
        full = torch.randn(M, N, generator=g).half()
        lo, hi = P.shard_range(N, world, rank)
        got = P.gather_columns(full[:, lo:hi].contiguous(), N)
        assert got.shape == (M, N) and torch.equal(got, full), f"rank {rank}: gather mismatch"
        out = torch.empty(M, N, dtype=torch.float16)
        assert P.gather_columns(full[:, lo:hi].contiguous(), N, out=out) is out and torch.equal(out, full)
        # blocked gather: same values, left in the collective's own [world, M, S] layout (sync, async, in-place block)
        for async_op in (False, True):
            bc = P.gather_columns_blocked(full[:, lo:hi].contiguous(), N, async_op=async_op)
            S = P.shard_size(N, world)
            assert tuple(bc.buf.shape) == (world, M, S) and tuple(bc.shape) == (M, N) and bc.dim() == 2
            assert torch.equal(bc.dense(), full), f"rank {rank}: blocked gather mismatch"
            for r in range(world):
                a, b = P.shard_range(N, world, r)
                assert torch.equal(bc.buf[r][:, : b - a], full[:, a:b])
        buf = P.gather_buffer(M, N, world, torch.float16, "cpu")
        buf[rank][:, : hi - lo].copy_(full[:, lo:hi])
        bc = P.gather_columns_blocked(buf[rank][:, : hi - lo], N, buf=buf)
        assert bc.buf is buf and torch.equal(bc.dense(), full)
        key = torch.full((5,), float(rank))
        key[rank] = 10.0
        red = P.allreduce_colmax(key.clone())
        want = torch.full((5,), float(world - 1))
        want[:world] = 10.0
        assert torch.equal(red, want), (rank, red)
        # per-rank shards of a per-output-channel table reassemble to the table
        table = torch.arange(N * 3, dtype=torch.float32).view(N, 3)
        mine = P.shard_rows(table, N, world, rank)
        parts = [torch.empty(P.shard_size(N, world), 3) for _ in range(world)]
        pad = torch.zeros(P.shard_size(N, world), 3)
        pad[: mine.shape[0]] = mine
        dist.all_gather(parts, pad)
        assert torch.equal(torch.cat(parts)[:N], table)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("N,M", [(64, 5), (1000, 3), (11008, 2)])
def test_gloo_world2_gather_and_allreduce(N, M):
    mp.spawn(_worker, args=(2, _free_port(), N, M), nprocs=2, join=True)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 8])
def test_column_shards_reproduce_unsharded_layer(world):
    from smoothquant_mixedprecision_b200.fake_quant import W4A4Linear
    torch.manual_seed(3)
    M, K, N, g = 200, 1024, 1000, 128
    lin = torch.nn.Linear(K, N).half().cuda()
    x = (torch.randn(M, K, device="cuda") * torch.exp(0.5 * torch.randn(K, device="cuda"))).half()
    imp = x.abs().float().mean(dim=0).cpu()
    for mode, frac in (("fused", 1.0), ("fused", 0.5), ("fused", 0.0), ("reference", None)):
        full = W4A4Linear.from_float(lin, "per_group", "per_group", importance=imp, salient_prop=0.05, group_size=g,
                                     mode=mode, sort_key=imp, int_fraction=frac)
        perm = None
        if mode == "reference":                       # freeze the dynamic permutation so that every shard shares it
            from smoothquant_mixedprecision_b200 import fake_quant as fq
            _, keep = full._device_indices(x.device)
            perm = fq._dynamic_perm(x, subset=keep)
        y = full(x, act_perm=perm) if mode == "reference" else full(x)
        parts = []
        for r in range(world):
            sh = copy.deepcopy(full).shard_(r, world, gather_output=False)
            lo, hi = P.shard_range(N, world, r)
            assert sh.out_features == hi - lo and sh.full_out_features == N
            parts.append(sh(x, act_perm=perm) if mode == "reference" else sh(x))
        assert torch.equal(torch.cat(parts, dim=1), y), mode


def test_blocked_columns_validates_its_buffer():
    buf = torch.zeros(2, 3, 16)
    assert tuple(P.BlockedColumns(buf, 20).shape) == (3, 20)
    assert torch.equal(P.BlockedColumns(buf, 32).dense(), torch.zeros(3, 32))
    with pytest.raises(ValueError):
        P.BlockedColumns(buf, 16)                 # the second block would hold no column
    with pytest.raises(ValueError):
        P.BlockedColumns(buf, 33)
    with pytest.raises(ValueError):
        P.BlockedColumns(buf.permute(1, 0, 2), 20)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("world,K,g,p", [(2, 4096, 128, 0.05), (8, 11008, 128, 0.05), (4, 1000, 64, 0.1),
                                         (2, 2048, 1024, 0.0), (8, 8192, 32, 0.05)])
def test_row_quantiser_reads_the_gather_layout_in_place(world, K, g, p, dtype):
    """mq_quant_rows_blocked on the [world, M, S] buffer an all-gather leaves behind == mq_quant_rows on the dense
    tensor, bit for bit (codes, scales, salient copy), with and without a permutation."""
    from smoothquant_mixedprecision_b200 import ops
    torch.manual_seed(world * 1000 + K)
    M = 333
    S = P.shard_size(K, world)
    x = (torch.randn(M, K, device="cuda") * torch.exp(0.5 * torch.randn(K, device="cuda"))).to(dtype)
    buf = torch.full((world, M, S), 7.0, dtype=dtype, device="cuda")       # poison in the padding: must never be read
    for r in range(world):
        lo, hi = P.shard_range(K, world, r)
        buf[r][:, : hi - lo] = x[:, lo:hi]
    bc = P.BlockedColumns(buf, K)
    assert torch.equal(bc.dense(), x)
    ks = int(round(K * p))
    order = torch.randperm(K, device="cuda")
    sal = torch.sort(order[:ks]).values.int() if ks else None
    for perm in (order[ks:].int(), None):
        Kn = K - ks
        if perm is None:                                       # contiguous mode: identity over the first Kn columns
            sal_c = torch.arange(Kn, K, device="cuda").int() if ks else None
            want = ops.quant_rows(x, None, sal_c, g, 4, Kn=Kn)
            got = ops.quant_rows(bc, None, sal_c, g, 4, Kn=Kn)
        else:
            want = ops.quant_rows(x, perm, sal, g, 4)
            got = ops.quant_rows(bc, perm, sal, g, 4)
        assert torch.equal(got.codes[:, : want.groups * g], want.codes[:, : want.groups * g])
        assert torch.equal(got.scales[:, :M], want.scales[:, :M])
        if ks:
            assert torch.equal(got.sal[:, :ks], want.sal[:, :ks])


@pytest.mark.gpu
def test_blocked_input_layer_matches_dense_input_layer():
    """A W4A4Linear fed a BlockedColumns produces the same y as the same layer fed the dense tensor (bit for bit), for
    prefill and decode row counts; a layer K1's blocked kernel does not cover falls back to .dense()."""
    from smoothquant_mixedprecision_b200.fake_quant import W4A4Linear, clear_act_cache
    torch.manual_seed(11)
    K, N, g, world = 2048, 512, 128, 4
    lin = torch.nn.Linear(K, N).half().cuda()
    imp = torch.rand(K)
    S = P.shard_size(K, world)
    for M in (8, 700):
        x = torch.randn(M, K, device="cuda").half()
        buf = torch.zeros(world, M, S, dtype=torch.float16, device="cuda")
        for r in range(world):
            lo, hi = P.shard_range(K, world, r)
            buf[r][:, : hi - lo] = x[:, lo:hi]
        for frac in (1.0, 0.5):
            mod = W4A4Linear.from_float(lin, "per_group", "per_group", importance=imp, salient_prop=0.05,
                                        group_size=g, mode="fused", sort_key=imp, int_fraction=frac)
            clear_act_cache()
            y = mod(x).clone()
            clear_act_cache()
            assert torch.equal(mod(P.BlockedColumns(buf, K)), y), (M, frac)
        # an output-quantised layer (OPT q/k/v: N == K) behind a blocked input runs the output quantiser as usual
        sq = torch.nn.Linear(K, K).half().cuda()
        modq = W4A4Linear.from_float(sq, "per_group", "per_group", quantize_output=True, importance=imp,
                                     salient_prop=0.05, group_size=g, mode="fused", sort_key=imp, int_fraction=1.0)
        clear_act_cache()
        yq = modq(x).clone()
        clear_act_cache()
        assert torch.equal(modq(P.BlockedColumns(buf, K)), yq), M


def _nccl_pipeline_worker(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import bench
        bench.HIDDEN, bench.INTER = 1024, 2752
        H, I = bench.HIDDEN, bench.INTER
        bench.SHAPES = {"q_proj": (H, H), "k_proj": (H, H), "v_proj": (H, H), "o_proj": (H, H), "gate_proj": (I, H),
                        "up_proj": (I, H), "down_proj": (H, I)}
        M = 640
        full, x0 = bench.build_stack(2, M, dev, seed=5)
        want = bench.run_stack(full, x0).clone()
        tpst, x1 = bench.build_stack(2, M, dev, seed=5, tp=(rank, world))
        assert torch.equal(x0, x1)
        for transport, asynchronous in (("nccl", True), ("nccl", False), ("pull", True), ("pull", False),
                                        ("push", True), ("push", False)):
            if transport != "nccl":
                P.enable_peer_gather(ring=3, mode=transport)
            P.ASYNC_GATHER = asynchronous
            for chunks in (1, 2, 3, 2, 1):      # the repeats re-use ring buffers (overwrite-after-pulled ordering)
                rows = [(i * M // chunks, (i + 1) * M // chunks) for i in range(chunks)]
                outs = bench.run_stack_tp(tpst, [x0[a:b] for a, b in rows])
                assert all(isinstance(o, P.BlockedColumns) for o in outs)
                got = torch.cat([o.dense() for o in outs])
                assert torch.equal(got, want), (rank, transport, asynchronous, chunks,
                                                float((got - want).abs().max()))
        # decode row counts take the decode kernel on every shard
        for mods in tpst:
            for m in mods.values():
                m.prepare_decode()
        for mods in full:
            for m in mods.values():
                m.prepare_decode()
        xd = x0[:16].contiguous()
        want_d = bench.run_stack(full, xd)
        got_d = bench.run_stack_tp(tpst, [xd])[0].dense()
        assert torch.equal(got_d, want_d), (rank, "decode")
        torch.cuda.synchronize()
    finally:
        P.ASYNC_GATHER, P.TRANSPORT = True, "nccl"
        dist.destroy_process_group()


@pytest.mark.gpu
def test_nccl_world2_column_parallel_pipeline_is_bit_identical():
    """Two GPUs, NCCL: the column-parallel pass over row chunks (in-place async all-gathers into [p, M, S], K1 reading
    that layout) reproduces the single-GPU stack bit for bit.  Skipped on a one-GPU box."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    mp.spawn(_nccl_pipeline_worker, args=(2, _free_port()), nprocs=2, join=True)
