"""mq_colstat_static: the no-allocation variant of the column-statistics kernel (scratch in a static __device__ array
inside the library, include/mq.h), called through the C ABI and compared with the existing mq_colstat entry point.

Collected last on purpose (the file name sorts after every other test file)."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
MQ_ERR_WORKSPACE = -4          # include/mq.h


def _call_both(x, kind):
    """(existing entry point with a caller workspace, static-scratch entry point) on the same input."""
    from smoothquant_mixedprecision_b200 import _lib, ops
    lib = _lib.lib
    M, K = x.shape
    want = torch.full((K,), -7.0, dtype=torch.float32, device=DEV)
    got = torch.full((K,), -9.0, dtype=torch.float32, device=DEV)
    ws = torch.empty(max(lib.mq_colstat_workspace_bytes(K, kind), 16), dtype=torch.uint8, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    with torch.cuda.device(x.device):
        rc = lib.mq_colstat(x.data_ptr(), M, K, x.stride(0), ops._DTYPES[x.dtype], kind, want.data_ptr(), ws.data_ptr(),
                            ws.numel(), st)
        assert rc == 0, lib.mq_last_error()
        rc = lib.mq_colstat_static(x.data_ptr(), M, K, x.stride(0), ops._DTYPES[x.dtype], kind, got.data_ptr(), st)
        assert rc == 0, lib.mq_last_error()
    torch.cuda.synchronize()
    return want, got


def _compare(want, got, kind, tag):
    from smoothquant_mixedprecision_b200 import _lib
    if kind in (_lib.MQ_STAT_ABSMAX, _lib.MQ_STAT_ARGMAX):
        assert torch.equal(got, want), tag                       # order-independent: identical
    else:
        # double accumulation per split, then one rounding to float: a different split count moves the sum by
        # ~1e-16 relative, i.e. at most the last float bit
        assert torch.allclose(got, want, rtol=1e-6, atol=0.0), (tag, float((got - want).abs().max()))


def test_static_scratch_colstat_matches_workspace_colstat_on_consecutive_calls():
    from smoothquant_mixedprecision_b200 import _lib
    lib = _lib.lib
    assert lib.mq_colstat_static_max_bytes() == 4 << 20
    g = torch.Generator(device=DEV).manual_seed(5)
    kinds = (_lib.MQ_STAT_ABSMAX, _lib.MQ_STAT_ABSMEAN, _lib.MQ_STAT_MEAN3STD, _lib.MQ_STAT_ARGMAX)
    # call 1 fills the scratch (64 splits x 4096 columns would need 6 MiB: the split count is reduced to fit), call 2
    # is smaller and runs over the leftovers of call 1, call 3 repeats call 1's shape on other data
    shapes = [(4096, 4096, torch.float16), (100, 1000, torch.bfloat16), (4096, 4096, torch.float16),
              (7, 33, torch.float32)]
    for i, (M, K, dt) in enumerate(shapes):
        x = (torch.randn(M, K, device=DEV, generator=g) * (1 + i)).to(dt)
        for kind in kinds:
            want, got = _call_both(x, kind)
            _compare(want, got, kind, (i, M, K, str(dt), kind))
    # against plain torch as well (absmax is exact)
    x = torch.randn(513, 777, device=DEV, generator=g).half()
    want, got = _call_both(x, _lib.MQ_STAT_ABSMAX)
    assert torch.equal(got, x.abs().amax(dim=0).float())


def test_static_scratch_colstat_allocates_nothing_and_rejects_oversize():
    from smoothquant_mixedprecision_b200 import _lib, ops
    lib = _lib.lib
    x = torch.randn(2048, 2048, device=DEV).half()
    out = torch.empty(2048, dtype=torch.float32, device=DEV)
    st = torch.cuda.current_stream().cuda_stream

    def call():
        with torch.cuda.device(x.device):
            return lib.mq_colstat_static(x.data_ptr(), 2048, 2048, 2048, ops._DTYPES[x.dtype], _lib.MQ_STAT_ABSMAX,
                                         out.data_ptr(), st)

    assert call() == 0, lib.mq_last_error()          # set-up call: resolves the scratch address for this device
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(5):
        assert call() == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0     # no device memory was taken by the calls
    assert torch.equal(out, x.abs().amax(dim=0).float())
    # beyond the maximum: K * 24 bytes > 4 MiB
    K = (4 << 20) // 24 + 1
    wide = torch.zeros(1, K, device=DEV, dtype=torch.float16)
    wout = torch.empty(K, dtype=torch.float32, device=DEV)
    with torch.cuda.device(x.device):
        rc = lib.mq_colstat_static(wide.data_ptr(), 1, K, K, ops._DTYPES[wide.dtype], _lib.MQ_STAT_ABSMAX,
                                   wout.data_ptr(), st)
    assert rc == MQ_ERR_WORKSPACE
    msg = lib.mq_last_error().decode()
    assert "static scratch" in msg and str(K) in msg
    # the largest width that fits still runs
    K = (4 << 20) // 24
    wide = torch.randn(3, K, device=DEV).half()
    want, got = _call_both(wide, _lib.MQ_STAT_ABSMAX)
    assert torch.equal(got, want)
