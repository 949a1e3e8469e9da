"""ctypes binding of libmq.so -- the C-ABI boundary of include/mq.h.

There is no CPU fallback: if the library is missing the import fails loudly, and every compute entry point needs a
CUDA device.  Build it with ``python -m smoothquant_mixedprecision_b200.build`` (or ``__graft_entry__.build()``).
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# MQ_LIB: an alternative build of the same library (kernel-variant experiments, tools/build_variant.sh); no fallback
LIB_PATH = os.environ.get("MQ_LIB") or os.path.join(HERE, "libmq.so")

MQ_F16, MQ_BF16, MQ_F32 = 0, 1, 2
MQ_STAT_ABSMAX, MQ_STAT_ABSMEAN, MQ_STAT_MEAN3STD, MQ_STAT_ARGMAX = 0, 1, 2, 3

_i64, _i32, _vp, _sz = C.c_int64, C.c_int, C.c_void_p, C.c_size_t

# name -> (restype, argtypes); mirrors include/mq.h declaration by declaration
SIGNATURES = {
    "mq_version": (_i32, []),
    "mq_last_error": (C.c_char_p, []),
    "mq_device_info": (_i32, [C.POINTER(_i32)] * 3),
    "mq_colstat_workspace_bytes": (_sz, [_i64, _i32]),
    "mq_colstat": (_i32, [_vp, _i64, _i64, _i64, _i32, _i32, _vp, _vp, _sz, _vp]),
    "mq_argsort_workspace_bytes": (_sz, [_i64]),
    "mq_colstat_static_max_bytes": (_sz, []),
    "mq_colstat_static": (_i32, [_vp, _i64, _i64, _i64, _i32, _i32, _vp, _vp]),
    "mq_argsort_stable": (_i32, [_vp, _vp, _i64, _i32, _vp, _vp, _sz, _vp]),
    "mq_quant_rows": (_i32, [_vp, _i64, _i64, _i64, _i32, _vp, _vp, _i64, _vp, _i64, _i32, _i32,
                             _vp, _i64, _vp, _i64, _vp, _i64, _vp]),
    "mq_quant_rows_split": (_i32, [_vp, _i64, _i64, _i64, _i32, _vp, _vp, _i64, _vp, _i64, _i32, _i32, _i64,
                                   _vp, _i64, _vp, _i64, _vp, _i64, _vp]),
    "mq_quant_rows_blocked": (_i32, [_vp, _i64, _i64, _i32, _i64, _i64, _i64, _i32, _vp, _i64, _vp, _i64, _i32, _i32,
                                     _vp, _i64, _vp, _i64, _vp, _i64, _vp]),
    "mq_norm_quant_rows": (_i32, [_vp, _i64, _i64, _i64, _i32, _i32, _vp, _vp, C.c_float, _vp, _i64, _vp, _i64, _i32,
                                  _i32, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _vp]),
    "mq_fakequant_rows": (_i32, [_vp, _i64, _i64, _i64, _i32, _vp, _i64, _i32, _i32, _vp, _i64, _vp]),
    "mq_fakequant_tensor": (_i32, [_vp, _i64, _i32, _i32, _vp, _vp, _sz, _vp]),
    "mq_pack_int4": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _vp]),
    "mq_unpack_int4": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _vp]),
    "mq_mixed_gemm": (_i32, [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64,
                             _vp, _i64, _i64, _i64, _i32, _i64, _i32, _vp, _i64, _vp]),
    "mq_decode_weight_bytes": (_i32, [_i64, _i64, _i32, _i64, C.POINTER(_sz), C.POINTER(_sz)]),
    "mq_pack_decode": (_i32, [_vp, _i64, _vp, _i64, _vp, _i64, _i64, _i64, _i32, _i64, _i32, _vp, _vp, _vp]),
    "mq_decode_gemm": (_i32, [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _i64,
                              _i32, _vp, _i64, _vp]),
    "mq_peer_gather": (_i32, [_vp, _vp, _i32, _i32, _sz, C.c_uint32, _vp, _i32]),
    "mq_peer_push": (_i32, [_vp, _vp, C.c_uint64, _i32, _i32, _sz, C.c_uint32, C.c_uint32, _i32, _vp, _i32]),
    "mq_peer_wait_arrived": (_i32, [_vp, _i32, _i32, C.c_uint32, _vp]),
    "mq_peer_publish_consumed": (_i32, [_vp, _i32, _i32, C.c_uint32, _i32, _vp]),
    "mq_peer_wait_pulled": (_i32, [_vp, _i32, _i32, C.c_uint32, _vp]),
    "mq_set_option": (_i32, [C.c_char_p, _i32]),
    "mq_quant_fault": (_i32, [C.POINTER(C.c_uint)]),
    "mq_gemm_fault": (_i32, [C.POINTER(C.c_uint)]),
    "mq_gemm_profile": (_i32, [C.POINTER(C.c_ulonglong), _i32]),
}


class MqError(RuntimeError):
    pass


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the CUDA library is not built and there is no CPU fallback. "
            "Run `python -m smoothquant_mixedprecision_b200.build`.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header / library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib.mq_last_error().decode("utf-8", "replace")
        raise MqError(f"{what or 'libmq'} failed with status {rc}: {msg}")


def set_option(name: str, value: int) -> None:
    check(lib.mq_set_option(name.encode(), int(value)), "mq_set_option")


def gemm_fault():
    buf = (C.c_uint * 8)()
    check(lib.mq_gemm_fault(buf), "mq_gemm_fault")
    return list(buf)


def quant_fault():
    buf = (C.c_uint * 8)()
    check(lib.mq_quant_fault(buf), "mq_quant_fault")
    return list(buf)


def gemm_profile(n_ctas: int = 148):
    """Per-CTA wait-cycle records of the last mixed GEMM launched with set_option("gemm_profile", 1)."""
    buf = (C.c_ulonglong * (8 * n_ctas))()
    check(lib.mq_gemm_profile(buf, n_ctas), "mq_gemm_profile")
    return [list(buf[i * 8:(i + 1) * 8]) for i in range(n_ctas)]
