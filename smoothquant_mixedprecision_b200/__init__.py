"""B200-native W4A4 MixedQuant linear path.  Public surface = the reference's (smoothquant/__init__.py:1-4)."""
from .smooth import smooth_lm


def __getattr__(name):
    # fake_quant pulls in the CUDA library (libmq.so) through ops/_lib; resolve it lazily so that host-only
    # utilities (smooth_lm) stay importable, while any use of the quantised path fails loudly without the library
    if name in ("quantize_model", "quantize_opt", "quantize_llama_like", "W4A4Linear"):
        from . import fake_quant
        return getattr(fake_quant, name)
    raise AttributeError(name)


__all__ = ["smooth_lm", "quantize_model"]
