"""Drop-in import path of the reference package (smoothquant/__init__.py:1-4): `from smoothquant import smooth_lm,
quantize_model` resolves to the B200-native implementation in smoothquant_mixedprecision_b200."""
from smoothquant_mixedprecision_b200.smooth import smooth_lm


def __getattr__(name):
    if name == "quantize_model":
        from smoothquant_mixedprecision_b200.fake_quant import quantize_model
        return quantize_model
    raise AttributeError(name)


__all__ = ["smooth_lm", "quantize_model"]
