"""`smoothquant.fake_quant` import path -> B200-native implementation (no CPU fallback)."""
from smoothquant_mixedprecision_b200.fake_quant import *  # noqa: F401,F403
from smoothquant_mixedprecision_b200.fake_quant import __all__  # noqa: F401
