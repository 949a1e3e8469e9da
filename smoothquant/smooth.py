"""`smoothquant.smooth` import path -> B200-native package (pure weight transform, see its docstring)."""
from smoothquant_mixedprecision_b200.smooth import *  # noqa: F401,F403
from smoothquant_mixedprecision_b200.smooth import __all__  # noqa: F401
