"""`smoothquant.model_size` import path (smoothquant/model_size.py:5, used by run_experiments.py:22, :178)."""
from smoothquant_mixedprecision_b200.model_size import get_model_size

__all__ = ["get_model_size"]
