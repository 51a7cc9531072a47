"""Optimizers for user-owned loops; the framework-owned loop uses the fused kernel in ``metisfl_b200.ops.optim``."""
from metisfl_b200.optim.fed_prox import FedProx  # noqa: F401
