"""Import-path parity with reference ``metisfl/models/keras/callbacks/performance_profiler.py``."""
from metisfl_b200.models.callbacks import PerformanceProfiler  # noqa: F401
