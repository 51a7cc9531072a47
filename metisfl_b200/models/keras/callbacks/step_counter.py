"""Import-path parity with reference ``metisfl/models/keras/callbacks/step_counter.py``."""
from metisfl_b200.models.callbacks import StepCounter  # noqa: F401
