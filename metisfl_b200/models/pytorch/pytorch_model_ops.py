"""Import-path parity with the reference (``metisfl/models/pytorch/pytorch_model_ops.py``)."""
from metisfl_b200.models.torch_model_ops import PyTorchModelOps, TorchModelOps  # noqa: F401
