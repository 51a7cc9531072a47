"""Model definition hooks (API parity with reference ``metisfl/models/model_def.py:8-23``).

In the reference a PyTorch model must implement its own ``fit`` / ``evaluate`` (``PyTorchDef``);
here the framework owns the training loop (``TorchModelOps``), so a plain ``nn.Module`` is enough.
``PyTorchDef`` stays available: a module deriving from it may override ``loss_fn`` / ``metrics_fn``
and, for full control, ``fit`` / ``evaluate`` (used instead of the built-in loop when defined)."""
import abc
from typing import Dict


class ModelDef:

    @abc.abstractmethod
    def get_model(self, *args, **kwargs):
        pass


class PyTorchDef(ModelDef):

    def get_model(self, *args, **kwargs):
        return self

    # optional overrides ------------------------------------------------------------------------
    def loss_fn(self, outputs, targets):  # pragma: no cover - default provided by TorchModelOps
        raise NotImplementedError

    def fit(self, dataset, epochs: int, *args, **kwargs) -> Dict:  # pragma: no cover
        raise NotImplementedError

    def evaluate(self, dataset, *args, **kwargs) -> Dict:  # pragma: no cover
        raise NotImplementedError
