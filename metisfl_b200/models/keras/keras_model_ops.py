"""Keras backend placeholder.

The reference's ``KerasModelOps`` (``models/keras/keras_model_ops.py:22-283``) needs TensorFlow <= 2.11,
which is not part of this environment; the B200 build standardises on the PyTorch backend
(``metisfl_b200.models.torch_model_ops.TorchModelOps``), which covers everything the Keras path did:
framework-owned training loop with a step budget (the ``StepCounter`` callback, without its
off-by-one, quirk Q11), per-batch / per-epoch wall-clock profiling (``PerformanceProfiler``), the five
``OptimizerConfig`` optimizers including FedProx, save / load of the model directory.
"""


class KerasModelOps(object):  # pragma: no cover - needs TensorFlow
    def __init__(self, *args, **kwargs):
        try:
            import tensorflow  # noqa: F401
        except ImportError as e:
            raise ImportError(
                "KerasModelOps needs TensorFlow, which is not installed here; use neural_engine='pytorch' "
                "(metisfl_b200.models.torch_model_ops.TorchModelOps).") from e
        raise NotImplementedError("The Keras backend is not part of the B200 build; use the PyTorch backend.")
