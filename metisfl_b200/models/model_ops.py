"""Abstract model backend (API parity with reference ``metisfl/models/model_ops.py:18-144``)."""
import abc
from typing import List

import numpy as np

from metisfl_b200.proto import metis_pb2, model_pb2
from metisfl_b200.utils.proto_messages_factory import ModelProtoMessages


class ModelOps(object):

    def __init__(self, model=None, he_scheme=None, *args, **kwargs):
        self._model = model
        self._he_scheme = he_scheme

    def get_model_weights_from_variables_pb(self, variables: List[model_pb2.Model.Variable]):
        """``Model.variables`` -> (names, trainables, values).  Ciphertext variables are decrypted with
        the learner's private key; plaintext ones are decoded in place (torch tensors, so that
        bfloat16 survives)."""
        assert all(isinstance(v, model_pb2.Model.Variable) for v in variables)
        names, trainables, values = [], [], []
        codec = ModelProtoMessages.TensorSpecProto
        for var in variables:
            names.append(var.name)
            trainables.append(var.trainable)
            if var.HasField("ciphertext_tensor"):
                assert self._he_scheme is not None, "Need encryption scheme to decrypt tensor."
                spec = var.ciphertext_tensor.tensor_spec
                flat = self._he_scheme.decrypt(spec.value, spec.length)
                values.append(codec.proto_tensor_spec_with_list_values_to_numpy_array_like(spec, flat))
            elif var.HasField("plaintext_tensor"):
                values.append(codec.proto_tensor_spec_to_torch_tensor(var.plaintext_tensor.tensor_spec))
            else:
                raise RuntimeError("Not a supported tensor type.")
        return names, trainables, values

    def cleanup(self):
        pass

    @abc.abstractmethod
    def load_model(self, model_dir=None, *args, **kwargs):
        pass

    @abc.abstractmethod
    def save_model(self, model_dir=None, *args, **kwargs):
        pass

    @abc.abstractmethod
    def set_model_weights(self, weights_names: List[str], weights_trainable: List[bool], weights_values, *args,
                          **kwargs):
        pass

    @abc.abstractmethod
    def get_model_weights(self, *args, **kwargs):
        """-> (names, trainables, values) of the FULL model state: parameters and buffers."""
        pass

    @abc.abstractmethod
    def train_model(self, train_dataset, learning_task_pb: metis_pb2.LearningTask,
                    hyperparameters_pb: metis_pb2.Hyperparameters, validation_dataset=None, test_dataset=None,
                    verbose=False, *args, **kwargs) -> metis_pb2.CompletedLearningTask:
        pass

    @abc.abstractmethod
    def evaluate_model(self, eval_dataset, batch_size=100, metrics=None, verbose=False, *args, **kwargs) -> dict:
        pass

    @abc.abstractmethod
    def infer_model(self, infer_dataset, batch_size=100, *args, **kwargs):
        pass

    @abc.abstractmethod
    def construct_optimizer(self, optimizer_config_pb: model_pb2.OptimizerConfig = None, *args, **kwargs):
        pass
