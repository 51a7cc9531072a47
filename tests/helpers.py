"""Shared test helpers: build serialized ``Model`` messages from numpy arrays."""
import numpy as np

from metisfl_b200.proto import model_pb2
from metisfl_b200.utils.proto_messages_factory import ModelProtoMessages


def model_bytes(arrays, names=None, trainable=True):
    names = names or [f"var{i}" for i in range(len(arrays))]
    m = ModelProtoMessages.construct_model_pb_from_np(list(arrays), names, [trainable] * len(arrays))
    return m.SerializeToString()


def model_arrays(blob):
    m = model_pb2.Model.FromString(blob)
    return [ModelProtoMessages.TensorSpecProto.proto_tensor_spec_to_numpy_array(v.plaintext_tensor.tensor_spec)
            for v in m.variables]


def one_to_ten(dtype):
    return np.arange(1, 11).astype(dtype)
