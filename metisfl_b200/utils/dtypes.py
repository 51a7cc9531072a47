"""dtype tables shared by the proto codec, the CPU oracle and the CUDA ops.

Wire numbering = ``metisfl.DType.Type`` (reference ``model.proto:16-27``) + the two extensions
FLOAT16 = 10 and BFLOAT16 = 11.
"""
import numpy as np
import torch

INT8, INT16, INT32, INT64, UINT8, UINT16, UINT32, UINT64, FLOAT32, FLOAT64, FLOAT16, BFLOAT16 = range(12)

WIRE_NAMES = ["INT8", "INT16", "INT32", "INT64", "UINT8", "UINT16", "UINT32", "UINT64",
              "FLOAT32", "FLOAT64", "FLOAT16", "BFLOAT16"]
WIRE_SIZE = [1, 2, 4, 8, 1, 2, 4, 8, 4, 8, 2, 2]

_TORCH = {
    torch.int8: INT8, torch.int16: INT16, torch.int32: INT32, torch.int64: INT64,
    torch.uint8: UINT8, torch.uint16: UINT16, torch.uint32: UINT32, torch.uint64: UINT64,
    torch.float32: FLOAT32, torch.float64: FLOAT64, torch.float16: FLOAT16, torch.bfloat16: BFLOAT16,
    torch.bool: UINT8,
}
_FROM_WIRE_TORCH = {v: k for k, v in _TORCH.items() if k is not torch.bool}

_NUMPY = {
    np.dtype("int8"): INT8, np.dtype("int16"): INT16, np.dtype("int32"): INT32, np.dtype("int64"): INT64,
    np.dtype("uint8"): UINT8, np.dtype("uint16"): UINT16, np.dtype("uint32"): UINT32,
    np.dtype("uint64"): UINT64, np.dtype("float32"): FLOAT32, np.dtype("float64"): FLOAT64,
    np.dtype("float16"): FLOAT16, np.dtype("bool"): UINT8,
}
_FROM_WIRE_NUMPY = {v: k for k, v in _NUMPY.items() if k != np.dtype("bool")}


def torch_to_wire(dt: torch.dtype) -> int:
    try:
        return _TORCH[dt]
    except KeyError:
        raise TypeError(f"torch dtype {dt} has no metisfl wire type") from None


def wire_to_torch(code: int) -> torch.dtype:
    return _FROM_WIRE_TORCH[int(code)]


def numpy_to_wire(dt) -> int:
    dt = np.dtype(dt).newbyteorder("=")
    try:
        return _NUMPY[dt]
    except KeyError:
        raise TypeError(f"numpy dtype {dt} has no metisfl wire type") from None


def wire_to_numpy(code: int):
    """numpy dtype for a wire code; BFLOAT16 has no numpy type and is carried as uint16 bits."""
    code = int(code)
    if code == BFLOAT16:
        return np.dtype("uint16")
    return _FROM_WIRE_NUMPY[code]
