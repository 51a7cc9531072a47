"""Small shared helpers: wire dtype tables, metric dictionary formatting, TLS configuration plumbing."""
import math
import os

import numpy as np
import pytest
import torch

from metisfl_b200.proto import metis_pb2
from metisfl_b200.utils import dtypes
from metisfl_b200.utils.formatting import DictionaryFormatter
from metisfl_b200.utils.ssl_configurator import SSLConfigurator


def test_wire_dtype_tables_round_trip():
    assert dtypes.WIRE_NAMES[dtypes.FLOAT32] == "FLOAT32" and dtypes.WIRE_SIZE[dtypes.BFLOAT16] == 2
    for code, name in enumerate(dtypes.WIRE_NAMES):
        t = dtypes.wire_to_torch(code)
        assert dtypes.torch_to_wire(t) == code and t.itemsize == dtypes.WIRE_SIZE[code], name
        n = dtypes.wire_to_numpy(code)
        assert n.itemsize == dtypes.WIRE_SIZE[code]
        if code != dtypes.BFLOAT16:                      # bf16 travels as raw uint16 bits on the numpy side
            assert dtypes.numpy_to_wire(n) == code
    assert dtypes.wire_to_numpy(dtypes.BFLOAT16) == np.dtype("uint16")
    assert dtypes.torch_to_wire(torch.bool) == dtypes.UINT8 and dtypes.numpy_to_wire(np.bool_) == dtypes.UINT8
    assert dtypes.numpy_to_wire(np.dtype(">f4")) == dtypes.FLOAT32   # byte order is carried separately on the wire
    with pytest.raises(TypeError):
        dtypes.torch_to_wire(torch.complex64)
    with pytest.raises(TypeError):
        dtypes.numpy_to_wire(np.complex64)


def test_metric_dictionaries_travel_as_strings():
    d = {"loss": [0.25], "accuracy": np.float32(0.5), "auc": float("nan"), "hist": np.array([1, 2]), 3: np.int64(7)}
    norm = DictionaryFormatter.normalize(d)
    assert norm["loss"] == 0.25 and norm["hist"] == [1, 2]
    s = DictionaryFormatter.stringify(norm)
    assert s["loss"] == "0.25" and s["accuracy"] == "0.5" and s["auc"] == "NaN" and s["3"] == "7"
    assert DictionaryFormatter.stringify({"x": float("inf")}, stringify_nan=False)["x"] == "inf"
    assert DictionaryFormatter.listify_values({"a": 1, "b": [2]}) == {"a": [1], "b": [2]}
    assert math.isnan(float(s["auc"]))


def test_default_certificates_and_ssl_config_plumbing(tmp_path):
    cert, key = SSLConfigurator.gen_default_certificates(as_stream=False)
    assert os.path.exists(cert) and os.path.exists(key)
    cert_s, key_s = SSLConfigurator.gen_default_certificates(as_stream=True)
    assert b"BEGIN CERTIFICATE" in cert_s and b"PRIVATE KEY" in key_s
    assert SSLConfigurator.load_file_as_stream(cert) == cert_s
    # a config with files -> the same material as streams; only the PUBLIC half is handed to peers
    cfg = metis_pb2.SSLConfig(enable_ssl=True)
    cfg.ssl_config_files.public_certificate_file = cert
    cfg.ssl_config_files.private_key_file = key
    pub, prv = SSLConfigurator.load_certificates_from_ssl_config_pb(cfg, as_stream=True)
    assert pub == cert_s and prv == key_s
    public_only = SSLConfigurator.gen_public_ssl_config_pb_as_stream(cfg)
    assert public_only.enable_ssl and public_only.ssl_config_stream.public_certificate_stream == cert_s
    assert not public_only.ssl_config_stream.private_key_stream
    off = metis_pb2.SSLConfig(enable_ssl=False)
    assert SSLConfigurator.load_certificates_from_ssl_config_pb(off, as_stream=True) == (None, None)
