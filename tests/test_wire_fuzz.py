"""Property tests of the hand-written C++ protobuf codec (`csrc/core/wire.cc`): it parses bytes that come from
learners, i.e. untrusted input.  (1) Whatever protobuf-python serialises, the C++ side reads back identically and
re-serialises to bytes protobuf-python parses to the same message; (2) arbitrary and mutated byte strings are either
parsed or rejected with an exception - never a crash, never a read past the buffer (run under ASan in tools/sanitize.sh
style builds to make the second half bite)."""
import numpy as np
from hypothesis import given, settings, strategies as st

from metisfl_b200 import _core
from metisfl_b200.proto import model_pb2

_DT = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11]
_SIZE = [1, 2, 4, 8, 1, 2, 4, 8, 4, 8, 2, 2]

variable = st.fixed_dictionaries({
    "name": st.text(min_size=0, max_size=24),
    "trainable": st.booleans(),
    "dtype": st.sampled_from(_DT),
    "dims": st.lists(st.integers(min_value=0, max_value=5), min_size=0, max_size=3),
    "slot": st.one_of(st.none(), st.tuples(st.integers(0, 15), st.integers(0, 2 ** 32 - 1),
                                           st.integers(0, 2 ** 40), st.integers(0, 2 ** 50))),
    "seed": st.integers(0, 2 ** 16),
})


def _build(vars_):
    m = model_pb2.Model()
    for v in vars_:
        n = int(np.prod(v["dims"])) if v["dims"] else 1
        var = m.variables.add(name=v["name"], trainable=v["trainable"])
        spec = var.plaintext_tensor.tensor_spec
        spec.length = n
        spec.dimensions.extend(v["dims"])
        spec.type.type = v["dtype"]
        spec.type.byte_order = 1
        if v["slot"] is None:
            spec.value = np.random.default_rng(v["seed"]).integers(0, 256, size=n * _SIZE[v["dtype"]], dtype=np.uint8).tobytes()
        else:
            r, s, o, ver = v["slot"]
            spec.device_slot.device_rank, spec.device_slot.slot = r, s
            spec.device_slot.offset_bytes, spec.device_slot.version = o, ver
    return m


@settings(max_examples=150, deadline=None)
@given(st.lists(variable, min_size=0, max_size=5))
def test_cxx_codec_round_trips_what_protobuf_python_writes(vars_):
    m = _build(vars_)
    blob = m.SerializeToString()
    parsed = _core.parse_model(blob)
    assert len(parsed) == len(m.variables)
    for got, want in zip(parsed, m.variables):
        spec = want.plaintext_tensor.tensor_spec
        assert got["name"] == want.name and got["trainable"] == want.trainable and not got["ciphertext"]
        assert got["length"] == spec.length and list(got["dimensions"]) == list(spec.dimensions)
        assert got["dtype"] == spec.type.type and got["value"] == spec.value
        if spec.HasField("device_slot"):
            assert got["device_slot"] == (spec.device_slot.device_rank, spec.device_slot.slot,
                                          spec.device_slot.offset_bytes, spec.device_slot.version)
        else:
            assert "device_slot" not in got
    again = model_pb2.Model.FromString(_core.build_model(parsed))
    assert again == m


@settings(max_examples=300, deadline=None)
@given(st.binary(min_size=0, max_size=200))
def test_arbitrary_bytes_parse_or_raise(blob):
    for fn in (_core.parse_model, _core.parse_federated_model):
        try:
            fn(blob)
        except RuntimeError:
            pass


@settings(max_examples=200, deadline=None)
@given(st.lists(variable, min_size=1, max_size=3), st.data())
def test_mutated_models_parse_or_raise(vars_, data):
    blob = bytearray(_build(vars_).SerializeToString())
    if not blob:
        return
    for _ in range(data.draw(st.integers(1, 4))):
        kind = data.draw(st.sampled_from(["flip", "truncate", "insert"]))
        if kind == "flip" and blob:
            i = data.draw(st.integers(0, len(blob) - 1))
            blob[i] = data.draw(st.integers(0, 255))
        elif kind == "truncate" and blob:
            del blob[data.draw(st.integers(0, len(blob) - 1)):]
        else:
            i = data.draw(st.integers(0, len(blob)))
            blob[i:i] = data.draw(st.binary(min_size=1, max_size=6))
    try:
        for v in _core.parse_model(bytes(blob)):
            assert len(v["value"]) <= len(blob)      # a view into the input, never beyond it
    except RuntimeError:
        pass
    fm = _core.build_federated_model(3, 7, bytes(blob))
    c, it, inner = _core.parse_federated_model(fm)
    assert (c, it, inner) == (3, 7, bytes(blob))


# ---- ciphertext blobs (the PWA rule hands learner-supplied ciphertexts to the host scheme) ------------------------
import struct

import pytest


@pytest.fixture(scope="module")
def ciphertext(tmp_path_factory):
    from metisfl_b200.encryption import fhe
    s = fhe.CKKS(4096, 52, backend="host")
    s.gen_crypto_context_and_keys(str(tmp_path_factory.mktemp("crypto")), seed=5)
    return s.encrypt(np.linspace(-1, 1, 100))


def _patched(blob, offset, fmt, value):
    b = bytearray(blob)
    struct.pack_into(fmt, b, offset, value)
    return bytes(b)


def test_hostile_ciphertext_headers_are_rejected(ciphertext):
    """CtHeader (64 bytes): magic, version u16 @4, log_n u8 @6, n_limbs u8 @7, n_ct u32 @8, slots u32 @12,
    n_values u64 @16, scale f64 @24, q[4] u64 @32.  Every field the scheme later divides by, shifts by or sizes a
    buffer with must be refused when it is out of range - with an exception, not a crash."""
    ok = _core.ckks_header(ciphertext)
    assert ok["n_ct"] == 1 and ok["log_n"] == 14
    _core.ckks_weighted_average([ciphertext, ciphertext], [0.5, 0.5])           # the well-formed case works
    hostile = [
        _patched(ciphertext, 32, "<Q", 0),                  # modulus 0: division by zero in the weight encoding
        _patched(ciphertext, 32, "<Q", 4),                  # even modulus
        _patched(ciphertext, 40, "<Q", 1 << 63),            # modulus beyond the 62-bit arithmetic
        _patched(ciphertext, 6, "<B", 63),                  # ring dimension 2^63
        _patched(ciphertext, 6, "<B", 0),
        _patched(ciphertext, 7, "<B", 0),                   # no limbs
        _patched(ciphertext, 7, "<B", 9),                   # more limbs than the header has room for
        _patched(ciphertext, 8, "<I", 2),                   # claims two ciphertexts, carries one
        _patched(ciphertext, 12, "<I", 0),                  # zero slots
        _patched(ciphertext, 12, "<I", 1 << 20),            # more slots than the ring has
        _patched(ciphertext, 24, "<d", float("nan")),       # scale
        _patched(ciphertext, 24, "<d", -1.0),
        ciphertext[:-8], ciphertext + b"\\0" * 8, ciphertext[:40], b"",
    ]
    for blob in hostile:
        for call in (lambda b: _core.ckks_header(b), lambda b: _core.ckks_rescale(b),
                     lambda b: _core.ckks_weighted_average([b, b], [0.5, 0.5]),
                     lambda b: _core.ckks_weighted_average([ciphertext, b], [0.5, 0.5])):
            with pytest.raises((RuntimeError, ValueError)):
                call(blob)


@settings(max_examples=60, deadline=None)
@given(st.data())
def test_mutated_ciphertext_headers_parse_or_raise(ciphertext, data):
    b = bytearray(ciphertext)
    for _ in range(data.draw(st.integers(1, 3))):
        b[data.draw(st.integers(0, 63))] = data.draw(st.integers(0, 255))
    try:
        _core.ckks_weighted_average([bytes(b), bytes(b)], [0.25, 0.75])
    except (RuntimeError, ValueError):
        pass
