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
