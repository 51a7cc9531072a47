"""Requests that carry a whole model, assembled and taken apart at the byte level.

In the reference every hop of a model is a protobuf object graph: the controller parses its serialized community
model, copies it into a ``RunTaskRequest`` and serializes that once per learner (``controller.cc:696-759``); a learner
wraps its model in ``CompletedLearningTask``, wraps that in ``MarkTaskCompletedRequest`` and serializes it
(``learner.py:197-206``); the controller parses the request and serializes the model again for its store
(``controller_servicer.cc:316-342``).  Each of those steps is a pass over every tensor byte.

A protobuf message on the wire is a concatenation of ``(tag, value)`` records in ANY order, and an embedded message is
just a length-delimited record.  So a request around an already-serialized model is ``serialized small fields + tag +
length + model bytes`` - one concatenation, no parse, no object copy - and the receiver finds the model record by
walking the top-level tags.  The gRPC layer passes such ``bytes`` through untouched (``_grpc_gen.py``).  What comes out
is bit-compatible with what protobuf would have produced up to field order, which parsers do not care about.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

_LEN = 2  # wire type of length-delimited records (strings, bytes, embedded messages)
MAX_TOP_LEVEL_RECORDS = 4096


def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _number(message_cls, field_name: str) -> int:
    return message_cls.DESCRIPTOR.fields_by_name[field_name].number


def record_header(field_number: int, payload_len: int) -> bytes:
    return _varint((field_number << 3) | _LEN) + _varint(payload_len)


def embed(message_cls, field_name: str, payload, rest=b"") -> bytes:
    """Serialized ``message_cls`` whose embedded-message (or bytes) field ``field_name`` is the already serialized
    ``payload`` and whose other fields are the serialized ``rest`` (a message of the same class with only those set)."""
    if not isinstance(rest, (bytes, bytearray, memoryview)):
        rest = rest.SerializeToString()
    return b"".join((rest, record_header(_number(message_cls, field_name), len(payload)), payload))


def split(buf) -> Dict[int, List[Tuple[int, object]]]:
    """Top-level records of a serialized message: ``{field number: [(wire type, value), ...]}``; length-delimited values
    are ``memoryview`` slices of ``buf`` (no copy), varints are ints, fixed-width values are memoryviews too.  Raises
    ``ValueError`` on truncated or malformed input."""
    mv = memoryview(buf)
    n, i = len(mv), 0
    out: Dict[int, List[Tuple[int, object]]] = {}

    def varint(i):
        shift = val = 0
        while True:
            if i >= n or shift > 63:
                raise ValueError("truncated or oversized varint")
            b = mv[i]
            i += 1
            val |= (b & 0x7F) << shift
            if not b & 0x80:
                return val, i
            shift += 7

    records = 0
    while i < n:
        records += 1
        if records > MAX_TOP_LEVEL_RECORDS:   # the requests taken apart here have a handful; this walk is Python
            raise ValueError("too many top-level records")
        key, i = varint(i)
        field, wt = key >> 3, key & 7
        if field == 0:
            raise ValueError("field number 0")
        if wt == 0:
            val, i = varint(i)
        elif wt == _LEN:
            ln, i = varint(i)
            if ln > n - i:
                raise ValueError("length-delimited record runs past the end of the message")
            val, i = mv[i:i + ln], i + ln
        elif wt in (1, 5):
            w = 8 if wt == 1 else 4
            if w > n - i:
                raise ValueError("fixed-width record runs past the end of the message")
            val, i = mv[i:i + w], i + w
        else:
            raise ValueError("unsupported wire type {}".format(wt))
        out.setdefault(field, []).append((wt, val))
    return out


def last_bytes(records: Dict[int, List[Tuple[int, object]]], message_cls, field_name: str):
    """The (last, as protobuf merges scalars) length-delimited value of ``field_name``, or None."""
    vals = [v for wt, v in records.get(_number(message_cls, field_name), ()) if wt == _LEN]
    return vals[-1] if vals else None


def merged_bytes(records, message_cls, field_name: str):
    """Embedded messages may legally arrive split over several records (protobuf merges them): one record is returned
    as is, several are concatenated - concatenation of serialized messages IS their merge."""
    vals = [v for wt, v in records.get(_number(message_cls, field_name), ()) if wt == _LEN]
    if not vals:
        return None
    return vals[0] if len(vals) == 1 else memoryview(b"".join(vals))


class WireCompletedTask(object):
    """A ``CompletedLearningTask`` whose model is already serialized (``construct_model_bytes_from_np``): what a learner
    hands to ``GRPCControllerClient.mark_task_completed`` so that the model is never copied into a message object."""

    def __init__(self, model_bytes: bytes, execution_metadata_pb=None, aux_metadata: str = ""):
        self.model_bytes = model_bytes
        self.execution_metadata = execution_metadata_pb
        self.aux_metadata = aux_metadata

    def _rest(self):
        from metisfl_b200.proto import metis_pb2
        rest = metis_pb2.CompletedLearningTask(aux_metadata=self.aux_metadata or "")
        if self.execution_metadata is not None:
            rest.execution_metadata.CopyFrom(self.execution_metadata)
        return rest.SerializeToString()

    def parts(self):
        """Serialized task as a list of byte strings (joined by the caller together with its own envelope)."""
        from metisfl_b200.proto import metis_pb2
        return [self._rest(), record_header(_number(metis_pb2.CompletedLearningTask, "model"), len(self.model_bytes)),
                self.model_bytes]

    def SerializeToString(self) -> bytes:
        return b"".join(self.parts())

    def to_pb(self):
        from metisfl_b200.proto import metis_pb2
        return metis_pb2.CompletedLearningTask.FromString(self.SerializeToString())

    @property
    def model(self):
        """The model as a message (parsed on demand; for callers that inspect a task instead of sending it)."""
        from metisfl_b200.proto import model_pb2
        return model_pb2.Model.FromString(self.model_bytes)


def mark_task_completed_request(learner_id: str, auth_token: str, task) -> bytes:
    """Serialized ``MarkTaskCompletedRequest`` around ``task`` (a ``WireCompletedTask`` or a ``CompletedLearningTask``
    message): one concatenation."""
    from metisfl_b200.proto import controller_pb2
    R = controller_pb2.MarkTaskCompletedRequest
    rest = R(learner_id=learner_id, auth_token=auth_token).SerializeToString()
    parts = task.parts() if isinstance(task, WireCompletedTask) else [task.SerializeToString()]
    return b"".join([rest, record_header(_number(R, "task"), sum(len(p) for p in parts))] + parts)
