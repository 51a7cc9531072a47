"""``model_pb2`` — message classes built at import time from ``schema.py`` (no protoc)."""
from metisfl_b200.proto import schema as _schema

_schema.export_messages(_schema.MODEL, globals())
