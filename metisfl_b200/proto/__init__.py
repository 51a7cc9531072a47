"""Wire contract: protobuf messages and gRPC service bindings (see ``schema.py``)."""
