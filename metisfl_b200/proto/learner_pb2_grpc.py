"""gRPC bindings for ``metisfl.LearnerService`` (4 RPCs, reference ``learner.proto:8-23``)."""
from metisfl_b200.proto import schema as _schema
from metisfl_b200.proto._grpc_gen import make_service as _make

LearnerServiceStub, LearnerServiceServicer, add_LearnerServiceServicer_to_server = _make(
    _schema.LEARNER, "LearnerService")
