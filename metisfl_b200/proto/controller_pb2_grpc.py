"""gRPC bindings for ``metisfl.ControllerService`` (12 RPCs, reference ``controller.proto:8-49``)."""
from metisfl_b200.proto import schema as _schema
from metisfl_b200.proto._grpc_gen import make_service as _make

ControllerServiceStub, ControllerServiceServicer, add_ControllerServiceServicer_to_server = _make(
    _schema.CONTROLLER, "ControllerService")
