"""The standard gRPC health-checking and server-reflection services on the controller's server (the reference enables
both, ``controller_servicer.cc:80-84``), called the way ``grpc_health_probe`` / ``grpcurl`` call them: by method path
with the public message layouts."""
import socket

import grpc
from google.protobuf import descriptor_pb2

from metisfl_b200.controller.service import ControllerServicer
from metisfl_b200.proto import metis_pb2
from metisfl_b200.proto.grpc_std_services import (HealthCheckRequest, HealthCheckResponse, ServerReflectionRequest,
                                                  ServerReflectionResponse)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _controller(port):
    p = metis_pb2.ControllerParams()
    p.server_entity.hostname, p.server_entity.port = "127.0.0.1", port
    p.global_model_specs.aggregation_rule.fed_avg.SetInParent()
    p.communication_specs.protocol = metis_pb2.CommunicationSpecs.SYNCHRONOUS
    p.model_store_config.in_memory_store.SetInParent()
    p.model_hyperparams.batch_size, p.model_hyperparams.epochs = 32, 1
    c = ControllerServicer(p)
    c.start()
    return c


def test_health_check_and_reflection_answer_like_the_standard_services():
    port = _free_port()
    ctrl = _controller(port)
    try:
        ch = grpc.insecure_channel(f"127.0.0.1:{port}")
        check = ch.unary_unary("/grpc.health.v1.Health/Check", request_serializer=HealthCheckRequest.SerializeToString,
                               response_deserializer=HealthCheckResponse.FromString)
        assert check(HealthCheckRequest(service=""), timeout=5).status == 1                           # SERVING
        assert check(HealthCheckRequest(service="metisfl.ControllerService"), timeout=5).status == 1
        try:
            check(HealthCheckRequest(service="no.such.Service"), timeout=5)
            raise AssertionError("unknown service must be NOT_FOUND")
        except grpc.RpcError as e:
            assert e.code() == grpc.StatusCode.NOT_FOUND
        # the raw wire form grpc_health_probe sends: an empty request is zero bytes, SERVING is 08 01
        raw = ch.unary_unary("/grpc.health.v1.Health/Check", request_serializer=lambda b: b,
                             response_deserializer=lambda b: b)
        assert raw(b"", timeout=5) == b"\x08\x01"

        info = ch.stream_stream("/grpc.reflection.v1alpha.ServerReflection/ServerReflectionInfo",
                                request_serializer=ServerReflectionRequest.SerializeToString,
                                response_deserializer=ServerReflectionResponse.FromString)
        reqs = [ServerReflectionRequest(list_services=""),
                ServerReflectionRequest(file_containing_symbol="metisfl.ControllerService"),
                ServerReflectionRequest(file_by_filename="no/such/file.proto")]
        answers = list(info(iter(reqs), timeout=10))
        names = {s.name for s in answers[0].list_services_response.service}
        assert {"metisfl.ControllerService", "grpc.health.v1.Health",
                "grpc.reflection.v1alpha.ServerReflection"} <= names
        files = [descriptor_pb2.FileDescriptorProto.FromString(b)
                 for b in answers[1].file_descriptor_response.file_descriptor_proto]
        assert files[0].package == "metisfl" and any(s.name == "ControllerService" for s in files[0].service)
        methods = {m.name for s in files[0].service for m in s.method}
        assert {"JoinFederation", "MarkTaskCompleted", "ReplaceCommunityModel"} <= methods
        assert len(files) >= 3                                       # the file and its imports (model, metis, ...)
        assert answers[2].WhichOneof("message_response") == "error_response"
        assert answers[2].error_response.error_code == grpc.StatusCode.NOT_FOUND.value[0]
        ch.close()
        # the operations helper built on the two services
        from metisfl_b200.utils import grpc_probe
        status, names = grpc_probe.probe(f"127.0.0.1:{port}", list_services=True)
        assert status == "SERVING" and "metisfl.ControllerService" in names
        assert grpc_probe.main([f"127.0.0.1:{port}", "--service", "metisfl.ControllerService"]) == 0
    finally:
        ctrl.stop()
    from metisfl_b200.utils import grpc_probe
    assert grpc_probe.main([f"127.0.0.1:{port}", "--timeout", "1"]) == 1      # the server is gone
