"""The two standard gRPC services the reference controller enables on its server
(``controller_servicer.cc:80-84``: ``EnableDefaultHealthCheckService(true)`` and the proto reflection plugin):

* ``grpc.health.v1.Health`` (``Check`` / ``Watch``) - what ``grpc_health_probe`` and Kubernetes probes call;
* ``grpc.reflection.v1alpha.ServerReflection`` - what ``grpcurl`` / ``grpc_cli`` use to list services and fetch schemas.

``grpcio-health-checking`` and ``grpcio-reflection`` are not part of this image, so both are built here the same way
as the MetisFL schema (``schema.py``): ``FileDescriptorProto`` objects registered in the descriptor pool, message
classes from ``message_factory``, generic RPC handlers.  Field numbers are those of the public ``health.proto`` and
``reflection.proto``.
"""
from __future__ import annotations

import threading

import grpc
from google.protobuf import descriptor_pb2 as dpb
from google.protobuf import descriptor_pool, message_factory

_FD = dpb.FieldDescriptorProto
_POOL = descriptor_pool.Default()

HEALTH_SERVICE = "grpc.health.v1.Health"
REFLECTION_SERVICE = "grpc.reflection.v1alpha.ServerReflection"


def _field(msg, name, number, typ, type_name=None, repeated=False, oneof=None):
    fd = msg.field.add(name=name, number=number, type=typ,
                       label=_FD.LABEL_REPEATED if repeated else _FD.LABEL_OPTIONAL)
    if type_name:
        fd.type_name = type_name
    if oneof is not None:
        fd.oneof_index = oneof
    return fd


def _health_file() -> dpb.FileDescriptorProto:
    f = dpb.FileDescriptorProto(name="grpc/health/v1/health.proto", package="grpc.health.v1", syntax="proto3")
    req = f.message_type.add(name="HealthCheckRequest")
    _field(req, "service", 1, _FD.TYPE_STRING)
    resp = f.message_type.add(name="HealthCheckResponse")
    e = resp.enum_type.add(name="ServingStatus")
    for k, v in (("UNKNOWN", 0), ("SERVING", 1), ("NOT_SERVING", 2), ("SERVICE_UNKNOWN", 3)):
        e.value.add(name=k, number=v)
    _field(resp, "status", 1, _FD.TYPE_ENUM, ".grpc.health.v1.HealthCheckResponse.ServingStatus")
    svc = f.service.add(name="Health")
    svc.method.add(name="Check", input_type=".grpc.health.v1.HealthCheckRequest",
                   output_type=".grpc.health.v1.HealthCheckResponse")
    svc.method.add(name="Watch", input_type=".grpc.health.v1.HealthCheckRequest",
                   output_type=".grpc.health.v1.HealthCheckResponse", server_streaming=True)
    return f


def _reflection_file() -> dpb.FileDescriptorProto:
    P = ".grpc.reflection.v1alpha."
    f = dpb.FileDescriptorProto(name="grpc/reflection/v1alpha/reflection.proto", package="grpc.reflection.v1alpha",
                                syntax="proto3")
    ext = f.message_type.add(name="ExtensionRequest")
    _field(ext, "containing_type", 1, _FD.TYPE_STRING)
    _field(ext, "extension_number", 2, _FD.TYPE_INT32)
    req = f.message_type.add(name="ServerReflectionRequest")
    req.oneof_decl.add(name="message_request")
    _field(req, "host", 1, _FD.TYPE_STRING)
    _field(req, "file_by_filename", 3, _FD.TYPE_STRING, oneof=0)
    _field(req, "file_containing_symbol", 4, _FD.TYPE_STRING, oneof=0)
    _field(req, "file_containing_extension", 5, _FD.TYPE_MESSAGE, P + "ExtensionRequest", oneof=0)
    _field(req, "all_extension_numbers_of_type", 6, _FD.TYPE_STRING, oneof=0)
    _field(req, "list_services", 7, _FD.TYPE_STRING, oneof=0)
    fdr = f.message_type.add(name="FileDescriptorResponse")
    _field(fdr, "file_descriptor_proto", 1, _FD.TYPE_BYTES, repeated=True)
    enr = f.message_type.add(name="ExtensionNumberResponse")
    _field(enr, "base_type_name", 1, _FD.TYPE_STRING)
    _field(enr, "extension_number", 2, _FD.TYPE_INT32, repeated=True)
    sr = f.message_type.add(name="ServiceResponse")
    _field(sr, "name", 1, _FD.TYPE_STRING)
    lsr = f.message_type.add(name="ListServiceResponse")
    _field(lsr, "service", 1, _FD.TYPE_MESSAGE, P + "ServiceResponse", repeated=True)
    er = f.message_type.add(name="ErrorResponse")
    _field(er, "error_code", 1, _FD.TYPE_INT32)
    _field(er, "error_message", 2, _FD.TYPE_STRING)
    resp = f.message_type.add(name="ServerReflectionResponse")
    resp.oneof_decl.add(name="message_response")
    _field(resp, "valid_host", 1, _FD.TYPE_STRING)
    _field(resp, "original_request", 2, _FD.TYPE_MESSAGE, P + "ServerReflectionRequest")
    _field(resp, "file_descriptor_response", 4, _FD.TYPE_MESSAGE, P + "FileDescriptorResponse", oneof=0)
    _field(resp, "all_extension_numbers_response", 5, _FD.TYPE_MESSAGE, P + "ExtensionNumberResponse", oneof=0)
    _field(resp, "list_services_response", 6, _FD.TYPE_MESSAGE, P + "ListServiceResponse", oneof=0)
    _field(resp, "error_response", 7, _FD.TYPE_MESSAGE, P + "ErrorResponse", oneof=0)
    svc = f.service.add(name="ServerReflection")
    svc.method.add(name="ServerReflectionInfo", input_type=P + "ServerReflectionRequest",
                   output_type=P + "ServerReflectionResponse", client_streaming=True, server_streaming=True)
    return f


def _register(fdp):
    try:
        return _POOL.FindFileByName(fdp.name)
    except KeyError:
        pass
    try:
        return _POOL.Add(fdp) or _POOL.FindFileByName(fdp.name)
    except TypeError:  # pragma: no cover - very old runtimes
        _POOL.AddSerializedFile(fdp.SerializeToString())
        return _POOL.FindFileByName(fdp.name)


_HEALTH_FD = _register(_health_file())
_REFL_FD = _register(_reflection_file())


def _cls(full_name):
    return message_factory.GetMessageClass(_POOL.FindMessageTypeByName(full_name))


HealthCheckRequest = _cls("grpc.health.v1.HealthCheckRequest")
HealthCheckResponse = _cls("grpc.health.v1.HealthCheckResponse")
ServerReflectionRequest = _cls("grpc.reflection.v1alpha.ServerReflectionRequest")
ServerReflectionResponse = _cls("grpc.reflection.v1alpha.ServerReflectionResponse")

SERVING, NOT_SERVING, SERVICE_UNKNOWN = 1, 2, 3


class HealthServicer:
    """``grpc.health.v1.Health``.  The empty service name stands for the whole server; ``set`` flips a status and wakes
    the ``Watch`` streams."""

    def __init__(self, services=()):
        self._cv = threading.Condition()
        self._status = {"": SERVING}
        self._shutdown = False
        for s in services:
            self._status[s] = SERVING

    def set(self, service: str, status: int):
        with self._cv:
            self._status[service] = status
            self._cv.notify_all()

    def enter_graceful_shutdown(self):
        with self._cv:
            for k in self._status:
                self._status[k] = NOT_SERVING
            self._shutdown = True
            self._cv.notify_all()

    def Check(self, request, context):
        with self._cv:
            st = self._status.get(request.service)
        if st is None:
            context.abort(grpc.StatusCode.NOT_FOUND, "unknown service")
        return HealthCheckResponse(status=st)

    def Watch(self, request, context):
        last = None
        while context.is_active():
            with self._cv:
                st = self._status.get(request.service, SERVICE_UNKNOWN)
                if st == last:
                    if self._shutdown:
                        return
                    self._cv.wait(timeout=0.5)
                    continue
                shutdown = self._shutdown
            last = st
            yield HealthCheckResponse(status=st)
            if shutdown:
                return


class ReflectionServicer:
    """``grpc.reflection.v1alpha.ServerReflection`` over the default descriptor pool."""

    def __init__(self, service_names):
        self._services = list(service_names)

    @staticmethod
    def _with_deps(fd):
        seen, order = set(), []

        def walk(d):
            if d.name in seen:
                return
            seen.add(d.name)
            for dep in d.dependencies:
                walk(dep)
            order.append(d)
        walk(fd)
        out = []
        for d in reversed(order):          # the requested file first, as the C++ plugin does
            p = dpb.FileDescriptorProto()
            d.CopyToProto(p)
            out.append(p.SerializeToString())
        return out

    def _answer(self, req):
        resp = ServerReflectionResponse(valid_host=req.host)
        resp.original_request.CopyFrom(req)
        which = req.WhichOneof("message_request")
        try:
            if which == "list_services":
                for s in self._services:
                    resp.list_services_response.service.add(name=s)
            elif which == "file_by_filename":
                resp.file_descriptor_response.file_descriptor_proto.extend(
                    self._with_deps(_POOL.FindFileByName(req.file_by_filename)))
            elif which == "file_containing_symbol":
                resp.file_descriptor_response.file_descriptor_proto.extend(
                    self._with_deps(_POOL.FindFileContainingSymbol(req.file_containing_symbol)))
            elif which == "all_extension_numbers_of_type":
                _POOL.FindMessageTypeByName(req.all_extension_numbers_of_type)   # proto3 schema: no extensions
                resp.all_extension_numbers_response.base_type_name = req.all_extension_numbers_of_type
            else:
                resp.error_response.error_code = grpc.StatusCode.UNIMPLEMENTED.value[0]
                resp.error_response.error_message = "request not supported"
        except KeyError as e:
            resp.error_response.error_code = grpc.StatusCode.NOT_FOUND.value[0]
            resp.error_response.error_message = "not found: {}".format(e)
        return resp

    def ServerReflectionInfo(self, request_iterator, context):
        for req in request_iterator:
            yield self._answer(req)


def add_standard_services(server, service_names):
    """Register health checking + reflection on ``server`` for the given fully-qualified service names.  Returns the
    ``HealthServicer`` (call ``enter_graceful_shutdown`` before stopping the server)."""
    names = list(service_names) + [HEALTH_SERVICE, REFLECTION_SERVICE]
    health = HealthServicer(service_names)
    refl = ReflectionServicer(names)
    server.add_generic_rpc_handlers((
        grpc.method_handlers_generic_handler(HEALTH_SERVICE, {
            "Check": grpc.unary_unary_rpc_method_handler(
                health.Check, request_deserializer=HealthCheckRequest.FromString,
                response_serializer=HealthCheckResponse.SerializeToString),
            "Watch": grpc.unary_stream_rpc_method_handler(
                health.Watch, request_deserializer=HealthCheckRequest.FromString,
                response_serializer=HealthCheckResponse.SerializeToString)}),
        grpc.method_handlers_generic_handler(REFLECTION_SERVICE, {
            "ServerReflectionInfo": grpc.stream_stream_rpc_method_handler(
                refl.ServerReflectionInfo, request_deserializer=ServerReflectionRequest.FromString,
                response_serializer=ServerReflectionResponse.SerializeToString)}),
    ))
    return health
