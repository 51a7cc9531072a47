"""Build gRPC client stubs / servicer bases from the service descriptors in ``schema.py``.

Replaces what ``grpc_tools.protoc`` would generate (reference: generated
``metisfl/proto/*_pb2_grpc.py``): method paths are ``/metisfl.<Service>/<Method>``.
"""
import grpc

from metisfl_b200.proto import schema


def _serializer(message_cls):
    """Requests assembled at the byte level (``proto/envelope.py``) go out as they are."""
    def serialize(message):
        if isinstance(message, (bytes, bytearray)):
            return bytes(message)
        return message.SerializeToString()
    return serialize


def make_service(file_name: str, service_name: str):
    sd = schema.file_descriptor(file_name).services_by_name[service_name]
    methods = [(m.name, schema.message_class(m.input_type.name),
                schema.message_class(m.output_type.name)) for m in sd.methods]
    full = sd.full_name

    class Stub:
        def __init__(self, channel):
            for name, req, resp in methods:
                setattr(self, name, channel.unary_unary(
                    f"/{full}/{name}",
                    request_serializer=_serializer(req),
                    response_deserializer=resp.FromString))

    def _unimplemented(name):
        def handler(self, request, context):
            context.set_code(grpc.StatusCode.UNIMPLEMENTED)
            context.set_details(f"Method {name} not implemented!")
            raise NotImplementedError(f"Method {name} not implemented!")
        handler.__name__ = name
        return handler

    Servicer = type(f"{service_name}Servicer", (object,),
                    {name: _unimplemented(name) for name, _, _ in methods})

    def add_to_server(servicer, server):
        # methods listed in ``servicer.RAW_REQUEST_METHODS`` get the request as the bytes that arrived (they take the
        # model record out themselves instead of paying for a full parse + re-serialization)
        raw = set(getattr(servicer, "RAW_REQUEST_METHODS", ()))
        handlers = {
            name: grpc.unary_unary_rpc_method_handler(
                getattr(servicer, name),
                request_deserializer=None if name in raw else req.FromString,
                response_serializer=resp.SerializeToString)
            for name, req, resp in methods
        }
        server.add_generic_rpc_handlers(
            (grpc.method_handlers_generic_handler(full, handlers),))

    Stub.__name__ = f"{service_name}Stub"
    return Stub, Servicer, add_to_server
