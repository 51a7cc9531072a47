"""``python -m metisfl_b200.utils.grpc_probe HOST:PORT [--service NAME] [--list]``

A small operations helper in the spirit of ``grpc_health_probe`` / ``grpcurl list``: asks a controller or learner
server for its health through ``grpc.health.v1.Health/Check`` and, with ``--list``, for the services it exposes through
the reflection service.  Exit code 0 = SERVING, 1 = anything else (usable as a container health check)."""
import argparse
import sys

import grpc

from metisfl_b200.proto.grpc_std_services import (HealthCheckRequest, HealthCheckResponse, ServerReflectionRequest,
                                                  ServerReflectionResponse)

_STATUS = {0: "UNKNOWN", 1: "SERVING", 2: "NOT_SERVING", 3: "SERVICE_UNKNOWN"}


def probe(target: str, service: str = "", list_services: bool = False, timeout: float = 5.0, certificate: bytes = None):
    """Returns ``(status_name, [service names] or None)``; raises ``grpc.RpcError`` when the server cannot be reached."""
    channel = (grpc.secure_channel(target, grpc.ssl_channel_credentials(certificate)) if certificate
               else grpc.insecure_channel(target))
    try:
        check = channel.unary_unary("/grpc.health.v1.Health/Check",
                                    request_serializer=HealthCheckRequest.SerializeToString,
                                    response_deserializer=HealthCheckResponse.FromString)
        status = _STATUS.get(check(HealthCheckRequest(service=service), timeout=timeout).status, "UNKNOWN")
        names = None
        if list_services:
            info = channel.stream_stream("/grpc.reflection.v1alpha.ServerReflection/ServerReflectionInfo",
                                         request_serializer=ServerReflectionRequest.SerializeToString,
                                         response_deserializer=ServerReflectionResponse.FromString)
            answer = next(iter(info(iter([ServerReflectionRequest(list_services="")]), timeout=timeout)))
            names = sorted(s.name for s in answer.list_services_response.service)
        return status, names
    finally:
        channel.close()


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("target", help="host:port of a controller or learner server")
    ap.add_argument("--service", default="", help="fully-qualified service name (default: the whole server)")
    ap.add_argument("--list", action="store_true", help="also list the services (server reflection)")
    ap.add_argument("--certificate", default=None, help="PEM file of the server certificate (TLS)")
    ap.add_argument("--timeout", type=float, default=5.0)
    a = ap.parse_args(argv)
    cert = open(a.certificate, "rb").read() if a.certificate else None
    try:
        status, names = probe(a.target, a.service, a.list, a.timeout, cert)
    except grpc.RpcError as e:
        print("{}: {}".format(a.target, e.code().name))
        return 1
    print("{}: {}".format(a.target, status))
    for n in names or []:
        print("  " + n)
    return 0 if status == "SERVING" else 1


if __name__ == "__main__":
    sys.exit(main())
