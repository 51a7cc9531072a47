"""gRPC channel / server helpers (API parity with reference ``metisfl/utils/grpc_services.py:13-107``):
unlimited message sizes both ways (whole models travel in one unary message on the compat path),
optional TLS from ``SSLConfig``, and a retrying request wrapper."""
import concurrent.futures
import time

import grpc

from metisfl_b200.proto.metis_pb2 import ServerEntity
from metisfl_b200.utils.metis_logger import MetisLogger
from metisfl_b200.utils.ssl_configurator import SSLConfigurator

_UNLIMITED = [("grpc.max_send_message_length", -1), ("grpc.max_receive_message_length", -1)]


class GRPCEndpoint(object):
    def __init__(self, server_entity: ServerEntity):
        self.server_entity = server_entity
        self.listening_endpoint = "{}:{}".format(server_entity.hostname, server_entity.port)


class GRPCChannelMaxMsgLength(object):
    def __init__(self, server_entity: ServerEntity):
        self.grpc_endpoint = GRPCEndpoint(server_entity)
        self.channel_opts = list(_UNLIMITED)
        public_certificate, _ = SSLConfigurator.load_certificates_from_ssl_config_pb(
            server_entity.ssl_config, as_stream=True)
        if public_certificate:
            creds = grpc.ssl_channel_credentials(public_certificate)
            self.channel = grpc.secure_channel(self.grpc_endpoint.listening_endpoint, creds, options=self.channel_opts)
        else:
            self.channel = grpc.insecure_channel(self.grpc_endpoint.listening_endpoint, options=self.channel_opts)


class GRPCServerClient(object):
    def __init__(self, server_entity: ServerEntity, max_workers=1):
        self.grpc_endpoint = GRPCEndpoint(server_entity)
        self.executor = concurrent.futures.ThreadPoolExecutor(max_workers=max_workers)
        self.executor_pool = []
        self._channel = self.get_channel()

    def get_channel(self):
        """Channel to the endpoint (one per client object; reopened on demand after shutdown)."""
        if getattr(self, "_channel", None) is None:
            self._channel = GRPCChannelMaxMsgLength(self.grpc_endpoint.server_entity).channel
        return self._channel

    def request_with_timeout(self, request_fn, request_timeout, request_retries):
        count_retries = 0
        response = None
        while count_retries < request_retries:
            try:
                response = request_fn(request_timeout)
            except grpc.RpcError as rpc_error:
                if rpc_error.code() == grpc.StatusCode.UNAVAILABLE and count_retries + 1 < request_retries:
                    MetisLogger.info("%s unavailable, retrying (%d/%d)", self.grpc_endpoint.listening_endpoint,
                                     count_retries + 1, request_retries)
                    time.sleep(min(10.0, 0.5 * (count_retries + 1)))
                else:
                    MetisLogger.warning("RPC to %s failed: %s", self.grpc_endpoint.listening_endpoint,
                                        rpc_error.code())
                    if count_retries + 1 >= request_retries:
                        raise
            else:
                break
            count_retries += 1
        return response

    def _submit(self, request_fn, request_retries, request_timeout, block):
        """Run ``request_fn`` through the retry wrapper, inline or on the client's worker pool."""
        if block:
            return self.request_with_timeout(request_fn, request_timeout, request_retries)
        fut = self.executor.submit(self.request_with_timeout, request_fn, request_timeout, request_retries)
        self.executor_pool.append(fut)
        return fut

    def shutdown(self):
        self.executor.shutdown(wait=True, cancel_futures=True)
        if self._channel is not None:
            self._channel.close()
            self._channel = None


class GRPCServerMaxMsgLength(object):
    def __init__(self, max_workers=None, server_entity: ServerEntity = None):
        self.grpc_endpoint = GRPCEndpoint(server_entity)
        self.channel_opts = list(_UNLIMITED)
        self.executor = concurrent.futures.ThreadPoolExecutor(max_workers=max_workers)
        self.server = grpc.server(self.executor, options=self.channel_opts)
        public_certificate, private_key = SSLConfigurator.load_certificates_from_ssl_config_pb(
            server_entity.ssl_config, as_stream=True)
        bind = "{}:{}".format("0.0.0.0" if server_entity.hostname not in ("localhost", "127.0.0.1") else
                              server_entity.hostname, server_entity.port)
        if public_certificate and private_key:
            creds = grpc.ssl_server_credentials(((private_key, public_certificate),))
            self.bound_port = self.server.add_secure_port(bind, creds)
        else:
            self.bound_port = self.server.add_insecure_port(bind)
