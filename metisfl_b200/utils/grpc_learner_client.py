"""Client of ``metisfl.LearnerService`` (API parity with reference
``metisfl/utils/grpc_learner_client.py:8-58``, plus the two task RPCs the controller's Python
transport needs)."""
from metisfl_b200.proto import learner_pb2_grpc
from metisfl_b200.utils import proto_messages_factory as proto_factory
from metisfl_b200.utils.grpc_services import GRPCServerClient


class GRPCLearnerClient(GRPCServerClient):

    def __init__(self, learner_server_entity, max_workers=1):
        super().__init__(learner_server_entity, max_workers)
        self._stub = learner_pb2_grpc.LearnerServiceStub(self._channel)

    def _call(self, rpc_name, request_pb, on_response, request_retries, request_timeout, block):
        def _request(_timeout=None):
            response = getattr(self._stub, rpc_name)(request_pb, timeout=_timeout)
            return on_response(response) if on_response else response
        return self._submit(_request, request_retries, request_timeout, block)

    def check_health_status(self, request_retries=1, request_timeout=None, block=True):
        req = proto_factory.ServiceCommonProtoMessages.construct_get_services_health_status_request_pb()
        return self._call("GetServicesHealthStatus", req, None, request_retries, request_timeout, block)

    def run_task(self, run_task_request_pb, request_retries=1, request_timeout=None, block=True):
        return self._call("RunTask", run_task_request_pb, None, request_retries, request_timeout, block)

    def evaluate_model(self, evaluate_model_request_pb, request_retries=1, request_timeout=None, block=True):
        return self._call("EvaluateModel", evaluate_model_request_pb, None, request_retries, request_timeout, block)

    def shutdown_learner(self, request_retries=1, request_timeout=None, block=True):
        req = proto_factory.ServiceCommonProtoMessages.construct_shutdown_request_pb()
        return self._call("ShutDown", req, lambda r: r.ack.status, request_retries, request_timeout, block)
