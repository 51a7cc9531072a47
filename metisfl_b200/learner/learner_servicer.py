"""gRPC face of a learner: ``LearnerService`` (4 RPCs).  Behavioural spec = reference
``metisfl/learner/learner_servicer.py:14-139``: ``RunTask`` returns immediately and pre-empts the
running task, ``EvaluateModel`` blocks, ``ShutDown`` stops serving, shuts the learner down, leaves
the federation, then releases the server; the servicer starts serving BEFORE it joins (:25-35)."""
import threading

from google.protobuf.timestamp_pb2 import Timestamp

from metisfl_b200.learner.learner import Learner
from metisfl_b200.proto import learner_pb2_grpc
from metisfl_b200.utils.grpc_services import GRPCServerMaxMsgLength
from metisfl_b200.utils.metis_logger import MetisLogger
from metisfl_b200.utils.proto_messages_factory import (LearnerServiceProtoMessages, ServiceCommonProtoMessages)


def _ack(status=True, message=None):
    t = Timestamp()
    t.GetCurrentTime()
    return ServiceCommonProtoMessages.construct_ack_pb(status=status, google_timestamp=t, message=message)


class LearnerServicer(learner_pb2_grpc.LearnerServiceServicer):

    def __init__(self, learner: Learner, servicer_workers=10, *args, **kwargs):
        self.learner = learner
        self.servicer_workers = servicer_workers
        self.__community_models_received = 0
        self.__model_evaluation_requests = 0
        self.__not_serving_event = threading.Event()
        self.__shutdown_event = threading.Event()
        self.__grpc_server = None
        self.__health = None

    def init_servicer(self):
        self.__grpc_server = GRPCServerMaxMsgLength(max_workers=self.servicer_workers,
                                                    server_entity=self.learner.learner_server_entity)
        learner_pb2_grpc.add_LearnerServiceServicer_to_server(self, self.__grpc_server.server)
        from metisfl_b200.proto.grpc_std_services import add_standard_services
        self.__health = add_standard_services(self.__grpc_server.server, ["metisfl.LearnerService"])
        self.__grpc_server.server.start()
        MetisLogger.info("Learner servicer listening on %s", self.learner.host_port_identifier())
        self.learner.join_federation()

    def wait_servicer(self):
        self.__shutdown_event.wait()
        self.__health.enter_graceful_shutdown()
        self.__grpc_server.server.stop(None)

    def _refuse(self, context):
        import grpc
        context.set_code(grpc.StatusCode.UNAVAILABLE)
        context.set_details("learner is shutting down")

    def EvaluateModel(self, request, context):
        if self.__not_serving_event.is_set():
            self._refuse(context)
            return LearnerServiceProtoMessages.construct_evaluate_model_response_pb(None)
        self.__model_evaluation_requests += 1
        evaluations_pb = self.learner.run_evaluation_task(
            request.model, request.batch_size, list(request.evaluation_dataset), request.metrics,
            cancel_running_tasks=False, block=True, verbose=False)
        return LearnerServiceProtoMessages.construct_evaluate_model_response_pb(evaluations_pb)

    def GetServicesHealthStatus(self, request, context):
        return ServiceCommonProtoMessages.construct_get_services_health_status_response_pb(
            services_status={"server": self.__grpc_server is not None and not self.__not_serving_event.is_set()})

    def RunTask(self, request, context):
        if self.__not_serving_event.is_set():
            self._refuse(context)
            return LearnerServiceProtoMessages.construct_run_task_response_pb(_ack(False))
        self.__community_models_received += 1
        from metisfl_b200.utils.fault_injector import FaultInjector
        if FaultInjector.instance().on_task(self.learner.host_port_identifier(), request.task.global_iteration) == "drop":
            return LearnerServiceProtoMessages.construct_run_task_response_pb(_ack(True))  # swallowed on purpose
        accepted = self.learner.run_learning_task(
            request.task, request.hyperparameters, request.federated_model.model,
            cancel_running_tasks=True, block=False, verbose=False)
        return LearnerServiceProtoMessages.construct_run_task_response_pb(_ack(bool(accepted)))

    def AggregateOnDevice(self, request, context):
        """Extension RPC (no reference counterpart): run the aggregation kernels over the device slots the controller
        names.  Answers with the device-timed duration, the quantifier counts and a reference to the new community
        model."""
        import grpc
        try:
            return self.learner.aggregate_on_device(request)
        except Exception as e:
            MetisLogger.error("AggregateOnDevice failed: %r", e)
            context.set_code(grpc.StatusCode.FAILED_PRECONDITION)
            context.set_details(repr(e))
            from metisfl_b200.proto import learner_pb2
            return learner_pb2.AggregateOnDeviceResponse(ack=_ack(False, repr(e)))

    def ShutDown(self, request, context):
        self.__not_serving_event.set()
        # answer first, then tear down (the caller must not wait for a possibly long drain)
        def _teardown():
            try:
                self.learner.shutdown(cancel_train_running_tasks=True, cancel_eval_running_tasks=False,
                                      cancel_infer_running_tasks=True)
            finally:
                try:
                    self.learner.leave_federation()
                except Exception as e:  # controller may already be gone
                    MetisLogger.warning("leave_federation failed: %r", e)
                self.learner.close()
                self.__shutdown_event.set()
        threading.Thread(target=_teardown, daemon=True).start()
        return ServiceCommonProtoMessages.construct_shutdown_response_pb(_ack(True))
