"""Client of ``metisfl.ControllerService`` (API parity with reference
``metisfl/utils/grpc_controller_client.py:11-297``: same method names, ``block`` /
``request_retries`` / ``request_timeout`` knobs; ``join_federation`` strips the private key and
persists / reloads credentials on ``ALREADY_EXISTS``).  Every call is one small closure handed to
``GRPCServerClient._submit`` instead of eleven copies of the scheduling boilerplate."""
import os

import grpc
from google.protobuf.json_format import MessageToDict  # noqa: F401  (re-exported for callers)

from metisfl_b200.proto import controller_pb2_grpc
from metisfl_b200.utils import proto_messages_factory as proto_factory
from metisfl_b200.utils.grpc_services import GRPCServerClient
from metisfl_b200.utils.metis_logger import MetisLogger
from metisfl_b200.utils.ssl_configurator import SSLConfigurator


class GRPCControllerClient(GRPCServerClient):

    def __init__(self, controller_server_entity, max_workers=1):
        super().__init__(controller_server_entity, max_workers)
        self._stub = controller_pb2_grpc.ControllerServiceStub(self._channel)

    def _call(self, rpc_name, request_pb, on_response, request_retries, request_timeout, block):
        def _request(_timeout=None):
            response = getattr(self._stub, rpc_name)(request_pb, timeout=_timeout)
            return on_response(response) if on_response else response
        return self._submit(_request, request_retries, request_timeout, block)

    def check_health_status(self, request_retries=1, request_timeout=None, block=True):
        req = proto_factory.ServiceCommonProtoMessages.construct_get_services_health_status_request_pb()
        return self._call("GetServicesHealthStatus", req, None, request_retries, request_timeout, block)

    def join_federation(self, learner_server_entity, learner_id_fp, auth_token_fp, train_dataset_size,
                        train_dataset_specs, validation_dataset_size, validation_dataset_specs, test_dataset_size,
                        test_dataset_specs, is_classification, is_regression, request_retries=1,
                        request_timeout=None, block=True, fabric_pb=None):
        # the controller only ever sees the learner's PUBLIC certificate
        public_entity = proto_factory.MetisProtoMessages.construct_server_entity_pb(
            hostname=learner_server_entity.hostname, port=learner_server_entity.port,
            ssl_config_pb=SSLConfigurator.gen_public_ssl_config_pb_as_stream(learner_server_entity.ssl_config))
        dataset_spec_pb = proto_factory.MetisProtoMessages.construct_dataset_spec_pb(
            num_training_examples=train_dataset_size, num_validation_examples=validation_dataset_size,
            num_test_examples=test_dataset_size, training_spec=train_dataset_specs,
            validation_spec=validation_dataset_specs, test_spec=test_dataset_specs,
            is_classification=is_classification, is_regression=is_regression)
        req = proto_factory.ControllerServiceProtoMessages.construct_join_federation_request_pb(
            server_entity_pb=public_entity, local_dataset_spec_pb=dataset_spec_pb)
        if fabric_pb is not None:  # this learner hosts a rank of the on-box fabric
            req.fabric.CopyFrom(fabric_pb)

        def _request(_timeout=None):
            try:
                response = self._stub.JoinFederation(req, timeout=_timeout)
            except grpc.RpcError as rpc_error:
                if rpc_error.code() == grpc.StatusCode.ALREADY_EXISTS and os.path.exists(learner_id_fp):
                    with open(learner_id_fp) as f1, open(auth_token_fp) as f2:
                        learner_id, auth_token = f1.read().strip(), f2.read().strip()
                    MetisLogger.info("Learner re-joined the federation with id %s", learner_id)
                    return learner_id, auth_token, True
                raise
            for path, value in ((learner_id_fp, response.learner_id), (auth_token_fp, response.auth_token)):
                with open(path, "w+") as f:
                    f.write(value.strip())
            MetisLogger.info("Joined the federation with assigned id %s", response.learner_id)
            return response.learner_id, response.auth_token, response.ack.status

        return self._submit(_request, request_retries, request_timeout, block)

    def leave_federation(self, learner_id, auth_token, request_retries=1, request_timeout=None, block=True):
        req = proto_factory.ControllerServiceProtoMessages.construct_leave_federation_request_pb(
            learner_id=learner_id, auth_token=auth_token)
        return self._call("LeaveFederation", req, lambda r: r.ack.status, request_retries, request_timeout, block)

    def mark_task_completed(self, learner_id, auth_token, completed_task_pb, request_retries=1,
                            request_timeout=None, block=True):
        # assembled around the serialized task (proto/envelope.py): no copy of the model into a request object
        from metisfl_b200.proto import envelope
        req = envelope.mark_task_completed_request(learner_id, auth_token, completed_task_pb)
        return self._call("MarkTaskCompleted", req, lambda r: r.ack.status, request_retries, request_timeout, block)

    def get_community_model_evaluation_lineage(self, num_backtracks, request_retries=1, request_timeout=None,
                                               block=True):
        req = proto_factory.ControllerServiceProtoMessages \
            .construct_get_community_model_evaluation_lineage_request_pb(num_backtracks)
        return self._call("GetCommunityModelEvaluationLineage", req, None, request_retries, request_timeout, block)

    def get_community_model_lineage(self, num_backtracks, request_retries=1, request_timeout=None, block=True):
        req = proto_factory.ControllerServiceProtoMessages.construct_get_community_model_lineage_request_pb(
            num_backtracks)
        return self._call("GetCommunityModelLineage", req, None, request_retries, request_timeout, block)

    def get_local_task_lineage(self, num_backtracks, learner_ids, request_retries=1, request_timeout=None,
                               block=True):
        req = proto_factory.ControllerServiceProtoMessages.construct_get_local_task_lineage_request_pb(
            num_backtracks, learner_ids)
        return self._call("GetLocalTaskLineage", req, None, request_retries, request_timeout, block)

    def get_participating_learners(self, request_retries=1, request_timeout=None, block=True):
        req = proto_factory.ControllerServiceProtoMessages.construct_get_participating_learners_request_pb()
        return self._call("GetParticipatingLearners", req, None, request_retries, request_timeout, block)

    def get_runtime_metadata(self, num_backtracks, request_retries=1, request_timeout=None, block=True):
        req = proto_factory.ControllerServiceProtoMessages.construct_get_runtime_metadata_lineage_request_pb(
            num_backtracks)
        return self._call("GetRuntimeMetadataLineage", req, None, request_retries, request_timeout, block)

    def replace_community_model(self, num_contributors, model_pb, request_retries=1, request_timeout=None,
                                block=True):
        federated_model_pb = proto_factory.ModelProtoMessages.construct_federated_model_pb(num_contributors, model_pb)
        req = proto_factory.ControllerServiceProtoMessages.construct_replace_community_model_request_pb(
            federated_model_pb)
        return self._call("ReplaceCommunityModel", req, lambda r: r.ack.status, request_retries, request_timeout,
                          block)

    def shutdown_controller(self, request_retries=1, request_timeout=None, block=True):
        req = proto_factory.ServiceCommonProtoMessages.construct_shutdown_request_pb()
        return self._call("ShutDown", req, lambda r: r.ack.status, request_retries, request_timeout, block)
