"""``python -m metisfl_b200.controller`` - same hex-encoded protobuf flags and defaults as the
reference launcher (``controller/__main__.py:12-121``): FedAvg / NUM_TRAINING_EXAMPLES, synchronous,
batch 100 x 5 epochs, VanillaSGD 0.01, in-memory store without eviction."""
import argparse

from metisfl_b200.controller.controller_instance import ControllerInstance
from metisfl_b200.proto import metis_pb2
from metisfl_b200.utils.proto_messages_factory import MetisProtoMessages, ModelProtoMessages


def _from_hex(cls, hex_str, default):
    if hex_str is None:
        return default
    return cls.FromString(bytes.fromhex(hex_str))


def init_controller(controller_server_entity_hex=None, global_model_specs_hex=None, communication_specs_hex=None,
                    model_hyperparameters_hex=None, model_store_config_hex=None, block=True):
    server_entity = _from_hex(metis_pb2.ServerEntity, controller_server_entity_hex,
                              MetisProtoMessages.construct_server_entity_pb("[::]", 50051))
    global_model_specs = _from_hex(
        metis_pb2.GlobalModelSpecs, global_model_specs_hex,
        MetisProtoMessages.construct_global_model_specs(
            MetisProtoMessages.construct_aggregation_rule_pb("FEDAVG", "NUMTRAININGEXAMPLES", None, None), 1))
    communication_specs = _from_hex(metis_pb2.CommunicationSpecs, communication_specs_hex,
                                    MetisProtoMessages.construct_communication_specs_pb("SYNCHRONOUS"))
    hyper = _from_hex(
        metis_pb2.ControllerParams.ModelHyperparams, model_hyperparameters_hex,
        MetisProtoMessages.construct_controller_modelhyperparams_pb(
            100, 5, ModelProtoMessages.construct_optimizer_config_pb(
                ModelProtoMessages.construct_vanilla_sgd_optimizer_pb(0.01)), 0.0))
    store = _from_hex(metis_pb2.ModelStoreConfig, model_store_config_hex,
                      MetisProtoMessages.construct_model_store_config_pb("InMemory", "NoEviction"))
    params = MetisProtoMessages.construct_controller_params_pb(server_entity, global_model_specs,
                                                               communication_specs, store, hyper)
    instance = ControllerInstance()
    instance.start(params)
    if block:
        instance.shutdown()
    return instance


def main():
    ap = argparse.ArgumentParser("metisfl_b200.controller")
    ap.add_argument("-e", "--controller_server_entity_protobuff_serialized_hexadecimal", type=str, default=None)
    ap.add_argument("-g", "--global_model_specs_protobuff_serialized_hexadecimal", type=str, default=None)
    ap.add_argument("-c", "--communication_specs_protobuff_serialized_hexadecimal", type=str, default=None)
    ap.add_argument("-m", "--model_hyperparameters_protobuff_serialized_hexadecimal", type=str, default=None)
    ap.add_argument("-s", "--model_store_config_protobuff_serialized_hexadecimal", type=str, default=None)
    a = ap.parse_args()
    init_controller(a.controller_server_entity_protobuff_serialized_hexadecimal,
                    a.global_model_specs_protobuff_serialized_hexadecimal,
                    a.communication_specs_protobuff_serialized_hexadecimal,
                    a.model_hyperparameters_protobuff_serialized_hexadecimal,
                    a.model_store_config_protobuff_serialized_hexadecimal)


if __name__ == "__main__":
    main()
