"""Launch command lines for the controller / learner processes (API parity with reference
``metisfl/utils/init_services_factory.py:1-85``).  Serialized protos travel as hex strings because
raw protobuf bytes do not survive shell quoting."""
import shlex
import sys


class MetisInitServicesCmdFactory(object):
    PACKAGE = "metisfl_b200"

    @classmethod
    def _cmd(cls, module, hex_flags, str_flags=()):
        parts = [shlex.quote(sys.executable or "python"), "-m", f"{cls.PACKAGE}.{module}"]
        parts += [f"--{flag}={blob.hex()}" for flag, blob in hex_flags if blob is not None]
        parts += [f"--{flag}={shlex.quote(str(value))}" for flag, value in str_flags if value not in (None, "")]
        return " ".join(parts) + " "

    @classmethod
    def init_controller_target(cls, controller_server_entity_pb_ser, global_model_specs_pb_ser,
                               communication_specs_pb_ser, model_hyperparameters_pb_ser, model_store_config_pb_ser):
        return cls._cmd("controller", [
            ("controller_server_entity_protobuff_serialized_hexadecimal", controller_server_entity_pb_ser),
            ("global_model_specs_protobuff_serialized_hexadecimal", global_model_specs_pb_ser),
            ("communication_specs_protobuff_serialized_hexadecimal", communication_specs_pb_ser),
            ("model_hyperparameters_protobuff_serialized_hexadecimal", model_hyperparameters_pb_ser),
            ("model_store_config_protobuff_serialized_hexadecimal", model_store_config_pb_ser)])

    @classmethod
    def init_learner_target(cls, learner_server_entity_pb_ser, controller_server_entity_pb_ser, he_scheme_pb_ser,
                            model_dir, train_dataset, train_dataset_recipe, validation_dataset="", test_dataset="",
                            validation_dataset_recipe="", test_dataset_recipe="", neural_engine="pytorch",
                            device=None, logical_learners=None, fabric=None):
        return cls._cmd("learner", [
            ("learner_server_entity_protobuff_serialized_hexadecimal", learner_server_entity_pb_ser),
            ("controller_server_entity_protobuff_serialized_hexadecimal", controller_server_entity_pb_ser),
            ("he_scheme_protobuff_serialized_hexadecimal", he_scheme_pb_ser)],
            [("neural_engine", neural_engine), ("model_dir", model_dir), ("train_dataset", train_dataset),
             ("validation_dataset", validation_dataset), ("test_dataset", test_dataset),
             ("train_dataset_recipe", train_dataset_recipe),
             ("validation_dataset_recipe", validation_dataset_recipe),
             ("test_dataset_recipe", test_dataset_recipe), ("device", device),
             ("logical_learners", logical_learners), ("fabric", fabric)])
