"""``python -m metisfl_b200.learner`` - same flags as the reference launcher
(``learner/__main__.py:88-137``: ``-l -c -f -e -m -t -v -s -u -w -z``) plus ``--device``."""
import argparse

from metisfl_b200.learner.learner import Learner
from metisfl_b200.learner.learner_servicer import LearnerServicer
from metisfl_b200.proto import metis_pb2


def _from_hex(cls, hex_str):
    return cls.FromString(bytes.fromhex(hex_str)) if hex_str else cls()


def init_learner(a, block=True):
    if (a.neural_engine or "").lower() == "echo":  # load-test backend: sleeps and echoes the model (no training)
        from metisfl_b200.learner.echo_learner import EchoLearner
        learner = EchoLearner(
            learner_server_entity=_from_hex(metis_pb2.ServerEntity,
                                            a.learner_server_entity_protobuff_serialized_hexadecimal),
            controller_server_entity=_from_hex(metis_pb2.ServerEntity,
                                               a.controller_server_entity_protobuff_serialized_hexadecimal),
            he_scheme_config_pb=metis_pb2.HESchemeConfig(), sleep_secs=a.echo_sleep_secs,
            learner_credentials_fp=a.credentials_dir)
        servicer = LearnerServicer(learner=learner, servicer_workers=5)
        servicer.init_servicer()
        if block:
            servicer.wait_servicer()
        return servicer
    learner = Learner(
        learner_server_entity=_from_hex(metis_pb2.ServerEntity, a.learner_server_entity_protobuff_serialized_hexadecimal),
        controller_server_entity=_from_hex(metis_pb2.ServerEntity,
                                           a.controller_server_entity_protobuff_serialized_hexadecimal),
        he_scheme_config_pb=_from_hex(metis_pb2.HESchemeConfig, a.he_scheme_protobuff_serialized_hexadecimal),
        nn_engine=a.neural_engine, model_dir=a.model_dir,
        train_dataset_fp=a.train_dataset, train_dataset_recipe_pkl=a.train_dataset_recipe,
        validation_dataset_fp=a.validation_dataset, validation_dataset_recipe_pkl=a.validation_dataset_recipe,
        test_dataset_fp=a.test_dataset, test_dataset_recipe_pkl=a.test_dataset_recipe, device=a.device,
        learner_credentials_fp=a.credentials_dir, fabric=a.fabric or None)
    servicer = LearnerServicer(learner=learner, servicer_workers=5)
    servicer.init_servicer()
    if block:
        servicer.wait_servicer()
    return servicer


def main():
    ap = argparse.ArgumentParser("metisfl_b200.learner")
    ap.add_argument("-l", "--learner_server_entity_protobuff_serialized_hexadecimal", type=str, default=None)
    ap.add_argument("-c", "--controller_server_entity_protobuff_serialized_hexadecimal", type=str, default=None)
    ap.add_argument("-f", "--he_scheme_protobuff_serialized_hexadecimal", type=str, default=None)
    ap.add_argument("-e", "--neural_engine", type=str, default="pytorch")
    ap.add_argument("-m", "--model_dir", type=str, default="")
    ap.add_argument("-t", "--train_dataset", type=str, default="")
    ap.add_argument("-v", "--validation_dataset", type=str, default="")
    ap.add_argument("-s", "--test_dataset", type=str, default="")
    ap.add_argument("-u", "--train_dataset_recipe", type=str, default="")
    ap.add_argument("-w", "--validation_dataset_recipe", type=str, default="")
    ap.add_argument("-z", "--test_dataset_recipe", type=str, default="")
    ap.add_argument("--device", type=str, default=None)
    ap.add_argument("--logical_learners", type=int, default=1)
    ap.add_argument("--fabric", type=str, default="",
                    help="host a rank of the on-box NVLink fabric: rank=R,world=W,init=tcp://127.0.0.1:PORT,rule=FedAvg")
    ap.add_argument("--credentials_dir", type=str, default="/tmp/metis/learner/")
    ap.add_argument("--echo_sleep_secs", type=float, default=3.0, help="task duration of the echo backend")
    init_learner(ap.parse_args())


if __name__ == "__main__":
    main()
