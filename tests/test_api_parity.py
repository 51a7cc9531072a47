"""Public Python API names of the reference stay importable, under both package names."""
import subprocess
import sys


def test_metisfl_import_paths():
    from metisfl.controller.controller import ControllerWrapper  # noqa: F401
    from metisfl.controller.controller_instance import ControllerInstance  # noqa: F401
    from metisfl.driver.driver_session import DriverSession
    from metisfl.encryption.fhe import CKKS  # noqa: F401
    from metisfl.learner.learner import Learner  # noqa: F401
    from metisfl.learner.learner_servicer import LearnerServicer  # noqa: F401
    from metisfl.models.model_dataset import (ModelDataset, ModelDatasetClassification,  # noqa: F401
                                              ModelDatasetRegression)
    from metisfl.models.model_def import ModelDef, PyTorchDef  # noqa: F401
    from metisfl.models.model_ops import ModelOps  # noqa: F401
    from metisfl.proto import (controller_pb2, controller_pb2_grpc, learner_pb2, learner_pb2_grpc,  # noqa: F401
                               metis_pb2, model_pb2, service_common_pb2)
    from metisfl.utils.fedenv_parser import FederationEnvironment  # noqa: F401
    from metisfl.utils.grpc_controller_client import GRPCControllerClient  # noqa: F401
    from metisfl.utils.grpc_learner_client import GRPCLearnerClient  # noqa: F401
    from metisfl.utils.proto_messages_factory import (ControllerServiceProtoMessages,  # noqa: F401
                                                      LearnerServiceProtoMessages, MetisProtoMessages,
                                                      ModelProtoMessages, ServiceCommonProtoMessages)
    import metisfl_b200.driver.driver_session as real
    assert DriverSession is real.DriverSession
    assert hasattr(controller_pb2_grpc, "ControllerServiceStub") and hasattr(learner_pb2_grpc, "LearnerServiceStub")
    assert hasattr(controller_pb2_grpc, "add_ControllerServiceServicer_to_server")
    w = ControllerWrapper()
    assert all(hasattr(w, m) for m in ("start", "shutdown", "shutdown_request_received", "wait"))


def test_cli_entry_points_have_the_reference_flags():
    out = subprocess.run([sys.executable, "-m", "metisfl.controller", "--help"], capture_output=True, text=True,
                         timeout=60).stdout
    for flag in ("-e", "-g", "-c", "-m", "-s"):
        assert f"{flag}," in out or f"{flag} " in out
    out = subprocess.run([sys.executable, "-m", "metisfl.learner", "--help"], capture_output=True, text=True,
                         timeout=60).stdout
    for flag in ("-l", "-c", "-f", "-e", "-m", "-t", "-v", "-s", "-u", "-w", "-z"):
        assert f"{flag}," in out or f"{flag} " in out
