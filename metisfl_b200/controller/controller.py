"""``ControllerWrapper`` - same four-call surface as the reference's pybind module
(``controller/pybind/controller_pybind.cc:51-67``: ``start(serialized ControllerParams)``,
``shutdown()``, ``shutdown_request_received()``, ``wait()``), backed by the native core
(``metisfl_b200._core``) plus the Python gRPC servicer."""
from metisfl_b200.controller.service import ControllerServicer
from metisfl_b200.proto.metis_pb2 import ControllerParams


class ControllerWrapper(object):

    def __init__(self):
        self._servicer = None

    def start(self, controller_params_serialized: bytes):
        params = ControllerParams.FromString(controller_params_serialized)
        self._servicer = ControllerServicer(params)
        self._servicer.start()

    @property
    def servicer(self):
        return self._servicer

    def shutdown(self):
        if self._servicer is not None:
            self._servicer.stop()
            self._servicer = None

    def shutdown_request_received(self) -> bool:
        return self._servicer is not None and self._servicer.shutdown_request_received()

    def wait(self):
        """Release everything after a ShutDown RPC was received."""
        self.shutdown()
