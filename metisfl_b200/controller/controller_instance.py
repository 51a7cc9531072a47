"""Process-level wrapper of the controller (API parity with reference
``controller/controller_instance.py:9-51``): start from a ``ControllerParams`` message, then block
until SIGINT/SIGTERM or a ``ShutDown`` RPC."""
import signal
import threading
import time

from metisfl_b200.controller import controller
from metisfl_b200.proto.metis_pb2 import ControllerParams


class ControllerInstance(object):

    def __init__(self):
        self._wrapper = controller.ControllerWrapper()
        self._shutdown_signal = False

    @property
    def servicer(self):
        return self._wrapper.servicer

    def start(self, controller_params_pb):
        assert isinstance(controller_params_pb, ControllerParams)
        self._wrapper.start(controller_params_pb.SerializeToString())

    def shutdown(self, instantly=False):
        def _on_signal(signum, frame):
            self._shutdown_signal = True

        if threading.current_thread() is threading.main_thread():
            signal.signal(signal.SIGTERM, _on_signal)
            signal.signal(signal.SIGINT, _on_signal)
        while True:
            if instantly or self._shutdown_signal:
                self._wrapper.shutdown()
                break
            if self._wrapper.shutdown_request_received():
                self._wrapper.wait()
                break
            time.sleep(0.01)
