"""Tracing hooks: NVTX ranges (visible in nsys / ncu timelines) and device timers.

The reference has no tracer; it records wall-clock runtime metadata in the controller.
The native core keeps those metadata fields; on the GPU side rounds are additionally timed with CUDA
events (never wall clock) and bracketed by NVTX ranges."""
import contextlib

import torch


@contextlib.contextmanager
def nvtx_range(name: str):
    pushed = False
    if torch.cuda.is_available():
        try:
            torch.cuda.nvtx.range_push(name)
            pushed = True
        except Exception:
            pushed = False
    try:
        yield
    finally:
        if pushed:
            torch.cuda.nvtx.range_pop()


class DeviceTimer:
    """CUDA-event stopwatch for named phases; ``report()`` synchronises once and returns milliseconds."""

    def __init__(self, device):
        self.enabled = torch.device(device).type == "cuda"
        self._pending = []
        self.totals = {}

    @contextlib.contextmanager
    def phase(self, name: str):
        if not self.enabled:
            yield
            return
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        with nvtx_range(name):
            yield
        e1.record()
        self._pending.append((name, e0, e1))

    def report(self, reset=True):
        for name, e0, e1 in self._pending:
            e1.synchronize()
            self.totals[name] = self.totals.get(name, 0.0) + e0.elapsed_time(e1)
        self._pending.clear()
        out = dict(self.totals)
        if reset:
            self.totals = {}
        return out
