"""Training-loop callbacks: ``StepCounter`` and ``PerformanceProfiler``.

Counterparts of the reference's Keras callbacks (``models/keras/callbacks/step_counter.py:6-56``,
``performance_profiler.py:5-63``), attached to the framework-owned PyTorch loop through
``TorchModelOps.train_model(..., callbacks=[...])``.  ``StepCounter`` stops after exactly ``total_steps`` batches
(the reference stops one batch late).  ``PerformanceProfiler`` records host wall-clock like the
reference and, on CUDA, device time per epoch from events (a per-batch synchronize would serialise the stream).
"""
import time

import torch


class Callback(object):
    stop_training = False

    def on_train_begin(self, logs=None): pass
    def on_train_end(self, logs=None): pass
    def on_epoch_begin(self, epoch, logs=None): pass
    def on_epoch_end(self, epoch, logs=None): pass
    def on_train_batch_begin(self, batch, logs=None): pass
    def on_train_batch_end(self, batch, logs=None): pass


class StepCounter(Callback):
    def __init__(self, total_steps):
        self.total_steps = int(total_steps)
        self.steps_counter = 0
        self.epochs_counter = 0

    def on_epoch_end(self, epoch, logs=None):
        self.epochs_counter += 1

    def on_train_batch_end(self, batch, logs=None):
        self.steps_counter += 1
        if self.steps_counter >= self.total_steps:
            self.stop_training = True


class PerformanceProfiler(Callback):
    def __init__(self, device=None):
        self.epochs_wall_clock_time_sec = []
        self.batches_wall_clock_time_sec = []
        self.epochs_device_time_ms = []
        self._cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        self._epoch_t0 = self._batch_t0 = None
        self._events = []

    def on_epoch_begin(self, epoch, logs=None):
        self._epoch_t0 = time.time()
        if self._cuda:
            e0 = torch.cuda.Event(enable_timing=True); e0.record()
            self._events.append([e0, None])

    def on_epoch_end(self, epoch, logs=None):
        if self._cuda:
            e1 = torch.cuda.Event(enable_timing=True); e1.record()
            self._events[-1][1] = e1
        self.epochs_wall_clock_time_sec.append(time.time() - self._epoch_t0)

    def on_train_batch_begin(self, batch, logs=None):
        self._batch_t0 = time.time()

    def on_train_batch_end(self, batch, logs=None):
        self.batches_wall_clock_time_sec.append(time.time() - self._batch_t0)

    def on_train_end(self, logs=None):
        if self._cuda:
            torch.cuda.synchronize()
            self.epochs_device_time_ms = [e0.elapsed_time(e1) for e0, e1 in self._events if e1 is not None]
