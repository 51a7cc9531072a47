"""Keras model backend.

Behavioural spec = the reference's ``KerasModelOps`` (``models/keras/keras_model_ops.py:22-283``): load / save a
SavedModel directory, export ALL ``layer.weights`` (trainable flag per variable), ``set_weights`` from the wire,
``train_model`` = ``model.fit`` bounded by a step budget (``StepCounter``) with per-batch / per-epoch profiling
(``PerformanceProfiler``), then evaluation on the three splits, ``construct_optimizer`` mutating the compiled
optimizer's hyper-parameters from ``OptimizerConfig`` (:245-283), FedProx as an optimizer (``fed_prox.py:50-60``).

TensorFlow is not part of this image, so nothing here imports it at module import time: the class works on any
object with the Keras ``Model`` surface (``fit`` / ``evaluate`` / ``predict`` / ``get_weights`` / ``set_weights`` /
``weights`` / ``optimizer``), and only ``load_model`` / ``save_model`` and the GPU memory-growth switch touch
``tensorflow`` itself.  The CPU test tier drives it with a small Keras-shaped model (``tests/test_keras_backend.py``);
with TensorFlow installed the same code drives real ``tf.keras`` models.  Differences from the reference, on purpose:
the step budget stops at ``total_steps`` (quirk Q11: the reference trains one batch more) and the evaluation metrics
are returned even when a split is empty (as ``{}``).
"""
from __future__ import annotations

import math
import os
from typing import List

import numpy as np

from metisfl_b200.models.callbacks import PerformanceProfiler, StepCounter
from metisfl_b200.models.model_dataset import ModelDataset
from metisfl_b200.models.model_ops import ModelOps
from metisfl_b200.models.model_proto_factory import ModelProtoFactory
from metisfl_b200.proto import metis_pb2, model_pb2
from metisfl_b200.utils.metis_logger import MetisLogger


def _tf():
    try:
        import tensorflow as tf
        return tf
    except ImportError as e:  # pragma: no cover - the image has no TensorFlow
        raise ImportError("this operation needs TensorFlow (load / save of a Keras SavedModel); the PyTorch backend "
                          "(metisfl_b200.models.torch_model_ops.TorchModelOps) needs nothing extra") from e


class _KerasCallbackAdapter:
    """Lets the framework-neutral callbacks of ``models/callbacks.py`` ride on ``model.fit(callbacks=[...])``: Keras only
    needs the hook names and a ``set_model`` / ``set_params`` pair; ``stop_training`` is forwarded to the model."""

    def __init__(self, inner):
        self.inner = inner
        self.model = None
        self.params = None

    def set_model(self, model):
        self.model = model

    def set_params(self, params):
        self.params = params

    def _forward(self, name, *args):
        fn = getattr(self.inner, name, None)
        if fn is not None:
            fn(*args)
        if getattr(self.inner, "stop_training", False) and self.model is not None:
            self.model.stop_training = True

    def on_train_begin(self, logs=None): self._forward("on_train_begin")
    def on_train_end(self, logs=None): self._forward("on_train_end")
    def on_epoch_begin(self, epoch, logs=None): self._forward("on_epoch_begin", epoch)
    def on_epoch_end(self, epoch, logs=None): self._forward("on_epoch_end", epoch)
    def on_train_batch_begin(self, batch, logs=None): self._forward("on_train_batch_begin", batch)
    def on_train_batch_end(self, batch, logs=None): self._forward("on_train_batch_end", batch)

    def __getattr__(self, name):  # every other Keras hook (test / predict batches ...) is a no-op
        if name.startswith("on_") or name.startswith("_implements"):
            return lambda *a, **k: False if name.startswith("_implements") else None
        raise AttributeError(name)


class KerasModelOps(ModelOps):

    def __init__(self, model_dir: str = None, he_scheme=None, model=None, keras_callbacks=None, *args, **kwargs):
        self._model_dir = model_dir
        self._keras_callbacks = list(keras_callbacks or [])
        if model is None:
            self._set_gpu_memory_growth()
            model = self.load_model(model_dir)
        super().__init__(model, he_scheme)

    @staticmethod
    def _set_gpu_memory_growth():  # keras_model_ops.py:22-32
        tf = _tf()
        for gpu in tf.config.list_physical_devices("GPU"):
            try:
                tf.config.experimental.set_memory_growth(gpu, True)
            except Exception as e:  # devices already initialised
                MetisLogger.warning("memory growth not set: %r", e)

    # -- persistence ----------------------------------------------------------------------------------------------
    def load_model(self, model_dir=None, *args, **kwargs):
        tf = _tf()
        d = model_dir or self._model_dir
        custom = {}
        try:
            from metisfl_b200.optim.fed_prox import FedProx
            custom["FedProx"] = FedProx
        except Exception:
            pass
        self._model = tf.keras.models.load_model(d, custom_objects=custom)
        return self._model

    def save_model(self, model_dir=None, *args, **kwargs):
        d = model_dir or self._model_dir
        os.makedirs(d, exist_ok=True)
        self._model.save(d)

    # -- weights ----------------------------------------------------------------------------------------------------
    def get_model_weights(self, *args, **kwargs):
        """All variables of all layers, in ``model.weights`` order (what ``set_weights`` takes back)."""
        names, trainables = [], []
        for v in self._model.weights:
            names.append(getattr(v, "name", "var_{}".format(len(names))))
            trainables.append(bool(getattr(v, "trainable", True)))
        values = [np.asarray(a) for a in self._model.get_weights()]
        return names, trainables, values

    def set_model_weights(self, weights_names: List[str], weights_trainable: List[bool], weights_values, *args,
                          **kwargs):
        current = self._model.get_weights()
        if len(weights_values) != len(current):
            raise ValueError("the incoming model has {} variables, the Keras model {}".format(
                len(weights_values), len(current)))
        cast = []
        for new, old in zip(weights_values, current):
            a = new.detach().cpu().float().numpy() if hasattr(new, "detach") else np.asarray(new)
            cast.append(a.astype(np.asarray(old).dtype).reshape(np.asarray(old).shape))
        self._model.set_weights(cast)

    # -- optimizer ----------------------------------------------------------------------------------------------------
    def construct_optimizer(self, optimizer_config_pb: model_pb2.OptimizerConfig = None, *args, **kwargs):
        """Overwrite the compiled optimizer's hyper-parameters with the federation's (keras_model_ops.py:245-283)."""
        opt = getattr(self._model, "optimizer", None)
        if optimizer_config_pb is None or opt is None:
            return opt
        which = optimizer_config_pb.WhichOneof("config")
        cfg = getattr(optimizer_config_pb, which) if which else None
        table = {"vanilla_sgd": {"learning_rate": "learning_rate"},
                 "momentum_sgd": {"learning_rate": "learning_rate", "momentum_factor": "momentum"},
                 "fed_prox": {"learning_rate": "learning_rate", "proximal_term": "mu"},
                 "adam": {"learning_rate": "learning_rate", "beta_1": "beta_1", "beta_2": "beta_2", "epsilon": "epsilon"},
                 "adam_weight_decay": {"learning_rate": "learning_rate", "weight_decay": "weight_decay"}}
        for field, attr in table.get(which, {}).items():
            value = getattr(cfg, field)
            if not value and field != "learning_rate":
                continue  # proto3 default: keep what the model was compiled with
            target = getattr(opt, attr, None)
            if hasattr(target, "assign"):
                target.assign(value)
            elif hasattr(opt, attr) or attr == "learning_rate":
                setattr(opt, attr, value)
        return opt

    # -- training -----------------------------------------------------------------------------------------------------
    def train_model(self, train_dataset: ModelDataset, learning_task_pb: metis_pb2.LearningTask,
                    hyperparameters_pb: metis_pb2.Hyperparameters, validation_dataset: ModelDataset = None,
                    test_dataset: ModelDataset = None, verbose=False, *args, **kwargs):
        if train_dataset is None:
            raise RuntimeError("Provided `dataset` for training is None.")
        self.construct_optimizer(hyperparameters_pb.optimizer)
        batch_size = hyperparameters_pb.batch_size or 32
        total_steps = int(learning_task_pb.num_local_updates)
        n = train_dataset.get_size()
        steps_per_epoch = max(1, math.ceil(n / batch_size))
        # steps -> epochs as the reference does (keras_model_ops.py:132-136); the StepCounter cuts the last epoch short
        epochs = math.ceil(total_steps / steps_per_epoch) if total_steps > steps_per_epoch else 1
        counter, profiler = StepCounter(total_steps), PerformanceProfiler(device="cpu")
        callbacks = [_KerasCallbackAdapter(counter), _KerasCallbackAdapter(profiler)] + self._keras_callbacks
        x, y = train_dataset.get_x(), train_dataset.get_y()
        fit_kwargs = dict(epochs=epochs, batch_size=batch_size, verbose=int(bool(verbose)), callbacks=callbacks)
        if validation_dataset is not None and validation_dataset.get_size():
            fit_kwargs["validation_data"] = (validation_dataset.get_x(), validation_dataset.get_y())
        history = self._model.fit(x, y, **fit_kwargs)
        hist = dict(getattr(history, "history", {}) or {})
        train_stats = {k: list(v) for k, v in hist.items() if not k.startswith("val_")}
        valid_stats = {k[4:]: list(v) for k, v in hist.items() if k.startswith("val_")}
        test_stats = {}
        if test_dataset is not None and test_dataset.get_size():
            test_stats = {k: [v] for k, v in self.evaluate_model(test_dataset, batch_size).items()}
        steps_done = counter.steps_counter or total_steps
        names, trainables, values = self.get_model_weights()
        bt, et = profiler.batches_wall_clock_time_sec, profiler.epochs_wall_clock_time_sec
        ms_per_batch = 1e3 * float(np.mean(bt)) if bt else 0.0
        ms_per_epoch = 1e3 * float(np.mean(et)) if et else ms_per_batch * steps_per_epoch
        msg = ModelProtoFactory.CompletedLearningTaskProtoMessage(
            weights_values=values, weights_trainable=trainables, weights_names=names, train_stats=train_stats,
            completed_epochs=steps_done / steps_per_epoch, global_iteration=learning_task_pb.global_iteration,
            validation_stats=valid_stats, test_stats=test_stats, completes_batches=steps_done, batch_size=batch_size,
            processing_ms_per_epoch=ms_per_epoch, processing_ms_per_batch=ms_per_batch)
        if self._model_dir:
            try:
                self.save_model(self._model_dir)  # the reference re-saves after every task (:180)
            except ImportError:
                pass
        return msg.construct_completed_learning_task_pb(he_scheme=self._he_scheme)

    # -- evaluation / inference -------------------------------------------------------------------------------------
    def evaluate_model(self, eval_dataset: ModelDataset, batch_size=100, metrics=None, verbose=False, *args,
                       **kwargs) -> dict:
        if eval_dataset is None or not eval_dataset.get_size():
            return {}
        res = self._model.evaluate(eval_dataset.get_x(), eval_dataset.get_y(), batch_size=batch_size or 100,
                                   verbose=int(bool(verbose)), return_dict=True)
        res = {k: float(v) for k, v in dict(res).items()}
        if metrics:
            res = {k: v for k, v in res.items() if k in metrics or k == "loss"} or res
        return res

    def infer_model(self, infer_dataset: ModelDataset, batch_size=100, *args, **kwargs):
        if infer_dataset is None:
            raise RuntimeError("Provided `dataset` for inference is None.")
        return np.asarray(self._model.predict(infer_dataset.get_x(), batch_size=batch_size or 100, verbose=0))
