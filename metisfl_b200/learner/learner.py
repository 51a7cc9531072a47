"""Learner: joins the federation, runs learning / evaluation tasks, reports results.

Behavioural spec = reference ``metisfl/learner/learner.py:21-417`` (credentials persisted to
``learner_id.txt`` / ``auth_token.txt`` :96-103; dataset recipes are cloudpickled callables
:134-182; ``run_learning_task`` cancels a running task, schedules the new one and reports with
``MarkTaskCompleted`` when it finishes :197-206,376-396; ``run_evaluation_task`` blocks :356-371;
``shutdown`` :398-417).

B200 design difference: the reference isolates every task in a *spawned process* (three pebble
pools) and re-creates the model backend and datasets per task.  With a GPU that would pay CUDA
context creation, cuDNN autotuning and CUDA-graph capture on every round, so here the model backend
lives for the life of the learner and tasks run on two long-lived worker threads (train / eval);
pre-emption is cooperative (a cancel flag checked between optimizer steps).
"""
from __future__ import annotations

import concurrent.futures
import inspect
import os
import threading

import cloudpickle

from metisfl_b200.learner.learner_evaluator import LearnerEvaluator
from metisfl_b200.learner.learner_trainer import LearnerTrainer
from metisfl_b200.models.model_dataset import (ModelDataset, ModelDatasetClassification, ModelDatasetRegression)
from metisfl_b200.models.model_proto_factory import ModelProtoFactory
from metisfl_b200.proto import learner_pb2, metis_pb2, model_pb2
from metisfl_b200.utils.grpc_controller_client import GRPCControllerClient
from metisfl_b200.utils.metis_logger import MetisLogger
from metisfl_b200.utils.proto_messages_factory import MetisProtoMessages


class Learner(object):

    def __init__(self, learner_server_entity: metis_pb2.ServerEntity,
                 controller_server_entity: metis_pb2.ServerEntity, he_scheme_config_pb: metis_pb2.HESchemeConfig,
                 nn_engine, model_dir, train_dataset_fp, train_dataset_recipe_pkl, validation_dataset_fp="",
                 validation_dataset_recipe_pkl="", test_dataset_fp="", test_dataset_recipe_pkl="",
                 recreate_queue_task_worker=False, learner_credentials_fp="/tmp/metis/learner/", device=None,
                 model_ops=None, fabric=None):
        self.learner_server_entity = learner_server_entity
        self._controller_server_entity = controller_server_entity
        self._he_scheme_config_pb = he_scheme_config_pb
        self._nn_engine = nn_engine
        self._model_dir = model_dir
        self._device = device
        if not train_dataset_recipe_pkl:
            raise AssertionError("Training dataset recipe is required.")
        self.train_dataset_recipe_pkl, self.train_dataset_fp = train_dataset_recipe_pkl, train_dataset_fp
        self.validation_dataset_recipe_pkl, self.validation_dataset_fp = \
            validation_dataset_recipe_pkl, validation_dataset_fp
        self.test_dataset_recipe_pkl, self.test_dataset_fp = test_dataset_recipe_pkl, test_dataset_fp
        self._train_pool = concurrent.futures.ThreadPoolExecutor(max_workers=1, thread_name_prefix="learner-train")
        self._eval_pool = concurrent.futures.ThreadPoolExecutor(max_workers=1, thread_name_prefix="learner-eval")
        self._train_future = None
        self._cancel_event = threading.Event()
        self._model_lock = threading.Lock()  # one task at a time touches the model backend
        self._model_ops = model_ops
        # a FabricSpec (or its text form) makes this learner a rank of the on-box NVLink fabric: models stay in HBM
        # and the messages carry device-slot references (learner/device_backend.py)
        self._fabric = fabric
        self._device_backend = None
        if fabric is not None:
            from metisfl_b200.learner.device_backend import DeviceBackend, FabricSpec
            spec = FabricSpec.parse(fabric) if isinstance(fabric, str) else fabric
            # collective: every rank of the fabric constructs its backend before any of them serves RPCs
            cert = b""
            ssl = learner_server_entity.ssl_config
            if ssl.enable_ssl:
                from metisfl_b200.utils.ssl_configurator import SSLConfigurator
                pub = SSLConfigurator.gen_public_ssl_config_pb_as_stream(ssl)
                cert = pub.ssl_config_stream.public_certificate_stream
            self._device_backend = DeviceBackend(
                model_dir, device, spec, certificate=cert, he_scheme_config_pb=he_scheme_config_pb,
                endpoint="{}:{}".format(learner_server_entity.hostname, learner_server_entity.port))
        self._datasets = None
        self._holdout_done = False
        self._he_scheme = None
        self._learner_controller_client = GRPCControllerClient(self._controller_server_entity, max_workers=1)
        self._credentials_dir = os.path.join(
            learner_credentials_fp, "{}_{}".format(learner_server_entity.hostname, learner_server_entity.port))
        os.makedirs(self._credentials_dir, exist_ok=True)
        self._learner_id = None
        self._auth_token = None
        self._learner_id_fp = os.path.join(self._credentials_dir, "learner_id.txt")
        self._auth_token_fp = os.path.join(self._credentials_dir, "auth_token.txt")

    # -- datasets ---------------------------------------------------------------------------------
    @staticmethod
    def _load_recipe(recipe):
        if callable(recipe):
            return recipe
        with open(recipe, "rb") as f:
            return cloudpickle.load(f)

    def _create_model_dataset_helper(self, dataset_recipe_pkl, dataset_fp=None, default_class=None):
        if not dataset_recipe_pkl and not default_class:
            raise RuntimeError("Neither the dataset recipe or the default class are specified. Exiting ...")
        if dataset_recipe_pkl:
            recipe_fn = self._load_recipe(dataset_recipe_pkl)
            if len(inspect.signature(recipe_fn).parameters) > 0:
                dataset = recipe_fn(dataset_fp) if dataset_fp else default_class()
            else:
                dataset = recipe_fn()
        else:
            dataset = default_class()
        assert isinstance(dataset, ModelDataset), \
            "The dataset needs to be an instance of: {}".format(ModelDataset.__name__)
        return dataset

    def _load_model_datasets(self):
        if self._datasets is None:  # recipes run once per learner, not once per task
            train = self._create_model_dataset_helper(self.train_dataset_recipe_pkl, self.train_dataset_fp)
            valid = self._create_model_dataset_helper(self.validation_dataset_recipe_pkl, self.validation_dataset_fp,
                                                      default_class=train.__class__)
            test = self._create_model_dataset_helper(self.test_dataset_recipe_pkl, self.test_dataset_fp,
                                                     default_class=train.__class__)
            self._datasets = (train, valid, test)
        return self._datasets

    def _apply_validation_holdout(self, pct: float):
        """``LearningTask.training_dataset_percentage_for_stratified_validation`` (``ValidationPercentage`` in the
        YAML): a learner without a validation set of its own holds out that fraction of its training examples -
        per class for classification labels - trains on the rest and reports the hold-out as its validation split.
        The reference plumbs the field down to the task (``controller.cc:730-731``) and never reads it (quirk Q7).
        Done once per learner; array datasets only (a ``DataLoader`` has no indexable examples to split)."""
        pct = float(pct)
        if 1.0 <= pct < 100.0:   # "10" is read as 10 %
            pct /= 100.0
        if self._holdout_done or not (0.0 < pct < 1.0):
            return
        self._holdout_done = True
        import numpy as np
        train, valid, test = self._load_model_datasets()
        x, y = train.get_x(), train.get_y()
        if valid.get_size() > 0 or not hasattr(x, "shape") or not hasattr(y, "shape") or len(x) != len(y) or len(y) < 2:
            return
        labels = np.asarray(y.detach().cpu() if hasattr(y, "detach") else y)
        rng = np.random.default_rng(0)
        is_cls = isinstance(train, ModelDatasetClassification) and labels.ndim == 1 and labels.dtype.kind in "iub"
        groups = [np.flatnonzero(labels == c) for c in np.unique(labels)] if is_cls else [np.arange(len(labels))]
        held = []
        for g in groups:
            k = int(round(float(pct) * len(g)))
            if len(groups) == 1:
                k = min(max(k, 1), len(g) - 1)
            held.append(rng.permutation(g)[:min(k, len(g) - 1)] if len(g) > 1 else g[:0])
        held = np.sort(np.concatenate(held))
        if held.size == 0:
            return
        keep = np.setdiff1d(np.arange(len(labels)), held)

        def part(idx):
            xs, ys = x[idx], y[idx]
            if is_cls:
                cls, cnt = np.unique(labels[idx], return_counts=True)
                return ModelDatasetClassification(x=xs, y=ys, size=len(idx),
                                                  examples_per_class={int(c): int(n) for c, n in zip(cls, cnt)})
            if isinstance(train, ModelDatasetRegression):
                return ModelDatasetRegression(x=xs, y=ys, size=len(idx))
            return ModelDataset(x=xs, y=ys, size=len(idx))
        self._datasets = (part(keep), part(held), test)
        MetisLogger.info("held out %d of %d training examples for validation", held.size, len(labels))

    def _load_model_datasets_size_specs_type_def(self):
        return [(d.get_size(), d.get_model_dataset_specifications(), type(d)) for d in self._load_model_datasets()]

    # -- model backend ----------------------------------------------------------------------------
    def _he_scheme_factory(self):
        cfg = self._he_scheme_config_pb
        if cfg is None or not cfg.enabled or cfg.WhichOneof("config") != "ckks_scheme_config":
            return None
        if self._he_scheme is None:
            from metisfl_b200.encryption import fhe
            scheme = fhe.CKKS(cfg.ckks_scheme_config.batch_size, cfg.ckks_scheme_config.scaling_factor_bits)
            scheme.load_crypto_context_from_file(cfg.crypto_context_file)
            scheme.load_public_key_from_file(cfg.public_key_file)
            scheme.load_private_key_from_file(cfg.private_key_file)
            self._he_scheme = scheme
        return self._he_scheme

    def _model_ops_factory(self, nn_engine=None):
        if self._model_ops is None:
            engine = (nn_engine or self._nn_engine or "pytorch").lower()
            if engine in ("keras", "tensorflow", "tf"):  # needs TensorFlow at load time (SavedModel directory)
                from metisfl_b200.models.keras.keras_model_ops import KerasModelOps
                self._model_ops = KerasModelOps(self._model_dir, he_scheme=self._he_scheme_factory())
                return self._model_ops
            if engine not in ("pytorch", "torch"):
                raise RuntimeError("Unknown neural engine '{}' (pytorch | keras | echo).".format(engine))
            from metisfl_b200.models.torch_model_ops import TorchModelOps
            self._model_ops = TorchModelOps(self._model_dir, he_scheme=self._he_scheme_factory(), device=self._device)
            self._model_ops.wire_tasks = True   # completed tasks leave as bytes built straight from the tensors
        return self._model_ops

    # -- federation membership --------------------------------------------------------------------
    def host_port_identifier(self):
        return "{}:{}".format(self.learner_server_entity.hostname, self.learner_server_entity.port)

    def join_federation(self):
        specs = self._load_model_datasets_size_specs_type_def()
        (n_tr, s_tr, cls_tr), (n_va, s_va, _), (n_te, s_te, _) = specs
        is_cls = issubclass(cls_tr, ModelDatasetClassification)
        is_reg = issubclass(cls_tr, ModelDatasetRegression)
        self._learner_id, self._auth_token, status = self._learner_controller_client.join_federation(
            self.learner_server_entity, self._learner_id_fp, self._auth_token_fp, n_tr, s_tr, n_va, s_va, n_te, s_te,
            is_cls, is_reg, request_retries=10, request_timeout=30, block=True,
            fabric_pb=self._device_backend.membership_pb() if self._device_backend is not None else None)
        return status

    def leave_federation(self):
        if self._learner_id is None:
            return False
        return self._learner_controller_client.leave_federation(self._learner_id, self._auth_token, block=True)

    # -- tasks (executed on the worker threads) ---------------------------------------------------
    def model_evaluate(self, model_pb: model_pb2.Model, batch_size: int, evaluation_datasets_pb, metrics_pb,
                       verbose=False):
        train, valid, test = self._load_model_datasets()
        E = learner_pb2.EvaluateModelRequest
        if self._device_backend is not None:
            want = {code: (ds if code in evaluation_datasets_pb else None)
                    for code, ds in ((E.TRAINING, train), (E.VALIDATION, valid), (E.TEST, test))}
            res = self._device_backend.evaluate(model_pb, want, batch_size, list(metrics_pb.metric))
            mk = lambda d: ModelProtoFactory.ModelEvaluationProtoMessage(d).construct_model_evaluation_pb()
            return MetisProtoMessages.construct_model_evaluations_pb(
                training_evaluation_pb=mk(res[E.TRAINING]), validation_evaluation_pb=mk(res[E.VALIDATION]),
                test_evaluation_pb=mk(res[E.TEST]))
        with self._model_lock:
            with LearnerEvaluator(self._model_ops_factory(), model_pb) as ev:
                res = {}
                for code, ds in ((E.TRAINING, train), (E.VALIDATION, valid), (E.TEST, test)):
                    res[code] = ev.evaluate_model(ds, batch_size, metrics_pb, verbose) \
                        if code in evaluation_datasets_pb else {}
        mk = lambda d: ModelProtoFactory.ModelEvaluationProtoMessage(d).construct_model_evaluation_pb()
        return MetisProtoMessages.construct_model_evaluations_pb(
            training_evaluation_pb=mk(res[E.TRAINING]), validation_evaluation_pb=mk(res[E.VALIDATION]),
            test_evaluation_pb=mk(res[E.TEST]))

    def model_infer(self, model_pb: model_pb2.Model, batch_size: int, infer_train=False, infer_test=False,
                    infer_valid=False, verbose=False):
        train, valid, test = self._load_model_datasets()
        out = {}
        with self._model_lock:
            with LearnerEvaluator(self._model_ops_factory(), model_pb) as ev:
                for flag, name, ds in ((infer_train, "train", train), (infer_valid, "valid", valid),
                                       (infer_test, "test", test)):
                    if flag:
                        out[name] = ev.infer_model(ds, batch_size, verbose)
        return out

    def model_train(self, learning_task_pb: metis_pb2.LearningTask, hyperparameters_pb: metis_pb2.Hyperparameters,
                    model_pb: model_pb2.Model, verbose=False, cancel_event=None):
        self._apply_validation_holdout(learning_task_pb.training_dataset_percentage_for_stratified_validation)
        train, valid, test = self._load_model_datasets()
        if self._device_backend is not None:
            with self._model_lock:
                return self._device_backend.train(train, learning_task_pb, hyperparameters_pb, model_pb, cancel_event)
        with self._model_lock:
            with LearnerTrainer(self._model_ops_factory(), model_pb) as tr:
                return tr.train_model(train, learning_task_pb, hyperparameters_pb, valid, test, verbose,
                                      cancel_event=cancel_event)

    def _mark_learning_task_completed(self, training_future):
        if training_future.cancelled():
            return
        exc = training_future.exception()
        if exc is not None:
            MetisLogger.error("learning task failed: %r", exc)
            return
        completed_task_pb = training_future.result()
        if completed_task_pb is None:  # pre-empted
            return
        self._learner_controller_client.mark_task_completed(
            self._learner_id, self._auth_token, completed_task_pb, request_retries=3, request_timeout=600,
            block=False)

    def run_evaluation_task(self, model_pb: model_pb2.Model, batch_size: int, evaluation_dataset_pb, metrics_pb,
                            cancel_running_tasks=False, block=False, verbose=False):
        fut = self._eval_pool.submit(self.model_evaluate, model_pb, batch_size, evaluation_dataset_pb, metrics_pb,
                                     verbose)
        return fut.result() if block else fut

    def aggregate_on_device(self, request_pb):
        """Controller-commanded aggregation over device slots (``LearnerService.AggregateOnDevice``)."""
        if self._device_backend is None:
            raise RuntimeError("this learner does not host a fabric rank")
        return self._device_backend.aggregate(request_pb)

    def run_inference_task(self):
        raise NotImplementedError("Not yet implemented.")

    def run_learning_task(self, learning_task_pb: metis_pb2.LearningTask,
                          hyperparameters_pb: metis_pb2.Hyperparameters, model_pb: model_pb2.Model,
                          cancel_running_tasks=False, block=False, verbose=False):
        if cancel_running_tasks and self._train_future is not None and not self._train_future.done():
            self._cancel_event.set()  # cooperative pre-emption of the running task
            if not self._train_future.cancel():
                try:
                    self._train_future.result(timeout=600)
                except Exception:
                    pass
        self._cancel_event = threading.Event()
        cancel = self._cancel_event

        def _task():
            res = self.model_train(learning_task_pb, hyperparameters_pb, model_pb, verbose, cancel)
            return None if cancel.is_set() else res

        fut = self._train_pool.submit(_task)
        fut.add_done_callback(self._mark_learning_task_completed)
        self._train_future = fut
        if block:
            fut.result()
            return True
        return not fut.cancelled()

    def shutdown(self, graceful=False, cancel_train_running_tasks=True, cancel_eval_running_tasks=True,
                 cancel_infer_running_tasks=True):
        if cancel_train_running_tasks:
            self._cancel_event.set()
        self._train_pool.shutdown(wait=True, cancel_futures=cancel_train_running_tasks)
        self._eval_pool.shutdown(wait=True, cancel_futures=cancel_eval_running_tasks and not graceful)
        return True

    def close(self):
        """Release the controller channel (after ``leave_federation``)."""
        self._learner_controller_client.shutdown()
        if self._device_backend is not None:
            self._device_backend.close()
