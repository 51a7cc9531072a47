"""PyTorch model backend that owns the training loop.

Replaces the reference's ``PyTorchModelOps`` (``models/pytorch/pytorch_model_ops.py:22-172``), fixing
quirk Q10: the full ``state_dict`` (parameters AND buffers) is exchanged, the model is placed on a
device, the optimizer in ``Hyperparameters`` is honoured (fused K2 kernel), and batch / epoch timing
is reported so ``NUM_COMPLETED_BATCHES`` scaling and semi-synchronous task sizing work.  Saved-model
layout: ``model_def.pkl`` (cloudpickled module) + ``model_weights.pt`` (state dict), as the
reference (``driver_session.py:56-63``).
"""
from __future__ import annotations

import math
import os
import time
from typing import List, Optional

import cloudpickle
import numpy as np
import torch
import torch.nn.functional as F

from metisfl_b200.models.model_dataset import ModelDataset
from metisfl_b200.models.model_ops import ModelOps
from metisfl_b200.models.model_proto_factory import ModelProtoFactory
from metisfl_b200.models.trainer import LocalTrainer, cross_entropy_loss
from metisfl_b200.ops.optim import OptimHyper
from metisfl_b200.parallel.arena import ModelArena
from metisfl_b200.proto import metis_pb2, model_pb2
from metisfl_b200.utils.metis_logger import MetisLogger

MODEL_DEF_FILE = "model_def.pkl"
MODEL_WEIGHTS_FILE = "model_weights.pt"


def save_torch_model(module: torch.nn.Module, model_dir: str):
    os.makedirs(model_dir, exist_ok=True)
    with open(os.path.join(model_dir, MODEL_DEF_FILE), "wb") as f:
        cloudpickle.dump(module, f)
    torch.save(module.state_dict(), os.path.join(model_dir, MODEL_WEIGHTS_FILE))


def load_torch_model(model_dir: str) -> torch.nn.Module:
    with open(os.path.join(model_dir, MODEL_DEF_FILE), "rb") as f:
        module = cloudpickle.load(f)
    weights = os.path.join(model_dir, MODEL_WEIGHTS_FILE)
    if os.path.exists(weights):
        module.load_state_dict(torch.load(weights, map_location="cpu"))
    return module


def optim_hyper_from_pb(cfg: Optional[model_pb2.OptimizerConfig]) -> OptimHyper:
    """``metisfl.OptimizerConfig`` oneof -> kernel hyper-parameters (``model.proto:116-152``)."""
    which = cfg.WhichOneof("config") if cfg is not None else None
    if which in (None, "vanilla_sgd"):
        o = cfg.vanilla_sgd if which else model_pb2.VanillaSGD(learning_rate=0.01)
        return OptimHyper("VanillaSGD", lr=o.learning_rate, l1=o.L1_reg, l2=o.L2_reg)
    if which == "momentum_sgd":
        return OptimHyper("MomentumSGD", lr=cfg.momentum_sgd.learning_rate, momentum=cfg.momentum_sgd.momentum_factor)
    if which == "fed_prox":
        return OptimHyper("FedProx", lr=cfg.fed_prox.learning_rate, mu=cfg.fed_prox.proximal_term)
    if which == "adam":
        a = cfg.adam
        return OptimHyper("Adam", lr=a.learning_rate, beta1=a.beta_1 or 0.9, beta2=a.beta_2 or 0.999,
                          eps=a.epsilon or 1e-7)
    a = cfg.adam_weight_decay
    return OptimHyper("AdamW", lr=a.learning_rate, weight_decay=a.weight_decay)


def _as_tensor(a):
    if a is None:
        return None
    if isinstance(a, torch.Tensor):
        return a
    return torch.from_numpy(np.ascontiguousarray(a))


class _Batches:
    """Uniform batch iterator over (x, y) arrays/tensors or a DataLoader; pins host memory for CUDA.

    ``cache`` (a dict owned by the caller, one per learner) keeps the pinned copy of an array dataset and a pinned
    gather buffer across tasks, so a task costs one host-side gather instead of a fresh pin + pageable H2D copies."""

    def __init__(self, dataset: ModelDataset, batch_size: int, shuffle: bool, pin: bool, seed: int = 0, cache=None,
                 x_dtype: Optional[torch.dtype] = None, channels_last: bool = False):
        """``x_dtype`` / ``channels_last`` (only with ``cache``): keep the cached copy of a floating-point ``x`` in the
        dtype and memory layout the device wants (e.g. bf16 NHWC), converted ONCE - a batch is then a plain DMA into
        the step's static input buffer, with no conversion kernel and half the bytes."""
        self.loader = None
        self.gather = None
        self.nhwc_shape = None
        x = dataset.get_x()
        if isinstance(x, torch.utils.data.DataLoader):
            self.loader = x
            self.n = len(x.dataset) if hasattr(x, "dataset") else dataset.get_size()
            self.batch_size = x.batch_size or batch_size
        elif isinstance(x, torch.utils.data.Dataset):
            self.loader = torch.utils.data.DataLoader(x, batch_size=batch_size, shuffle=shuffle, pin_memory=pin)
            self.n = len(x)
            self.batch_size = batch_size
        else:
            pin = bool(pin and torch.cuda.is_available())
            key = (id(dataset), pin, x_dtype, bool(channels_last))
            ent = cache.get(key) if cache is not None else None
            if ent is None:
                tx, ty = _as_tensor(x), _as_tensor(dataset.get_y())
                if ty is not None and ty.dtype in (torch.int32, torch.int16, torch.uint8, torch.int8):
                    ty = ty.long()
                nhwc = None
                if cache is not None and tx.is_floating_point():
                    if x_dtype is not None and tx.dtype != x_dtype:
                        tx = tx.to(x_dtype)
                    if channels_last and tx.dim() == 4:
                        # stored as [N, H*W*C] rows (physically NHWC); batches are handed out as NCHW views of it
                        n_, c_, h_, w_ = tx.shape
                        nhwc = (h_, w_, c_)
                        tx = tx.permute(0, 2, 3, 1).contiguous().view(n_, -1)
                if pin:
                    tx = tx.pin_memory()
                    ty = ty.pin_memory() if ty is not None else None
                ent = {"x": tx, "y": ty, "gx": None, "gy": None, "nhwc": nhwc, "ref": dataset}  # ref: id() stays unique
                if cache is not None:
                    cache[key] = ent
            self.x, self.y, self.nhwc_shape = ent["x"], ent["y"], ent["nhwc"]
            if shuffle and pin and cache is not None:
                if ent["gx"] is None:
                    ent["gx"] = torch.empty_like(self.x).pin_memory()
                    ent["gy"] = torch.empty_like(self.y).pin_memory() if self.y is not None else None
                self.gather = (ent["gx"], ent["gy"])
            self.n = self.x.shape[0]
            self.batch_size = max(1, min(batch_size, self.n))
        self.shuffle = shuffle
        self.gen = torch.Generator().manual_seed(seed)

    @property
    def steps_per_epoch(self) -> int:
        return max(1, math.ceil(self.n / self.batch_size)) if self.n else 0

    def __iter__(self):
        if self.loader is not None:
            yield from self.loader
            return
        if self.shuffle:
            perm = torch.randperm(self.n, generator=self.gen)
            if self.gather is not None:   # shuffled rows land in pinned memory: the H2D copies stay asynchronous
                x = torch.index_select(self.x, 0, perm, out=self.gather[0])
                y = torch.index_select(self.y, 0, perm, out=self.gather[1]) if self.y is not None else None
            else:
                x, y = self.x[perm], self.y[perm] if self.y is not None else None
        else:
            x, y = self.x, self.y
        for i in range(0, self.n, self.batch_size):
            xb = x[i:i + self.batch_size]
            if self.nhwc_shape is not None:   # NHWC rows -> NCHW view with channels_last strides (no copy)
                h_, w_, c_ = self.nhwc_shape
                xb = xb.view(xb.shape[0], h_, w_, c_).permute(0, 3, 1, 2)
            yield xb, (y[i:i + self.batch_size] if y is not None else None)


class TorchModelOps(ModelOps):

    def __init__(self, model_dir: str = "/tmp/metis/model", he_scheme=None, device=None,
                 model_dtype: torch.dtype = None, model: torch.nn.Module = None, loss_fn=None,
                 use_cuda_graph: Optional[bool] = None, *args, **kwargs):
        self._model_dir = model_dir
        module = model if model is not None else load_torch_model(model_dir)
        super().__init__(module, he_scheme)
        if device is None:
            device = "cuda" if torch.cuda.is_available() else "cpu"
        self.device = torch.device(device)
        if model_dtype is None:
            model_dtype = torch.bfloat16 if self.device.type == "cuda" else torch.float32
        self.model_dtype = model_dtype
        if loss_fn is None:
            loss_fn = module.loss_fn if _overrides(module, "loss_fn") else cross_entropy_loss
        self.loss_fn = loss_fn
        if self.device.type == "cuda":
            from metisfl_b200.ops import conv as _conv
            _conv.use_overlapped_wgrad(module)  # weight gradients overlap the dgrad chain on a side stream
        if self.device.type == "cuda" and model_dtype == torch.bfloat16:
            from metisfl_b200.ops import gemm as _gemm
            _gemm.use_b200_linear(module)  # Linear layers on the tcgen05 GEMM (K10)
        self.arena = ModelArena(module, self.device, model_dtype)
        # variable-shaped last batches: the CUDA graph is used only for full batches
        self.trainer = LocalTrainer(self.arena, OptimHyper("VanillaSGD", lr=0.01), loss_fn,
                                    use_cuda_graph=use_cuda_graph)

    # -- persistence ------------------------------------------------------------------------------
    def load_model(self, model_dir=None, *args, **kwargs):
        module = load_torch_model(model_dir or self._model_dir)
        names, trainables, values = [], [], []
        for k, v in module.state_dict().items():
            names.append(k); trainables.append(True); values.append(v)
        self.set_model_weights(names, trainables, values)
        return self._model

    def save_model(self, model_dir=None, *args, **kwargs):
        d = model_dir or self._model_dir
        os.makedirs(d, exist_ok=True)
        torch.save({k: v.detach().float().cpu() if v.is_floating_point() else v.detach().cpu()
                    for k, v in self._model.state_dict().items()}, os.path.join(d, MODEL_WEIGHTS_FILE))

    # -- weights ----------------------------------------------------------------------------------
    def get_model_weights(self, *args, **kwargs):
        # one device->host copy of the whole arena, then per-variable views of the host copy
        host = self.arena.state.detach().to("cpu")
        seg = self.arena.layout.views(host)
        names, trainables, values = [], [], []
        for v in self.arena.layout.variables:
            names.append(v.name)
            trainables.append(v.trainable)
            values.append(self.arena.layout.var_view(seg, v))
        return names, trainables, values

    def set_model_weights(self, weights_names: List[str], weights_trainable: List[bool], weights_values, *args,
                          **kwargs):
        by_name = dict(zip(weights_names, weights_values))
        views = {n: v for n, v, _ in self.arena.named_variables()}
        missing = [n for n in views if n not in by_name]
        if missing and len(weights_values) == len(views):  # positional fallback (foreign naming)
            by_name = dict(zip(views.keys(), weights_values))
            missing = []
        if missing:
            raise KeyError(f"model weights are missing variables: {missing[:5]}")
        with torch.no_grad():
            for n, view in views.items():
                src = _as_tensor(by_name[n])
                view.copy_(src.to(view.device, non_blocking=True).reshape(view.shape).to(view.dtype))
        self.arena.sync_master_from_model()

    def set_model_weights_from_pb(self, model_pb: model_pb2.Model):
        """Wire model -> arena in bulk: one staging copy + one cast/scatter launch per dtype pair (K12)."""
        from metisfl_b200.ops import pack
        if any(v.HasField("ciphertext_tensor") for v in model_pb.variables):
            names, trainables, values = self.get_model_weights_from_variables_pb(model_pb.variables)
            return self.set_model_weights(names, trainables, values)
        pack.unpack_model_into_arena(model_pb, self.arena)

    def get_model_pb(self) -> model_pb2.Model:
        """Arena -> wire model with a single device->host copy."""
        from metisfl_b200.ops import pack
        return pack.arena_to_model_pb(self.arena)

    def construct_optimizer(self, optimizer_config_pb: model_pb2.OptimizerConfig = None, *args, **kwargs):
        h = optim_hyper_from_pb(optimizer_config_pb)
        self.trainer.set_hyper(h)
        return h

    # -- training ---------------------------------------------------------------------------------
    def train_model(self, train_dataset: ModelDataset, learning_task_pb: metis_pb2.LearningTask,
                    hyperparameters_pb: metis_pb2.Hyperparameters, validation_dataset: ModelDataset = None,
                    test_dataset: ModelDataset = None, verbose=False, cancel_event=None, callbacks=None, *args,
                    **kwargs):
        if train_dataset is None:
            raise RuntimeError("Provided `dataset` for training is None.")
        self.construct_optimizer(hyperparameters_pb.optimizer)
        batch_size = hyperparameters_pb.batch_size or 32
        total_steps = int(learning_task_pb.num_local_updates)
        module = self._model
        if _overrides(module, "fit"):  # user-owned loop (reference PyTorchDef contract)
            return self._train_with_user_fit(train_dataset, learning_task_pb, batch_size, total_steps)
        batches = _Batches(train_dataset, batch_size, shuffle=True, pin=self.device.type == "cuda",
                           seed=learning_task_pb.global_iteration)
        spe = batches.steps_per_epoch
        tr = self.trainer
        tr.reset_optimizer_state()
        if tr.hyper.kind == 2:
            tr.snapshot_anchor()
        train_stats, valid_stats, test_stats = {}, {}, {}
        steps_done, epochs_done, epoch_ms = 0, 0, []
        t_task = time.perf_counter()
        cancelled = False
        callbacks = list(callbacks or [])
        for cb in callbacks:
            cb.on_train_begin()
        while steps_done < total_steps and spe > 0 and not cancelled:
            t_epoch = time.perf_counter()
            running, n_run = None, 0
            for cb in callbacks:
                cb.on_epoch_begin(epochs_done)
            for x, y in batches:
                if steps_done >= total_steps or any(cb.stop_training for cb in callbacks):
                    break
                for cb in callbacks:
                    cb.on_train_batch_begin(steps_done)
                if cancel_event is not None and cancel_event.is_set():
                    cancelled = True
                    break
                x = x.to(self.device, non_blocking=True)
                y = y.to(self.device, non_blocking=True)
                if tr.use_graph and x.shape[0] != batches.batch_size:
                    loss = self._eager_step(x, y)  # ragged last batch: outside the captured graph
                else:
                    loss = tr.train_step(x, y)
                running = loss.detach().float().clone() if running is None else running + loss.detach().float()
                n_run += 1
                steps_done += 1
                for cb in callbacks:
                    cb.on_train_batch_end(steps_done - 1)
            for cb in callbacks:
                cb.on_epoch_end(epochs_done)
            if self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
            epoch_ms.append((time.perf_counter() - t_epoch) * 1e3)
            epochs_done += 1
            if any(cb.stop_training for cb in callbacks):
                break
            if n_run:
                train_stats.setdefault("loss", []).append(float(running) / n_run)
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        total_ms = (time.perf_counter() - t_task) * 1e3
        for cb in callbacks:
            cb.on_train_end()
        # end-of-task evaluation of the locally trained model (last epoch entry of each split)
        metrics = list(learning_task_pb.metrics.metric)
        for ds, stats in ((train_dataset, train_stats), (validation_dataset, valid_stats), (test_dataset, test_stats)):
            if ds is not None and ds.get_size() and not cancelled:
                ev = self.evaluate_model(ds, batch_size, metrics)
                for k, v in ev.items():
                    if stats is train_stats and k == "loss":
                        continue
                    stats.setdefault(k, []).append(v)
        # a task that is only going to be sent (the Learner sets ``wire_tasks``) is serialized straight out of the arena
        wire = getattr(self, "wire_tasks", False)
        names, trainables, values = ([], [], []) if (wire and self._he_scheme is None) else self.get_model_weights()
        ms_per_batch = total_ms / max(1, steps_done)
        msg = ModelProtoFactory.CompletedLearningTaskProtoMessage(
            weights_values=values, weights_trainable=trainables, weights_names=names,
            train_stats=train_stats, completed_epochs=steps_done / max(1, spe),
            global_iteration=learning_task_pb.global_iteration, validation_stats=valid_stats, test_stats=test_stats,
            completes_batches=steps_done, batch_size=batch_size,
            processing_ms_per_epoch=ms_per_batch * spe, processing_ms_per_batch=ms_per_batch)
        if verbose:
            MetisLogger.info("task done: %d steps, %.2f ms/batch, loss %s", steps_done, ms_per_batch,
                             train_stats.get("loss", [None])[-1])
        if wire and self._he_scheme is None:
            from metisfl_b200.ops import pack
            from metisfl_b200.proto.envelope import WireCompletedTask
            return WireCompletedTask(pack.arena_to_model_bytes(self.arena), msg.construct_task_execution_metadata_pb())
        if wire:
            return msg.construct_completed_learning_task_wire(he_scheme=self._he_scheme)
        return msg.construct_completed_learning_task_pb(he_scheme=self._he_scheme)

    def _eager_step(self, x, y):
        tr = self.trainer
        self._model.train()
        if x.is_floating_point():
            x = x.to(self.arena.model_dtype)
        loss = tr._fwd_bwd(x, y)
        tr.apply_optimizer()
        return loss

    def _train_with_user_fit(self, train_dataset, learning_task_pb, batch_size, total_steps):
        batches = _Batches(train_dataset, batch_size, shuffle=True, pin=False)
        spe = max(1, batches.steps_per_epoch)
        epochs = max(1, math.ceil(total_steps / spe))
        t0 = time.perf_counter()
        stats = self._model.fit(batches, epochs=epochs) or {}
        total_ms = (time.perf_counter() - t0) * 1e3
        self.arena.sync_master_from_model()
        names, trainables, values = self.get_model_weights()
        steps = epochs * spe
        msg = ModelProtoFactory.CompletedLearningTaskProtoMessage(
            weights_values=values, weights_trainable=trainables, weights_names=names, train_stats=stats,
            completed_epochs=epochs, global_iteration=learning_task_pb.global_iteration, completes_batches=steps,
            batch_size=batch_size, processing_ms_per_epoch=total_ms / epochs, processing_ms_per_batch=total_ms / steps)
        if getattr(self, "wire_tasks", False):
            return msg.construct_completed_learning_task_wire(he_scheme=self._he_scheme)
        return msg.construct_completed_learning_task_pb(he_scheme=self._he_scheme)

    # -- evaluation / inference -----------------------------------------------------------------
    @torch.no_grad()
    def evaluate_model(self, eval_dataset: ModelDataset, batch_size=100, metrics=None, verbose=False, *args,
                       **kwargs) -> dict:
        if eval_dataset is None or not eval_dataset.get_size():
            return {}
        if _overrides(self._model, "evaluate"):
            return dict(self._model.evaluate(_Batches(eval_dataset, batch_size or 100, False, False)) or {})
        batches = _Batches(eval_dataset, batch_size or 100, shuffle=False, pin=self.device.type == "cuda")
        loss_sum = torch.zeros((), device=self.device)
        correct = torch.zeros((), device=self.device)
        seen, classification = 0, False
        self._model.eval()
        for x, y in batches:
            x = x.to(self.device, non_blocking=True)
            y = y.to(self.device, non_blocking=True)
            if x.is_floating_point():
                x = x.to(self.arena.model_dtype)
            out = self._model(x)
            loss_sum += self.loss_fn(out, y).float() * x.shape[0]
            if out.dim() == 2 and not y.is_floating_point():
                classification = True
                correct += (out.argmax(dim=-1) == y).sum()
            seen += x.shape[0]
        self._model.train()
        res = {"loss": float(loss_sum) / max(1, seen)}
        if classification:
            res["accuracy"] = float(correct) / max(1, seen)
        if metrics:
            res = {k: v for k, v in res.items() if k in metrics or k == "loss"} or res
        return res

    @torch.no_grad()
    def infer_model(self, infer_dataset: ModelDataset, batch_size=100, *args, **kwargs):
        if infer_dataset is None:
            raise RuntimeError("Provided `dataset` for inference is None.")
        outs = []
        self._model.eval()
        for x, _ in _Batches(infer_dataset, batch_size or 100, shuffle=False, pin=False):
            x = x.to(self.device)
            if x.is_floating_point():
                x = x.to(self.arena.model_dtype)
            outs.append(self._model(x).float().cpu())
        self._model.train()
        return torch.cat(outs).numpy() if outs else np.empty((0,))


def _overrides(module, name) -> bool:
    """True if the user's module defines ``name`` itself (not the PyTorchDef placeholder)."""
    from metisfl_b200.models.model_def import PyTorchDef
    fn = getattr(type(module), name, None)
    return fn is not None and fn is not getattr(PyTorchDef, name, None) and fn is not getattr(torch.nn.Module, name, None)


# name kept for code written against the reference
PyTorchModelOps = TorchModelOps
