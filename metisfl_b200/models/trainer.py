"""Framework-owned local training loop on a flat model arena.

The reference delegates local training to ``model.fit`` (Keras, ``keras_model_ops.py:117-197``) or
to a user-written PyTorch loop (``pytorch_model_ops.py:83-131``) and never places the model on a
device (quirk Q10).  Here the learner owns the loop: forward/backward run in PyTorch on the
arena-backed module (captured once into a CUDA graph, since the per-layer launches of a
CIFAR-sized network are launch-bound on a B200), and the optimizer step - fused with zeroing the
gradients and with publishing the new parameters into the peer-visible slot - is the
hand-written K2 kernel (``csrc/kernels/optim.cu``).
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F

from metisfl_b200.ops import _lib
from metisfl_b200.ops import conv as _conv_ops
from metisfl_b200.ops.optim import OptimHyper, optimizer_step
from metisfl_b200.parallel.arena import ModelArena


def cross_entropy_loss(logits, target):
    return F.cross_entropy(logits.float(), target)


class LocalTrainer:
    def __init__(self, arena: ModelArena, hyper: OptimHyper, loss_fn: Callable = cross_entropy_loss,
                 use_cuda_graph: Optional[bool] = None):
        self.arena = arena
        self.module = arena.module
        self.hyper = hyper
        self.loss_fn = loss_fn
        self.device = arena.device
        self.use_graph = self.device.type == "cuda" if use_cuda_graph is None else use_cuda_graph
        self.step_count = 0
        self._graph = None
        self._graph_key = None
        self._static_x = self._static_y = self._static_loss = None
        self.launches_per_step = 0
        self.graph_launches = 0
        self._alloc_state()

    # -- optimizer state -----------------------------------------------------------------------
    def _alloc_state(self):
        k = self.hyper.kind
        self.s1: List[Optional[torch.Tensor]] = []
        self.s2: List[Optional[torch.Tensor]] = []
        self.anchor: List[Optional[torch.Tensor]] = []
        for s, m in zip(self.arena.layout.segments, self.arena.master):
            need1 = s.trainable and k in (1, 3, 4)
            need2 = s.trainable and k in (3, 4)
            needa = s.trainable and k == 2
            self.s1.append(torch.zeros_like(m) if need1 else None)
            self.s2.append(torch.zeros_like(m) if need2 else None)
            self.anchor.append(torch.zeros_like(m) if needa else None)

    def set_hyper(self, hyper: OptimHyper):
        if hyper.kind != self.hyper.kind:
            self.hyper = hyper
            self._alloc_state()
        self.hyper = hyper

    def reset_optimizer_state(self):
        self.step_count = 0
        for t in self.s1 + self.s2:
            if t is not None:
                t.zero_()

    def snapshot_anchor(self):
        """FedProx: remember the community weights the task started from (theta*)."""
        for a, m in zip(self.anchor, self.arena.master):
            if a is not None:
                a.copy_(m)

    # -- forward/backward ----------------------------------------------------------------------
    def _fwd_bwd(self, x, y):
        out = self.module(x)
        loss = self.loss_fn(out, y)
        loss.backward()
        if self.device.type == "cuda":
            _conv_ops.join()  # weight gradients issued on the side stream (ops/conv.py)
        return loss.detach()

    def _capture(self, x, y):
        arena = self.arena
        saved = arena.state.clone()
        self._static_x = torch.empty_like(x)
        self._static_y = torch.empty_like(y)
        self._static_x.copy_(x)
        self._static_y.copy_(y)
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(3):  # warm-up: cuDNN autotune, lazy inits, autograd graph caches
                self._fwd_bwd(self._static_x, self._static_y)
        torch.cuda.current_stream(self.device).wait_stream(side)
        for g in arena.grad:
            if g is not None:
                g.zero_()
        self._graph = torch.cuda.CUDAGraph()
        launches0 = _lib.LAUNCHES
        # thread_local: a learner's evaluation thread may be launching / allocating on this device at the same time
        with torch.cuda.graph(self._graph, capture_error_mode="thread_local"):
            self._static_loss = self._fwd_bwd(self._static_x, self._static_y)
        self.graph_launches = _lib.LAUNCHES - launches0  # library kernels (K10 / K14 ...) every replay launches
        # undo what warm-up and capture did to BatchNorm statistics etc.
        arena.state.copy_(saved)
        for g in arena.grad:
            if g is not None:
                g.zero_()
        self._graph_key = (tuple(x.shape), x.dtype, tuple(y.shape), y.dtype)

    def _x_dtype(self, x):
        return self.arena.model_dtype if x.is_floating_point() else x.dtype

    def forward_backward(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """Accumulates gradients into the flat grad buffers; returns the (device) loss."""
        self.module.train()
        if x.dtype != self._x_dtype(x):
            x = x.to(self._x_dtype(x))
        if self.use_graph:
            key = (tuple(x.shape), x.dtype, tuple(y.shape), y.dtype)
            if self._graph is None or key != self._graph_key:
                self._capture(x, y)
            if x.data_ptr() != self._static_x.data_ptr():
                self._static_x.copy_(x, non_blocking=True)
            if y.data_ptr() != self._static_y.data_ptr():
                self._static_y.copy_(y, non_blocking=True)
            self._graph.replay()
            _lib.LAUNCHES += self.graph_launches
            return self._static_loss
        return self._fwd_bwd(x, y)

    def input_buffers(self, x_like: torch.Tensor, y_like: torch.Tensor):
        """Static device buffers the CUDA graph reads; H2D copies can land in them directly."""
        if self.use_graph:
            key = (tuple(x_like.shape), self._x_dtype(x_like), tuple(y_like.shape), y_like.dtype)
            if self._graph is None or key != self._graph_key:
                self._capture(x_like.to(self.device, self._x_dtype(x_like)), y_like.to(self.device))
            return self._static_x, self._static_y
        return None, None

    # -- optimizer -------------------------------------------------------------------------------
    def apply_optimizer(self, publish_views: Optional[Sequence[torch.Tensor]] = None, publish_scale: float = 1.0):
        """K2 over every trainable segment; zeroes grads; optionally publishes in the same pass."""
        self.step_count += 1
        arena = self.arena
        n = 0
        for i, s in enumerate(arena.layout.segments):
            if not s.trainable:
                continue
            master = arena.master[i]
            model = None if master is arena.seg[i] else arena.seg[i]
            pub = publish_views[i] if publish_views is not None else None
            optimizer_step(master, model, arena.grad[i], self.hyper, self.s1[i], self.s2[i], self.anchor[i],
                           publish=pub, step=self.step_count, publish_scale=publish_scale, zero_grad=True)
            n += 1
        self.launches_per_step = n
        return n

    def train_step(self, x, y, publish_views=None, publish_scale: float = 1.0) -> torch.Tensor:
        loss = self.forward_backward(x, y)
        self.apply_optimizer(publish_views, publish_scale)
        return loss

    @torch.no_grad()
    def evaluate(self, x, y):
        """Returns (loss, #correct) as device tensors."""
        self.module.eval()
        if x.dtype != self._x_dtype(x):
            x = x.to(self._x_dtype(x))
        out = self.module(x)
        loss = self.loss_fn(out, y)
        correct = (out.argmax(dim=-1) == y).sum() if out.dim() == 2 and not y.is_floating_point() else None
        self.module.train()
        return loss, correct
