"""Event-driven simulation of heterogeneous learners for the asynchronous protocols (config #3).

The reference exercises asynchrony with real learner processes of different speeds
(``examples/config/fashionmnist/test_localhost_asynchronous_vanillasgd.yaml``); on one box with k
logical learners per GPU the same behaviour is reproduced deterministically: every learner has a
task duration, a learner's task reads the community model that is current when the task *starts*,
and its result is folded when its completion time comes - with the staleness (community updates since
it started) feeding FedAsync's weight or FedRec's recency swap.  Every rank replays the same global
event order, so the cross-GPU folds (peer reads of the finishing learner's slot) stay in lockstep.
"""
from __future__ import annotations

import heapq
from typing import Callable, Dict, List, Sequence

import torch

from metisfl_b200.ops import aggregation as agg_ops
from metisfl_b200.parallel.engine import FederationEngine, staleness_factor
from metisfl_b200.parallel.fabric import _CudaMem


class AsyncFederation:
    def __init__(self, engine: FederationEngine, durations: Sequence[float],
                 batches_fn: Callable[[int, int], Sequence]):
        """``durations[g]`` = simulated seconds per task of global learner g; ``batches_fn(g, n)`` =
        batches of its n-th task."""
        assert engine.cfg.rule in ("FedAsync", "FedRec")
        self.e = engine
        self.durations = list(durations)
        assert len(self.durations) == engine.num_learners
        self.batches_fn = batches_fn
        self.task_no = [0] * engine.num_learners
        self.started_at_version = [0] * engine.num_learners
        self.staleness_log: List[int] = []
        self.losses: Dict[int, float] = {}
        self.now = 0.0
        self.events = []
        nb = engine.layout.total_bytes
        # typed views of EVERY learner's slot (peer-mapped memory is addressable by torch kernels too)
        self.views = []
        for g in range(engine.num_learners):
            r, j = divmod(g, engine.k)
            if r == engine.rank or not engine.fabric.peer_mapped:
                self.views.append(engine.slot_views[j] if r == engine.rank else None)
            else:
                raw = torch.as_tensor(_CudaMem(engine.slots.ptrs[r] + j * nb, nb), device=engine.device)
                self.views.append(engine.layout.views(raw))
        self._rec_S = None
        self._rec_prev = [None] * engine.num_learners
        self._rec_w = [0.0] * engine.num_learners
        self._rec_z = 0.0
        for g in range(engine.num_learners):  # everybody starts from the initial community model
            self._start(g)

    def _start(self, g: int):
        e = self.e
        r, j = divmod(g, e.k)
        self.started_at_version[g] = e.global_iteration
        if r == e.rank:
            loss = e.train_learner(j, self.batches_fn(g, self.task_no[g]))
            if loss is not None:
                self.losses[g] = loss
        self.task_no[g] += 1
        heapq.heappush(self.events, (self.now + self.durations[g], g))

    def _fold(self, g: int):
        e, cfg = self.e, self.e.cfg
        staleness = e.global_iteration - self.started_at_version[g]
        self.staleness_log.append(staleness)
        x_views = self.views[g]
        e.fabric.device_barrier(0)  # the finishing learner's slot is complete and visible
        if cfg.rule == "FedAsync":
            a = cfg.fedasync_alpha * staleness_factor(cfg.staleness_fn, staleness, cfg.staleness_a, cfg.staleness_b)
            for si, sg in enumerate(e.layout.segments):
                c = e.community_views[si]
                if sg.exact:
                    c.copy_(x_views[si])
                else:
                    agg_ops.weighted_sum([c, x_views[si]], [1.0 - a, a], out=c)
                e.gpu_launches += 1
        else:  # FedRec: swap this learner's previous contribution for the new one
            if self._rec_S is None:
                self._rec_S = [torch.zeros(sg.numel, dtype=torch.float32, device=e.device) if not sg.exact else None
                               for sg in e.layout.segments]
            w_new = float(e.examples[g]) if cfg.scaling != "NumParticipants" else 1.0
            prev, w_old = self._rec_prev[g], self._rec_w[g]
            self._rec_z += w_new - (w_old if prev is not None else 0.0)
            for si, sg in enumerate(e.layout.segments):
                c = e.community_views[si]
                if sg.exact:
                    c.copy_(x_views[si])
                else:
                    agg_ops.rolling_update(c, self._rec_S[si], x_views[si], w_new,
                                           prev[si] if prev is not None else None, w_old, self._rec_z)
                e.gpu_launches += 1
            self._rec_prev[g] = [v.clone() for v in x_views]
            self._rec_w[g] = w_new
        e.fabric.device_barrier(1)  # every rank has read the slot before its owner trains into it again
        e.global_iteration += 1

    def run(self, n_updates: int):
        """Process ``n_updates`` completion events (= community updates)."""
        for _ in range(n_updates):
            self.now, g = heapq.heappop(self.events)
            self._fold(g)
            self._start(g)
        return {"updates": n_updates, "mean_staleness": sum(self.staleness_log) / max(1, len(self.staleness_log)),
                "max_staleness": max(self.staleness_log or [0]), "sim_time": self.now}
