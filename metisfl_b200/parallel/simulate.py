"""Event-driven simulation of heterogeneous learners for the asynchronous protocols (config #3).

The reference exercises asynchrony with real learner processes of different speeds
(``examples/config/fashionmnist/test_localhost_asynchronous_vanillasgd.yaml``; the asynchronous scheduler
re-schedules only the caller, ``asynchronous_scheduler.h:12-19``).  On one box with k logical learners per GPU
the order of completions is made deterministic: every learner has a simulated task duration, completions are
processed in simulated-time order, and each completion is one community update (FedAsync mixing with a
staleness weight, or FedRec's recency swap, ``federated_recency.cc:27-94``).

What actually runs is asynchronous.  Each GPU trains its own learners back to back on its stream; nothing on a
GPU waits for another GPU's training.  The folds are executed by ONE GPU (the executor) on a separate stream, in
the agreed order: its kernel (K3, ``csrc/kernels/async_fold.cu``) waits on the owner's "published" counter, reads
the slot once over NVLink and multicasts the new community model and its version to all GPUs.  A learner's next
task waits only for ITS OWN previous fold (a version counter / CUDA event) and then installs whatever community
version is newest on its GPU - at least the one the simulated order promises, possibly a newer one, as in a real
asynchronous federation.  The staleness used for FedAsync's weight is the simulated one (updates between a task's
start and its fold in the agreed order).
"""
from __future__ import annotations

import heapq
from typing import Callable, Dict, List, Sequence

from metisfl_b200.parallel.engine import FederationEngine


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
        for g in range(engine.num_learners):  # everybody starts from the initial community model
            self._start(g)

    def _start(self, g: int):
        e = self.e
        r, j = divmod(g, e.k)
        self.started_at_version[g] = e.version
        if r == e.rank:
            loss = e.train_learner(j, self.batches_fn(g, self.task_no[g]))
            if loss is not None:
                self.losses[g] = loss
        self.task_no[g] += 1
        heapq.heappush(self.events, (self.now + self.durations[g], g))

    def _fold(self, g: int):
        staleness = self.e.version - self.started_at_version[g]
        self.staleness_log.append(staleness)
        self.e.fold_async(g, staleness=staleness, task_no=self.task_no[g])

    def run(self, n_updates: int):
        """Process ``n_updates`` completion events (= community updates)."""
        for _ in range(n_updates):
            self.now, g = heapq.heappop(self.events)
            self._fold(g)
            self._start(g)
        return {"updates": n_updates, "mean_staleness": sum(self.staleness_log) / max(1, len(self.staleness_log)),
                "max_staleness": max(self.staleness_log or [0]), "sim_time": self.now}
