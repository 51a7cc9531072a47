#!/usr/bin/env python
"""BASELINE.json config #3: CIFAR-10-shaped ResNet-18, 64 simulated heterogeneous learners on 8 x B200
(8 logical learners per GPU), asynchronous FedAsync or FedRec with staleness weighting.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/resnet_async.py --rule FedAsync
"""
import argparse

from _common import device_time_ms, finish, init

import torch

from metisfl_b200.models.zoo import ResNet18
from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
from metisfl_b200.parallel.simulate import AsyncFederation

ap = argparse.ArgumentParser()
ap.add_argument("--rule", default="FedAsync", choices=["FedAsync", "FedRec"])
ap.add_argument("--learners-per-gpu", type=int, default=8)
ap.add_argument("--updates", type=int, default=128)
ap.add_argument("--local-steps", type=int, default=2)
ap.add_argument("--batch", type=int, default=128)
a = ap.parse_args()
rank, world, dev = init()
cfg = EngineConfig(learners_per_gpu=a.learners_per_gpu, rule=a.rule, staleness_fn="polynomial", fedasync_alpha=0.6)
eng = FederationEngine(ResNet18, cfg, device=dev)
L = eng.num_learners
g = torch.Generator().manual_seed(7)
pool = [(torch.randn(a.batch, 3, 32, 32, generator=g).to(torch.bfloat16).contiguous(memory_format=torch.channels_last).to(dev),
         torch.randint(0, 10, (a.batch,), generator=g).to(dev)) for _ in range(4)]
# heterogeneous speeds: task durations spread over a 4x range, deterministic and identical on every rank
durations = [1.0 + 3.0 * ((i * 37) % L) / max(1, L - 1) for i in range(L)]
sim = AsyncFederation(eng, durations, lambda gl, n: [pool[(gl + n + s) % len(pool)] for s in range(a.local_steps)])
sim.run(max(8, L))  # warm-up (CUDA graph capture, cuDNN autotune, every learner has folded once)
eng.sync_folds()
stats = {}


def timed():
    stats.update(sim.run(a.updates))
    if eng.fold_stream is not None:  # the executor's folds run on their own stream: the clock stops after the last one
        torch.cuda.current_stream(dev).wait_stream(eng.fold_stream)


ms = device_time_ms(timed, dev, world)
eng.sync_folds()
health = eng.async_health()
finish(rank, world, {
    "config": "#3 ResNet-18 async", "rule": a.rule, "learners": L, "gpus": world, "community_updates": a.updates,
    "ms_per_update": ms / a.updates, "updates_per_s": a.updates / (ms * 1e-3),
    "aggregated_params_per_s": a.updates * eng.layout.num_params / (ms * 1e-3),
    "mean_staleness": stats.get("mean_staleness"), "max_staleness": stats.get("max_staleness"),
    "local_steps": a.local_steps, "batch": a.batch, "executor_health": health,
    "mode": "one-sided folds by the executor GPU (K3), multicast push" if eng.community.mc_ptr else
            "one-sided folds by the executor GPU (K3), peer stores"}, engine=eng)
