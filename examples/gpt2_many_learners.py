#!/usr/bin/env python
"""BASELINE.json config #5: GPT-2 medium (355 M) FedAvg with 1024 simulated learners multiplexed over
8 x B200 (128 logical learners per GPU), streaming aggregation: no per-learner model is kept.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/gpt2_many_learners.py
"""
import argparse

from _common import device_time_ms, finish, init

import torch

from metisfl_b200.models import zoo
from metisfl_b200.ops.optim import OptimHyper
from metisfl_b200.parallel.engine import EngineConfig, FederationEngine

ap = argparse.ArgumentParser()
ap.add_argument("--learners-per-gpu", type=int, default=128)
ap.add_argument("--model", default="gpt2-medium", choices=["gpt2-medium", "gpt2-small"])
ap.add_argument("--seq-len", type=int, default=1024)
ap.add_argument("--batch", type=int, default=4)
ap.add_argument("--local-steps", type=int, default=1)
ap.add_argument("--rounds", type=int, default=1)
a = ap.parse_args()
rank, world, dev = init()
factory = zoo.gpt2_medium if a.model == "gpt2-medium" else zoo.gpt2_small
cfg = EngineConfig(learners_per_gpu=2, optimizer=OptimHyper("AdamW", lr=1e-4, weight_decay=0.01), channels_last=False)
eng = FederationEngine(factory, cfg, loss_fn=zoo.lm_loss, device=dev)
g = torch.Generator().manual_seed(11 + rank)
pool = [torch.randint(0, 50257, (a.batch, a.seq_len), generator=g).to(dev) for _ in range(4)]
batches = lambda gl: [(pool[(gl + s) % len(pool)],) * 2 for s in range(a.local_steps)]
eng.run_round_streaming(2, batches)  # warm-up
out = {}
ms = device_time_ms(lambda: [out.update(eng.run_round_streaming(a.learners_per_gpu, batches)) for _ in range(a.rounds)],
                    dev, world)
n_learners = a.learners_per_gpu * world
finish(rank, world, {
    "config": "#5 GPT-2 many learners", "model": a.model, "params": eng.layout.num_params, "learners": n_learners,
    "gpus": world, "round_time_s": ms * 1e-3 / a.rounds,
    "aggregated_params_per_s": n_learners * eng.layout.num_params / (ms * 1e-3 / a.rounds),
    "peak_mem_gib": torch.cuda.max_memory_allocated(dev) / 2**30 if dev.type == "cuda" else None,
    "last_loss": out.get("loss")}, engine=eng)
