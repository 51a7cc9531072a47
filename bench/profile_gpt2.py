#!/usr/bin/env python
"""Kernel-level breakdown of one GPT-2-medium learner step (batch 4 x 1024 tokens) of config #5:
forward + backward + fused AdamW, eager (torch.profiler) and as the CUDA graph the engine replays."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from metisfl_b200.models import zoo  # noqa: E402
from metisfl_b200.models.trainer import LocalTrainer  # noqa: E402
from metisfl_b200.ops import gemm  # noqa: E402
from metisfl_b200.ops.optim import OptimHyper  # noqa: E402
from metisfl_b200.parallel.arena import ModelArena  # noqa: E402

dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = zoo.gpt2_medium() if os.environ.get("MODEL", "medium") == "medium" else zoo.gpt2_small()
if os.environ.get("B200_LINEAR", "1") == "1":
    gemm.use_b200_linear(model)
arena = ModelArena(model, dev, torch.bfloat16)
x = torch.randint(0, 50257, (4, 1024), device=dev)
hyper = OptimHyper("AdamW", lr=1e-4, weight_decay=0.01)
eager = LocalTrainer(arena, hyper, zoo.lm_loss, use_cuda_graph=False)
for _ in range(2):
    eager.train_step(x, x)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(2):
        eager.train_step(x, x)
    torch.cuda.synchronize()
rows = []
for e in prof.key_averages():
    t = getattr(e, "device_time_total", 0) or getattr(e, "cuda_time_total", 0)
    if e.device_type == torch.autograd.DeviceType.CUDA and t > 0:
        rows.append((t / 2.0, e.count / 2.0, e.key[:110]))
rows.sort(reverse=True)
graph = LocalTrainer(arena, hyper, zoo.lm_loss, use_cuda_graph=True)
for _ in range(3):
    graph.train_step(x, x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    graph.train_step(x, x)
e1.record(); e1.synchronize()
print(json.dumps({"sum_kernel_ms_per_step": sum(r[0] for r in rows) / 1e3, "kernels_per_step": sum(r[1] for r in rows),
                  "graph_replay_ms_per_step": e0.elapsed_time(e1) / 5,
                  "top": [{"ms": round(r[0] / 1e3, 3), "n": r[1], "name": r[2]} for r in rows[:30]]}))
