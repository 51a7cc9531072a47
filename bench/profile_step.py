#!/usr/bin/env python
"""Kernel-level breakdown of one local training step (eager, torch.profiler / CUPTI) of the flagship ResNet-18,
plus the CUDA-graph replay time of the same step.  Output: JSON with the top kernels by total device time."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from metisfl_b200.models.zoo import ResNet18  # noqa: E402
from metisfl_b200.ops.optim import OptimHyper  # noqa: E402
from metisfl_b200.parallel.arena import ModelArena  # noqa: E402
from metisfl_b200.models.trainer import LocalTrainer  # noqa: E402

batch = int(os.environ.get("BATCH", "128"))
conv = os.environ.get("B200_CONV", "0") == "1"
dev = torch.device("cuda", 0)
torch.manual_seed(0)
import metisfl_b200.models.zoo.resnet as _rn  # noqa: E402
_rn.PARALLEL_SHORTCUT = os.environ.get("PARALLEL_SHORTCUT", "1") == "1"
torch.backends.cudnn.benchmark = os.environ.get("CUDNN_BENCHMARK", "1") == "1"
model = ResNet18()
if conv:
    from metisfl_b200.ops.conv import use_b200_conv
    use_b200_conv(model)
elif os.environ.get("OVERLAP_WGRAD", "1") == "1":
    from metisfl_b200.ops.conv import use_overlapped_wgrad
    use_overlapped_wgrad(model)
arena = ModelArena(model, dev, torch.bfloat16)
x = torch.randn(batch, 3, 32, 32, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 10, (batch,), device=dev)

eager = LocalTrainer(arena, OptimHyper("MomentumSGD", lr=0.01, momentum=0.9), use_cuda_graph=False)
for _ in range(3):
    eager.train_step(x, y)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(5):
        eager.train_step(x, y)
    torch.cuda.synchronize()
rows = []
for e in prof.key_averages():
    t = getattr(e, "device_time_total", 0) or getattr(e, "cuda_time_total", 0)
    if e.device_type == torch.autograd.DeviceType.CUDA and t > 0:
        rows.append((t / 5.0, e.count / 5.0, e.key[:110]))
rows.sort(reverse=True)
total = sum(r[0] for r in rows)

graph = LocalTrainer(arena, OptimHyper("MomentumSGD", lr=0.01, momentum=0.9), use_cuda_graph=True)
for _ in range(5):
    graph.train_step(x, y)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    graph.train_step(x, y)
e1.record(); e1.synchronize()
print(json.dumps({"batch": batch, "b200_conv": conv, "sum_kernel_us_per_step": total, "kernels_per_step": sum(r[1] for r in rows),
                  "graph_replay_ms_per_step": e0.elapsed_time(e1) / 20,
                  "top": [{"us": round(r[0], 1), "n": r[1], "name": r[2]} for r in rows[:45]]}, indent=0))
