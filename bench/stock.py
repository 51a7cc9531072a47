"""Stock arm of the headline benchmark: what a user gets WITHOUT this framework's kernels on the same box.

The reference itself cannot be installed here, so the north star names the stand-in: a plain PyTorch /
cuDNN ResNet-18 learner plus a ``torch.distributed`` NCCL FedAvg (``buf *= w; all_reduce``) - baseline B3 of
SURVEY section 6.2, measured for the whole ROUND, in the same process and under the same clocks as the product arm.
Nothing of metisfl_b200 is on this path: standard ``nn.Conv2d`` / ``nn.BatchNorm2d`` / ``F.relu`` modules in
channels_last bf16, ``torch.optim.SGD`` (fused), cuDNN autotuning, and - so that launch latency is not what is being
compared - the local step captured into a CUDA graph with stock ``torch.cuda.CUDAGraph``.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


class _Block(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.shortcut = None
        if stride != 1 or cin != cout:
            self.shortcut = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return F.relu(out + (x if self.shortcut is None else self.shortcut(x)))


class StockResNet18(nn.Module):
    def __init__(self, num_classes=10, width=64):
        super().__init__()
        self.conv1 = nn.Conv2d(3, width, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(width)
        blocks, cin = [], width
        for cout, stride in [(width, 1), (width * 2, 2), (width * 4, 2), (width * 8, 2)]:
            blocks += [_Block(cin, cout, stride), _Block(cout, cout, 1)]
            cin = cout
        self.layers = nn.Sequential(*blocks)
        self.fc = nn.Linear(cin, num_classes)

    def forward(self, x):
        x = F.relu(self.bn1(self.conv1(x)))
        x = self.layers(x)
        return self.fc(F.adaptive_avg_pool2d(x, 1).flatten(1))


class StockLearner:
    """One learner of the stock arm: bf16 channels_last module, fused SGD, optional CUDA-graph step."""

    def __init__(self, device, batch_shape, lr=0.02, momentum=0.9, graph=True):
        self.device = device
        self.model = StockResNet18().to(device=device, dtype=torch.bfloat16).to(memory_format=torch.channels_last)
        try:
            self.opt = torch.optim.SGD(self.model.parameters(), lr=lr, momentum=momentum, fused=True)
        except (TypeError, RuntimeError):
            self.opt = torch.optim.SGD(self.model.parameters(), lr=lr, momentum=momentum, foreach=True)
        self.x = torch.zeros(batch_shape, device=device, dtype=torch.bfloat16).contiguous(
            memory_format=torch.channels_last)
        self.y = torch.zeros(batch_shape[0], device=device, dtype=torch.int64)
        self.tensors = [p for p in self.model.parameters()] + [b for b in self.model.buffers() if b.is_floating_point()]
        self.graph = None
        self.loss = None
        if graph:
            try:
                self._capture()
            except Exception as e:  # keep the arm alive: eager is still a valid stock number
                self.graph = None
                self.capture_error = repr(e)

    def _step(self):
        self.opt.zero_grad(set_to_none=False)
        loss = F.cross_entropy(self.model(self.x).float(), self.y)
        loss.backward()
        self.opt.step()
        return loss.detach()

    def _capture(self):
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(3):
                self._step()
        torch.cuda.current_stream(self.device).wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.loss = self._step()
        self.graph = g

    def step(self, x, y):
        self.x.copy_(x, non_blocking=True)
        self.y.copy_(y, non_blocking=True)
        if self.graph is not None:
            self.graph.replay()
            return self.loss
        return self._step()

    def fedavg(self, weight, world):
        """``buf *= w; all_reduce(buf)`` over the flattened state, written back into the module."""
        import torch.distributed as dist
        flat = torch._utils._flatten_dense_tensors([t.data for t in self.tensors])
        flat.mul_(weight)
        if world > 1:
            dist.all_reduce(flat)
        for t, v in zip(self.tensors, torch._utils._unflatten_dense_tensors(flat, [t.data for t in self.tensors])):
            t.data.copy_(v)
