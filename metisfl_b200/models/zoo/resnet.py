"""ResNet-18 for 32x32 inputs (CIFAR-10 variant: 3x3 stem, no max-pool), 11 173 962 parameters
plus BatchNorm running statistics.  The reference's closest model is a 1.6 M-parameter Keras CNN
with BatchNorm (``examples/keras/models/cifar_cnn.py:21-53``); ResNet-18 is the north-star config
#2 (BASELINE.json).  ``channels_last`` keeps convolutions on the NHWC tensor-core path; every BatchNorm (+ residual add) (+ ReLU)
group is one call of the fused kernel family K14 (``metisfl_b200/ops/bn.py``) on CUDA.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from metisfl_b200.ops.bn import bn_act


PARALLEL_SHORTCUT = True  # downsample branch (1x1 conv + BN) on a second stream, concurrent with conv1 / bn1
_SHORTCUT_STREAMS = {}


def _shortcut_stream(device):
    s = _SHORTCUT_STREAMS.get(device)
    if s is None:
        s = _SHORTCUT_STREAMS[device] = torch.cuda.Stream(device=device)
    return s


class BasicBlock(nn.Module):
    expansion = 1

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
        if self.shortcut is not None and x.is_cuda and PARALLEL_SHORTCUT:
            # the two branches are independent and each is a chain of short kernels: run them side by side (inside
            # a CUDA graph capture this becomes a fork / join; autograd replays the same streams in backward)
            main = torch.cuda.current_stream(x.device)
            side = _shortcut_stream(x.device)
            side.wait_stream(main)
            with torch.cuda.stream(side):
                skip = bn_act(self.shortcut[1], self.shortcut[0](x), relu=False)
            x.record_stream(side)
            out = bn_act(self.bn1, self.conv1(x))
            out = self.conv2(out)
            main.wait_stream(side)
            skip.record_stream(main)
            return bn_act(self.bn2, out, residual=skip)
        out = bn_act(self.bn1, self.conv1(x))
        skip = x if self.shortcut is None else bn_act(self.shortcut[1], self.shortcut[0](x), relu=False)
        return bn_act(self.bn2, self.conv2(out), residual=skip)  # BN + residual add + ReLU in one pass (K14)


class ResNet18(nn.Module):
    def __init__(self, num_classes=10, width=64):
        super().__init__()
        self.conv1 = nn.Conv2d(3, width, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(width)
        cfg = [(width, 1), (width * 2, 2), (width * 4, 2), (width * 8, 2)]
        blocks, cin = [], width
        for cout, stride in cfg:
            blocks += [BasicBlock(cin, cout, stride), BasicBlock(cout, cout, 1)]
            cin = cout
        self.layers = nn.Sequential(*blocks)
        self.fc = nn.Linear(cin, num_classes)

    def forward(self, x):
        x = bn_act(self.bn1, self.conv1(x))
        x = self.layers(x)
        x = F.adaptive_avg_pool2d(x, 1).flatten(1)
        return self.fc(x)
