"""Neuro-imaging models of the reference zoo (MRI volumes 91 x 109 x 91), as ``nn.Module``s.

* ``BrainAge2DCNN`` - slices of one volume form the batch of a 2-D encoder (32-64-128-256-256 conv blocks with
  affine-free batch norm, max-pool, ReLU), 1x1 convs to a 32-d slice embedding, mean over slices, 64-unit dense,
  age regression with the output bias initialised at 62.68 (reference ``brainage_cnns.py:17-66``).
* ``BrainAge3DCNN`` - the 3-D variant (``brainage_cnns.py:84-113``).
* ``AlzheimersDisease2DCNN`` / ``AlzheimersDisease3DCNN`` - binary classifiers (``alzheimers_disease_cnns.py:15-47,
  67-93``).

Volumes are large (0.9 M voxels); the tests use reduced volumes, the shapes are parameters.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from metisfl_b200.models.model_def import PyTorchDef

AGE_BIAS = 62.68


def _block(conv, bn, pool, cin, cout, affine):
    return nn.Sequential(conv(cin, cout, 3, padding=1), bn(cout, affine=affine), pool(2), nn.ReLU(inplace=True))


class BrainAge2DCNN(nn.Module, PyTorchDef):
    """Input [B, D, H, W] (a batch of volumes); every volume's W axis is treated as the slice/batch axis."""

    def __init__(self, widths=(32, 64, 128, 256, 256), in_channels: int = None):
        super().__init__()
        chans = [1] + list(widths)
        self.encoder = nn.Sequential(*[_block(nn.Conv2d, nn.BatchNorm2d, nn.MaxPool2d, chans[i], chans[i + 1], False)
                                       for i in range(len(widths))])
        self.post1 = nn.Sequential(nn.Conv2d(widths[-1], 64, 1), nn.BatchNorm2d(64, affine=False), nn.ReLU(inplace=True))
        self.post2 = nn.Conv2d(64, 32, 1)
        self.fc1 = nn.Linear(32, 64)
        self.fc2 = nn.Linear(64, 1)
        nn.init.constant_(self.fc2.bias, AGE_BIAS)

    def forward(self, vol):
        b, d, h, w = vol.shape
        slices = vol.permute(0, 3, 1, 2).reshape(b * w, 1, d, h)       # slices become the conv batch
        z = self.post1(self.encoder(slices))
        z = self.post2(F.adaptive_avg_pool2d(z, 1)).flatten(1)          # [b*w, 32]
        z = z.view(b, w, -1).mean(dim=1)                                # mean slice embedding
        return self.fc2(F.relu(self.fc1(z))).squeeze(-1)

    def loss_fn(self, outputs, targets):
        return F.mse_loss(outputs.float(), targets.float())


class BrainAge3DCNN(nn.Module, PyTorchDef):
    """Input [B, D, H, W] or [B, 1, D, H, W]."""

    def __init__(self, widths=(32, 64, 128, 256, 256)):
        super().__init__()
        chans = [1] + list(widths)
        self.encoder = nn.Sequential(*[_block(nn.Conv3d, nn.BatchNorm3d, nn.MaxPool3d, chans[i], chans[i + 1], False)
                                       for i in range(len(widths))])
        self.post = nn.Sequential(nn.Conv3d(widths[-1], 64, 1), nn.BatchNorm3d(64, affine=False), nn.ReLU(inplace=True))
        self.reg = nn.Conv3d(64, 1, 1)
        nn.init.constant_(self.reg.bias, AGE_BIAS)

    def forward(self, vol):
        if vol.dim() == 4:
            vol = vol.unsqueeze(1)
        z = F.adaptive_avg_pool3d(self.post(self.encoder(vol)), 1)
        return self.reg(z).flatten()

    def loss_fn(self, outputs, targets):
        return F.mse_loss(outputs.float(), targets.float())


class AlzheimersDisease2DCNN(nn.Module, PyTorchDef):
    """Input [B, D, H, W]: the depth axis is folded into the batch, features are averaged over it."""

    def __init__(self, widths=(32, 64, 128, 256, 256)):
        super().__init__()
        chans = [1] + list(widths)
        self.encoder = nn.Sequential(*[_block(nn.Conv2d, nn.BatchNorm2d, nn.MaxPool2d, chans[i], chans[i + 1], True)
                                       for i in range(len(widths))])
        self.fc1 = nn.Linear(widths[-1], 64)
        self.fc2 = nn.Linear(64, 1)

    def forward(self, vol):
        b, d, h, w = vol.shape
        z = self.encoder(vol.reshape(b * d, 1, h, w))
        z = F.adaptive_avg_pool2d(z, 1).flatten(1).view(b, d, -1).mean(dim=1)
        return self.fc2(F.relu(self.fc1(z))).squeeze(-1)

    def loss_fn(self, outputs, targets):
        return F.binary_cross_entropy_with_logits(outputs.float(), targets.float())


class AlzheimersDisease3DCNN(nn.Module, PyTorchDef):
    def __init__(self, widths=(64, 64, 128, 256), dropout: float = 0.2):
        super().__init__()
        chans = [1] + list(widths)
        layers = []
        for i in range(len(widths)):
            layers += [nn.Conv3d(chans[i], chans[i + 1], 3, bias=False), nn.BatchNorm3d(chans[i + 1]),
                       nn.ReLU(inplace=True), nn.MaxPool3d(2)]
        self.encoder = nn.Sequential(*layers)
        self.fc1 = nn.Linear(widths[-1], 512)
        self.drop = nn.Dropout(dropout)
        self.fc2 = nn.Linear(512, 1)

    def forward(self, vol):
        if vol.dim() == 4:
            vol = vol.unsqueeze(1)
        z = F.adaptive_avg_pool3d(self.encoder(vol), 1).flatten(1)
        return self.fc2(self.drop(F.relu(self.fc1(z)))).squeeze(-1)

    def loss_fn(self, outputs, targets):
        return F.binary_cross_entropy_with_logits(outputs.float(), targets.float())
