"""Small vision classifiers of the reference's example zoo, as ``nn.Module``s.

* ``CifarCNN``   - 6 conv (64,64,128,128,128,128) + 2 BatchNorm + dense 512 -> 10, 1 610 314 parameters
  (reference ``examples/keras/models/cifar_cnn.py:14-40``; the count is the one quoted in ``controller.cc:602``).
* ``MnistFC`` / ``FashionMnistFC`` - 784-128-128-10 dense nets (``mnist_fc.py``, ``fashion_mnist_fc.py``).
* ``MelanomaFC`` - frozen convolutional trunk + 8-unit head + sigmoid output (``melanoma_fc.py:12-32``); the
  reference downloads ImageNet Xception weights, which is impossible offline, so the trunk here is a small frozen
  random-feature extractor with the same head - the federated contract (only the head's tensors are trainable) is
  what the example exercises.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from metisfl_b200.models.model_def import PyTorchDef
from metisfl_b200.models.zoo.mlp import FashionMnistMLP


class CifarCNN(nn.Module):
    def __init__(self, num_classes: int = 10, dropout: float = 0.2):
        super().__init__()

        def block(cin, cout, bn):
            layers = [nn.BatchNorm2d(cin)] if bn else []
            layers += [nn.Conv2d(cin, cout, 3, padding=1), nn.ReLU(inplace=True),
                       nn.Conv2d(cout, cout, 3, padding=1), nn.ReLU(inplace=True), nn.MaxPool2d(2)]
            return nn.Sequential(*layers)

        self.features = nn.Sequential(block(3, 64, False), block(64, 128, True), block(128, 128, True))
        self.fc1 = nn.Linear(128 * 4 * 4, 512)
        self.drop = nn.Dropout(dropout)
        self.fc2 = nn.Linear(512, num_classes)

    def forward(self, x):
        x = self.features(x)
        # NHWC flatten order so the dense layer sees the same element order as the Keras model
        x = x.permute(0, 2, 3, 1).flatten(1)
        return self.fc2(self.drop(F.relu(self.fc1(x))))


class MnistFC(FashionMnistMLP):
    pass


class FashionMnistFC(FashionMnistMLP):
    pass


class MelanomaFC(nn.Module, PyTorchDef):
    def __init__(self, image_size=(64, 64), trunk_width: int = 64):
        super().__init__()
        self.trunk = nn.Sequential(
            nn.Conv2d(3, trunk_width, 3, stride=2, padding=1), nn.ReLU(inplace=True),
            nn.Conv2d(trunk_width, trunk_width * 2, 3, stride=2, padding=1), nn.ReLU(inplace=True),
            nn.AdaptiveAvgPool2d(1), nn.Flatten())
        for p in self.trunk.parameters():
            p.requires_grad_(False)
        self.head = nn.Sequential(nn.Linear(trunk_width * 2, 8), nn.ReLU(inplace=True), nn.Dropout(0.7), nn.Linear(8, 1))

    def forward(self, x):
        return self.head(self.trunk(x)).squeeze(-1)

    def loss_fn(self, outputs, targets):
        return F.binary_cross_entropy_with_logits(outputs.float(), targets.float())
