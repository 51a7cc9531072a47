"""Dense classifiers / regressors.

``FashionMnistMLP`` matches the shape of the reference's Keras example (784-128-128-10 = 118 282
parameters, ``examples/keras/models/fashion_mnist_fc.py:18-25``); ``HousingMLP`` is the
parametric-width/depth regressor of its scalability harness
(``examples/keras/models/housing_mlp.py:8-43``).
"""
import torch
import torch.nn as nn


class FashionMnistMLP(nn.Module):
    def __init__(self, in_features=784, hidden=(128, 128), num_classes=10):
        super().__init__()
        layers, d = [], in_features
        for h in hidden:
            layers += [nn.Linear(d, h), nn.ReLU()]
            d = h
        layers.append(nn.Linear(d, num_classes))
        self.net = nn.Sequential(*layers)

    def forward(self, x):
        return self.net(x.flatten(1))


class HousingMLP(nn.Module):
    def __init__(self, in_features=13, width=100, depth=10):
        super().__init__()
        layers, d = [], in_features
        for _ in range(depth):
            layers += [nn.Linear(d, width), nn.ReLU()]
            d = width
        layers.append(nn.Linear(d, 1))
        self.net = nn.Sequential(*layers)

    def forward(self, x):
        return self.net(x).squeeze(-1)
