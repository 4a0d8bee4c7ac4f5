"""The reference's default model (reference src/model.py:7-24): a 62,006-parameter LeNet-style CNN.

conv(3->6,5x5) -> ReLU -> maxpool2 -> conv(6->16,5x5) -> ReLU -> maxpool2 -> flatten 400
-> fc120 -> ReLU -> fc84 -> ReLU -> fc10.

Same attribute names, parameter shapes, ``state_dict`` keys and initialisation order as the
reference (so a given seed yields identical initial weights and its checkpoints load both ways),
but ``forward`` runs five fused kernels instead of eleven library calls:
conv+bias+ReLU+pool (x2), linear+bias+ReLU (x2), linear+bias.
"""
from __future__ import annotations

import torch
from torch import nn

from ..ops import functional as F


class MLModel(nn.Module):
    _mlt_native = True

    def __init__(self, num_classes: int = 10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 6, 5)
        self.pool = nn.MaxPool2d(2, 2)
        self.conv2 = nn.Conv2d(6, 16, 5)
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, 84)
        self.fc3 = nn.Linear(84, num_classes)

    def forward(self, x):
        x = F.conv_relu_pool(x, self.conv1.weight, self.conv1.bias, relu=True, pool=True)
        x = F.conv_relu_pool(x, self.conv2.weight, self.conv2.bias, relu=True, pool=True)
        x = x.view(-1, 16 * 5 * 5)
        x = F.linear_act(x, self.fc1.weight, self.fc1.bias, "relu")
        x = F.linear_act(x, self.fc2.weight, self.fc2.bias, "relu")
        x = F.linear_act(x, self.fc3.weight, self.fc3.bias, None)
        return x
