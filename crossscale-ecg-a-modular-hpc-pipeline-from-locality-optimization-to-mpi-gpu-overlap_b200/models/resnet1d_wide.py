"""Widened ResNet-1D ECG classifier (~24 M parameters at the default widths).

Not in the reference; it is the bandwidth-bound aggregation stress model named by
``BASELINE.json`` ("Widened ResNet-1D ECG model (~25M params) FedAvg over 8xB200").  Convolutions go
through ``ops.conv1d`` (tcgen05 implicit GEMM on CUDA, ``F.conv1d`` on CPU); BN + ReLU (+ residual)
go through ``ops.norm_act`` which is one fused kernel on CUDA.
"""
from __future__ import annotations

from typing import Sequence

import torch
import torch.nn as nn

from ..ops.conv1d import Conv1dB200
from ..ops.norm_act import BNReLU1d, avg_pool2, conv_bn, global_avg_pool1d


class BasicBlock1D(nn.Module):
    """conv-BN-ReLU-conv-BN(+skip)-ReLU; ``down`` halves the length with a 2:1 average pool first (all
    convolutions stay stride 1, which is what the shifted-descriptor implicit GEMM implements)."""

    def __init__(self, cin: int, cout: int, k: int = 7, down: bool = False):
        super().__init__()
        self.pool = down
        self.conv1 = Conv1dB200(cin, cout, k, padding=k // 2, bias=False)
        self.bn1 = BNReLU1d(cout, relu=True)
        self.conv2 = Conv1dB200(cout, cout, k, padding=k // 2, bias=False)
        self.bn2 = BNReLU1d(cout, relu=True)  # ReLU applied after the residual add (fused)
        self.down = None
        if cin != cout:
            self.down = Conv1dB200(cin, cout, 1, padding=0, bias=False)
            self.down_bn = BNReLU1d(cout, relu=False)

    def forward(self, x):
        if self.pool:
            x = avg_pool2(x)
        idn = x if self.down is None else conv_bn(self.down, self.down_bn, x)
        y = conv_bn(self.conv1, self.bn1, x)
        return conv_bn(self.conv2, self.bn2, y, residual=idn)


class ResNet1DWide(nn.Module):
    def __init__(self, num_classes: int = 5, widths: Sequence[int] = (128, 256, 512, 768),
                 blocks_per_stage: int = 2, stem: int = 64, k: int = 7):
        super().__init__()
        self.stem = Conv1dB200(1, stem, 7, padding=3, bias=False)
        self.stem_bn = BNReLU1d(stem, relu=True)
        layers, cin = [], stem
        for i, w in enumerate(widths):
            for b in range(blocks_per_stage):
                layers.append(BasicBlock1D(cin, w, k=k, down=(b == 0 and i > 0)))
                cin = w
        self.layers = nn.Sequential(*layers)
        self.head = nn.Linear(cin, num_classes)

    def forward(self, x):
        y = self.stem_bn(self.stem(x))
        y = self.layers(y)
        return self.head(global_avg_pool1d(y))


def resnet1d_wide_25m(num_classes: int = 5) -> ResNet1DWide:
    return ResNet1DWide(num_classes=num_classes)


def resnet1d_small(num_classes: int = 5) -> ResNet1DWide:
    """Small variant for tests."""
    return ResNet1DWide(num_classes=num_classes, widths=(16, 32), blocks_per_stage=1, stem=16, k=5)
