"""The reference's MNIST CNN (ref:main.py:20-45), same parameter/buffer names and shapes so
checkpoints interchange: conv1[32,1,3,3] conv2[64,32,3,3] fc1[128,9216] fc2[10,128] batchnorm[128].

Forward semantics are the reference's: conv-ReLU-conv-ReLU-maxpool2-Dropout2d(.25)-flatten-fc1-
BatchNorm1d-ReLU-Dropout2d(.5) (element-wise on a 2-D tensor, ref:main.py:42)-fc2-log_softmax.
On CUDA the layers run the hand-written kernels with these fusions: bias+ReLU in the conv epilogues,
ReLU in the BatchNorm kernel; ``fused_step`` (parallel/engine.py) goes further and captures the whole
training step in a CUDA graph.
"""
from __future__ import annotations

import torch
from torch import nn

from .. import ops


class ConvNet(nn.Module):
    def __init__(self):
        super().__init__()
        # registration order matches ref:main.py:23-29 so that seeded init draws the same numbers
        self.conv1 = ops.Conv2d(1, 32, 3, 1, fuse_relu=True)
        self.conv2 = ops.Conv2d(32, 64, 3, 1, fuse_relu=True)
        self.dropout1 = ops.Dropout2d(0.25)
        self.dropout2 = ops.Dropout2d(0.5)
        self.fc1 = ops.Linear(9216, 128)
        self.fc2 = ops.Linear(128, 10)
        self.batchnorm = ops.BatchNorm1d(128, fuse_relu=True)
        self.pool = ops.MaxPool2d(2, out_nchw=True)

    def features(self, x: torch.Tensor) -> torch.Tensor:
        """Everything up to the logits."""
        x = self.conv1(x)            # conv + bias + ReLU            (ref:main.py:32-33)
        x = self.conv2(x)            # conv + bias + ReLU            (ref:main.py:34-35)
        x = self.pool(x)             # max_pool2d(2)                 (ref:main.py:36)
        x = self.dropout1(x)         # channel dropout p=.25         (ref:main.py:37)
        x = torch.flatten(x, 1)      # NCHW flatten order            (ref:main.py:38)
        x = self.fc1(x)              #                               (ref:main.py:39)
        x = self.batchnorm(x)        # BatchNorm1d + ReLU            (ref:main.py:40-41)
        x = self.dropout2(x)         # element-wise dropout p=.5     (ref:main.py:42)
        return self.fc2(x)           #                               (ref:main.py:43)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.log_softmax(self.features(x), dim=1)   # ref:main.py:44
