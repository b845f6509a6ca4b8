"""LeNet-style MNIST CNN (architecture parity: lab/tutorial_1a/hfl_complete.py:39-64).

conv(1->32,3) ReLU conv(32->64,3) ReLU maxpool2 dropout.25 flatten fc(9216->128) ReLU
dropout.5 fc(128->10) log_softmax.  1 199 882 parameters, no buffers.  Parameter names
(`conv1`, `conv2`, `fc1`, `fc2`) match the reference so its ``state_dict`` checkpoints load.

This ``nn.Module`` is the *oracle and CPU plumbing* implementation.  On a B200 the packed
client engine (``fl/packed.py``) trains many of these at once with the hand-written grouped
kernels in ``csrc/`` over the same flat parameter layout (`MNIST_LAYOUT`).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

# (name, shape) in nn.Module.parameters() order -- the flat arena layout the kernels assume.
MNIST_LAYOUT = [
    ("conv1.weight", (32, 1, 3, 3)), ("conv1.bias", (32,)),
    ("conv2.weight", (64, 32, 3, 3)), ("conv2.bias", (64,)),
    ("fc1.weight", (128, 9216)), ("fc1.bias", (128,)),
    ("fc2.weight", (10, 128)), ("fc2.bias", (10,)),
]
MNIST_NUM_PARAMS = 1_199_882


class MnistCnn(nn.Module):
    def __init__(self, p_drop1: float = 0.25, p_drop2: float = 0.5):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 32, kernel_size=3)
        self.conv2 = nn.Conv2d(32, 64, kernel_size=3)
        self.fc1 = nn.Linear(64 * 12 * 12, 128)
        self.fc2 = nn.Linear(128, 10)
        self.dropout1 = nn.Dropout(p_drop1)
        self.dropout2 = nn.Dropout(p_drop2)

    def features(self, x: torch.Tensor) -> torch.Tensor:
        h = F.relu(self.conv2(F.relu(self.conv1(x))))
        return self.dropout1(F.max_pool2d(h, 2)).flatten(1)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        h = self.dropout2(F.relu(self.fc1(self.features(x))))
        return F.log_softmax(self.fc2(h), dim=1)
