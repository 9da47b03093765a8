"""`Net` and `Cifarnet` running entirely on this framework's hand-written sm_100a kernels
(conv / max-pool / ReLU-fused epilogues / tcgen05 GEMM for the linear layers) — same parameters,
`state_dict` keys and flat layout as the torch-op versions in `small.py`, so they are drop-in
replacements (`select_model("cifarnet-native" | "convnet-native", ...)`). Off-GPU every op falls
back to its torch equivalent."""

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import nn as gnn

__all__ = ["NativeCifarnet", "NativeNet"]


class NativeCifarnet(nn.Module):
    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 6, 5)
        self.pool = nn.MaxPool2d(2, 2)
        self.conv2 = nn.Conv2d(6, 16, 5)
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, 84)
        self.fc3 = nn.Linear(84, num_classes)

    def forward(self, x):
        x = gnn.max_pool2d(gnn.conv2d(x, self.conv1.weight, self.conv1.bias, relu=True), 2)
        x = gnn.max_pool2d(gnn.conv2d(x, self.conv2.weight, self.conv2.bias, relu=True), 2)
        x = gnn.linear(x.flatten(1), self.fc1.weight, self.fc1.bias, relu=True)
        x = gnn.linear(x, self.fc2.weight, self.fc2.bias, relu=True)
        return gnn.linear(x, self.fc3.weight, self.fc3.bias)


class NativeNet(nn.Module):
    """MNIST convnet; dropout masks come from torch's Philox generator (elementwise, no library GEMM/conv)."""

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 10, kernel_size=5)
        self.conv2 = nn.Conv2d(10, 20, kernel_size=5)
        self.conv2_drop = nn.Dropout2d()
        self.fc1 = nn.Linear(320, 50)
        self.fc2 = nn.Linear(50, num_classes)

    def forward(self, x):
        x = torch.relu(gnn.max_pool2d(gnn.conv2d(x, self.conv1.weight, self.conv1.bias), 2))
        x = torch.relu(gnn.max_pool2d(self.conv2_drop(gnn.conv2d(x, self.conv2.weight, self.conv2.bias)), 2))
        x = gnn.linear(x.flatten(1), self.fc1.weight, self.fc1.bias, relu=True)
        x = F.dropout(x, training=self.training)
        return F.log_softmax(gnn.linear(x, self.fc2.weight, self.fc2.bias), dim=1)
