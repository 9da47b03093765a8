"""Small reference networks: MNIST `Net`, `Cifarnet`, `CNN`/`CNNet`, `PimaNet`, `LeNet`.

Architectures (layer shapes, parameter order, attribute names → identical `state_dict` keys and
flat-vector layout) follow /root/reference/pytorch_impl/libs/garfieldpp/models/nets.py:40-144,
pimanet.py:4-18 and lenet.py. Parameter counts with 10 classes: Net 21 840, Cifarnet 62 006,
CNN 5 852 170, PimaNet (1 output) 4 801.
"""

import torch
import torch.nn as nn
import torch.nn.functional as F

__all__ = ["Net", "Cifarnet", "CNN", "CNNet", "PimaNet", "LeNet"]


class Net(nn.Module):
    """MNIST convnet: 2 conv(5×5) + 2 FC, log-softmax output (pair it with the `nll` loss)."""

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 10, kernel_size=5)
        self.conv2 = nn.Conv2d(10, 20, kernel_size=5)
        self.conv2_drop = nn.Dropout2d()
        self.fc1 = nn.Linear(320, 50)
        self.fc2 = nn.Linear(50, num_classes)
        # parameter-free layers as modules (no state_dict keys): the device engines route modules to this library's kernels
        self.pool, self.relu = nn.MaxPool2d(2), nn.ReLU()

    def forward(self, x):
        x = self.relu(self.pool(self.conv1(x)))
        x = self.relu(self.pool(self.conv2_drop(self.conv2(x))))
        x = self.relu(self.fc1(x.flatten(1)))
        x = F.dropout(x, training=self.training)
        return F.log_softmax(self.fc2(x), dim=1)


class Cifarnet(nn.Module):
    """LeNet-style CIFAR-10 network (the PyTorch tutorial classifier)."""

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 6, 5)
        self.pool = nn.MaxPool2d(2, 2)
        self.conv2 = nn.Conv2d(6, 16, 5)
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, 84)
        self.fc3 = nn.Linear(84, num_classes)
        self.relu = nn.ReLU()          # a module, so that the device engines can route it to this library's kernel

    def forward(self, x):
        x = self.pool(self.relu(self.conv1(x)))
        x = self.pool(self.relu(self.conv2(x)))
        x = self.relu(self.fc1(x.flatten(1)))
        x = self.relu(self.fc2(x))
        return self.fc3(x)


def _conv_stage(cin, cmid, cout, dropout=None):
    layers = [nn.Conv2d(cin, cmid, 3, padding=1), nn.BatchNorm2d(cmid), nn.ReLU(inplace=True),
              nn.Conv2d(cmid, cout, 3, padding=1), nn.ReLU(inplace=True), nn.MaxPool2d(2, 2)]
    if dropout:
        layers.append(nn.Dropout2d(p=dropout))
    return layers


class CNN(nn.Module):
    """Six-conv CIFAR network with three FC layers (5.85 M parameters)."""

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv_layer = nn.Sequential(*_conv_stage(3, 32, 64), *_conv_stage(64, 128, 128, dropout=0.05),
                                        *_conv_stage(128, 256, 256))
        self.fc_layer = nn.Sequential(nn.Dropout(p=0.1), nn.Linear(4096, 1024), nn.ReLU(inplace=True),
                                      nn.Linear(1024, 512), nn.ReLU(inplace=True), nn.Dropout(p=0.1),
                                      nn.Linear(512, num_classes))

    def forward(self, x):
        return self.fc_layer(self.conv_layer(x).flatten(1))


def CNNet(num_classes=10):
    return CNN(num_classes=num_classes)


class PimaNet(nn.Module):
    """8-64-64-1 sigmoid MLP for the Pima diabetes table (binary cross-entropy loss)."""

    def __init__(self, num_classes=1):
        super().__init__()
        self.net = nn.Sequential(nn.Linear(8, 64), nn.ReLU(), nn.Linear(64, 64), nn.ReLU(),
                                 nn.Linear(64, num_classes), nn.Sigmoid())

    def forward(self, x):
        return self.net(x)


class LeNet(nn.Module):
    """Classic LeNet-5 on 3×32×32 inputs."""

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 6, 5)
        self.conv2 = nn.Conv2d(6, 16, 5)
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, 84)
        self.fc3 = nn.Linear(84, num_classes)
        self.pool, self.relu = nn.MaxPool2d(2), nn.ReLU()

    def forward(self, x):
        x = self.pool(self.relu(self.conv1(x)))
        x = self.pool(self.relu(self.conv2(x)))
        x = self.relu(self.fc1(x.flatten(1)))
        x = self.relu(self.fc2(x))
        return self.fc3(x)
