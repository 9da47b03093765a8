"""`ModelManager`: the TensorFlow-side model table under the names its applications use.

Parity target: tensorflow_impl/libs/model.py:44-103 (`Small, MobileNetV2, Resnet50, Resnet200, VGG, DenseNet, Cifarnet,
Inception, CNN`). TensorFlow is not available here, so every entry is a torch module: the three
hand-written Keras nets are restated layer for layer (`Small`: flatten-128-classes; `CNN`: conv128-pool-conv64-pool-64-
classes; `Cifarnet`: two 5×5×64 "same" convolutions with 3×3 pooling, 384-192-classes) with logits as outputs (the
softmax lives in the loss), and the Keras application nets map to this framework's zoo / torchvision counterparts.
"""

from __future__ import annotations

import torch
from torch import nn

__all__ = ["ModelManager"]


class _Small(nn.Module):
    def __init__(self, in_shape, num_classes):
        super().__init__()
        features = 1
        for s in in_shape:
            features *= s
        self.net = nn.Sequential(nn.Flatten(), nn.Linear(features, 128), nn.ReLU(), nn.Linear(128, num_classes))

    def forward(self, x):
        return self.net(x)


def _flat_features(body, in_shape):
    with torch.no_grad():
        return body(torch.zeros(1, *in_shape)).numel()


class _KerasCNN(nn.Module):
    def __init__(self, in_shape, num_classes):
        super().__init__()
        self.body = nn.Sequential(nn.Conv2d(in_shape[0], 128, 3), nn.ReLU(), nn.MaxPool2d(2, ceil_mode=True),
                                  nn.Conv2d(128, 64, 3), nn.ReLU(), nn.MaxPool2d(2, ceil_mode=True), nn.Flatten())
        self.head = nn.Sequential(nn.Linear(_flat_features(self.body, in_shape), 64), nn.Linear(64, num_classes))

    def forward(self, x):
        return self.head(self.body(x))


class _KerasCifarnet(nn.Module):
    def __init__(self, in_shape, num_classes):
        super().__init__()
        self.body = nn.Sequential(nn.Conv2d(in_shape[0], 64, 5, padding=2), nn.ReLU(), nn.MaxPool2d(3),
                                  nn.Conv2d(64, 64, 5, padding=2), nn.ReLU(), nn.MaxPool2d(3), nn.Flatten())
        self.head = nn.Sequential(nn.Linear(_flat_features(self.body, in_shape), 384), nn.ReLU(), nn.Linear(384, 192), nn.ReLU(),
                                  nn.Linear(192, num_classes))
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.Linear)) and m is not self.head[-1]:
                nn.init.kaiming_uniform_(m.weight, nonlinearity="relu")      # kernel_initializer='he_uniform'

    def forward(self, x):
        return self.head(self.body(x))


class ModelManager:
    """`ModelManager(model, info).model` — `info` carries `num_classes` and the CHW `image_shape` (a dict or any object
    with those attributes), the role of the reference's tfds `info` object."""

    NAMES = ("Small", "MobileNetV2", "Resnet50", "Resnet200", "VGG", "DenseNet", "Cifarnet", "Inception", "CNN")

    def __init__(self, model, info):
        get = (lambda key, default: info.get(key, default)) if isinstance(info, dict) else (lambda key, default: getattr(info, key, default))
        self.num_classes = int(get("num_classes", 10))
        self.image_shape = tuple(get("image_shape", (3, 32, 32)))
        self.info = info
        assert model in self.NAMES, "Selected model not available"
        self.model = self._build(model)

    def _build(self, name):
        from ..models import cifar_zoo as zoo
        shape, classes = self.image_shape, self.num_classes
        if name == "Small":
            return _Small(shape, classes)
        if name == "CNN":
            return _KerasCNN(shape, classes)
        if name == "Cifarnet":
            return _KerasCifarnet(shape, classes)
        if name == "MobileNetV2":
            return zoo.MobileNetV2(num_classes=classes)
        if name == "DenseNet":
            return zoo.DenseNet121(num_classes=classes)
        import torchvision.models as tv
        if name == "Resnet50":
            return tv.resnet50(num_classes=classes)
        if name == "Resnet200":          # the reference builds ResNet152V2 under this name (model.py:53)
            return tv.resnet152(num_classes=classes)
        if name == "VGG":
            return tv.vgg16(num_classes=classes)
        return tv.inception_v3(num_classes=classes, aux_logits=False, init_weights=False)   # "Inception"
