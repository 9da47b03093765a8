"""Name → object factories and helpers of the `garfieldpp` library.

API parity with /root/reference/pytorch_impl/libs/garfieldpp/tools.py:47-193: `select_loss`,
`select_model`, `select_optimizer`, `adjust_learning_rate`, `get_bytes_com`, `convert_to_gbit`,
the RPC helpers `_call_method/_remote_method_sync/_remote_method_async`, and the process-global
`server_instance` / `worker_instance` with the blocking `get_server()` / `get_worker()`.

B200-native differences: `select_model` never wraps in `DataParallel` (one process per GPU);
`select_optimizer` returns a `FlatOptimizer` (one fused kernel over the flat arena) when given a
`FlatModel`, and the stock `torch.optim` class when given a plain module.
"""

from __future__ import annotations

import threading
import time

import torch
import torch.nn as nn

from .. import models as _zoo
from ..ops.optim import FlatOptimizer, OPTIMIZERS
from ..parallel.arena import FlatModel

__all__ = ["select_loss", "select_model", "select_optimizer", "adjust_learning_rate", "get_bytes_com",
           "convert_to_gbit", "get_server", "get_worker", "set_server", "set_worker", "model_names",
           "num_classes_of", "_call_method", "_remote_method_sync", "_remote_method_async"]

_LOSSES = {"nll": nn.NLLLoss, "cross-entropy": nn.CrossEntropyLoss, "binary-cross-entropy": nn.BCELoss}
_NUM_CLASSES = {"cifar10": 10, "cifar100": 100, "mnist": 10, "imagenet": 1000, "pima": 1}


def _tv(name):
    def build(num_classes):
        import torchvision
        ctor = getattr(torchvision.models, name)
        if name == "inception_v3":
            return ctor(num_classes=num_classes, init_weights=True)
        return ctor(num_classes=num_classes)
    return build


_MODELS = {
    "convnet": _zoo.Net, "cifarnet": _zoo.Cifarnet, "cnn": _zoo.CNNet, "resnet18": _zoo.ResNet18,
    # ImageNet-style torchvision variants, as in the reference (tools.py:69-75)
    "resnet34": _tv("resnet34"), "resnet50": _tv("resnet50"), "resnet152": _tv("resnet152"),
    "inception": _tv("inception_v3"), "vgg16": _tv("vgg16"), "vgg19": _tv("vgg19"),
    "preactresnet18": _zoo.PreActResNet18, "googlenet": _zoo.GoogLeNet, "densenet121": _zoo.DenseNet121,
    "resnext29": _zoo.ResNeXt29_2x64d, "mobilenet": _zoo.MobileNet, "mobilenetv2": _zoo.MobileNetV2,
    "dpn92": _zoo.DPN92, "shufflenetg2": _zoo.ShuffleNetG2, "senet18": _zoo.SENet18,
    "efficientnetb0": _zoo.EfficientNetB0, "regnetx200": _zoo.RegNetX_200MF, "pimanet": _zoo.PimaNet,
    # extras beyond the reference's table (local CIFAR-style variants and the rest of the zoo)
    "resnet50-cifar": _zoo.ResNet50, "resnet34-cifar": _zoo.ResNet34, "vgg16-cifar": _zoo.VGG16, "lenet": _zoo.LeNet,
    "cifarnet-native": _zoo.NativeCifarnet, "convnet-native": _zoo.NativeNet,  # hand-written sm_100a kernels end to end
    "shufflenetv2": _zoo.ShuffleNetV2, "pnasneta": _zoo.PNASNetA, "regnety400": _zoo.RegNetY_400MF,
}


def model_names():
    return tuple(_MODELS)


def num_classes_of(dataset):
    base = dataset[len("synthetic-"):] if dataset.startswith("synthetic-") else dataset
    if base not in _NUM_CLASSES:
        raise KeyError("The specified dataset is undefined, available datasets are: %s" % list(_NUM_CLASSES))
    return _NUM_CLASSES[base]


def select_loss(loss_fn):
    """Loss module by name: nll, cross-entropy, binary-cross-entropy."""
    if loss_fn not in _LOSSES:
        raise KeyError("The selected loss function is undefined, available losses are: %s" % list(_LOSSES))
    return _LOSSES[loss_fn]()


def select_model(model, device, dataset):
    """Instantiate `model` with the class count of `dataset` on `device`."""
    if model not in _MODELS:
        raise KeyError("The specified model is undefined, available models are: %s" % list(_MODELS))
    return _MODELS[model](num_classes=num_classes_of(dataset)).to(device)


def select_optimizer(model, optimizer, *args, **kwargs):
    """Optimizer by name (sgd, adam, adamw, rmsprop, adagrad) for a module or a `FlatModel`."""
    if optimizer not in OPTIMIZERS:
        raise KeyError("The selected optimizer is undefined, the available optimizers are: %s" % list(OPTIMIZERS))
    if isinstance(model, FlatModel):
        return FlatOptimizer(model.flat, optimizer, *args, **kwargs)
    table = {"sgd": torch.optim.SGD, "adam": torch.optim.Adam, "adamw": torch.optim.AdamW,
             "rmsprop": torch.optim.RMSprop, "adagrad": torch.optim.Adagrad}
    return table[optimizer](model.parameters(), *args, **kwargs)


def adjust_learning_rate(optimizer, lr):
    for group in optimizer.param_groups:
        group["lr"] = lr


def get_bytes_com():
    """Bytes sent + received on all NICs so far (psutil), 0 when psutil is missing."""
    try:
        import psutil
    except ImportError:
        return 0
    io = psutil.net_io_counters()
    return io.bytes_sent + io.bytes_recv


def convert_to_gbit(value):
    return value / 1024. / 1024. / 1024. * 8


# -- process-global role instances (looked up by remote peers) ------------------------------------

server_instance, worker_instance = None, None
_registered = threading.Condition()


def set_server(obj):
    global server_instance
    with _registered:
        server_instance = obj
        _registered.notify_all()


def set_worker(obj):
    global worker_instance
    with _registered:
        worker_instance = obj
        _registered.notify_all()


def get_server(timeout=None):
    """The `Server` created in this process (blocks until it exists)."""
    with _registered:
        _registered.wait_for(lambda: server_instance is not None, timeout)
        return server_instance


def get_worker(timeout=None):
    """The `Worker` created in this process (blocks until it exists)."""
    with _registered:
        _registered.wait_for(lambda: worker_instance is not None, timeout)
        return worker_instance


# -- RPC helpers (compatibility transport) -----------------------------------------------------------

def _call_method(method, rref, *args, **kwargs):
    return method(rref.local_value(), *args, **kwargs)


def _remote_method_sync(method, rref, *args, **kwargs):
    from torch.distributed.rpc import rpc_sync
    return rpc_sync(rref.owner(), _call_method, args=[method, rref] + list(args), kwargs=kwargs)


def _remote_method_async(method, rref, *args, **kwargs):
    from torch.distributed.rpc import rpc_async
    return rpc_async(rref.owner(), _call_method, args=[method, rref] + list(args), kwargs=kwargs)
