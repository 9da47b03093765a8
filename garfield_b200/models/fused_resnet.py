"""Route the BatchNorm / ReLU / residual-add epilogues of torchvision-style ResNets through the fused
channels-last kernels (`ops.nn.fused_batch_norm`): per bottleneck block 3 BatchNorm + 3 ReLU + 1 add
launches become 3 launches forward and 3 backward. Parameters, buffers and `state_dict` keys are untouched —
only `forward` of the block (and the stem) is replaced, and any situation the fused kernel does not
cover (eval mode, NCHW tensors, CPU) falls back to the standard ops inside `fused_batch_norm`."""

import types

import torch

from ..ops.nn import fused_batch_norm

__all__ = ["fuse_resnet_blocks"]

import os

# GARFIELD_FORK_SKIP=1: sum the skip-connection gradient inside conv1's data-gradient epilogue (one add kernel less per block).
# Off by default: measured SLOWER on the flagship (2.40 vs 2.32 ms per step at 8 workers on one B200) although it removes 12
# launches — the data-gradient kernel then waits for the skip gradient's producer as well, which shortens the overlap window of the
# side-stream weight gradients. Numerically identical to the separate add (tests/test_grouped.py).
FORK_SKIP = [os.environ.get("GARFIELD_FORK_SKIP", "0") != "0"]


def _first_conv(block, x):
    """(conv1(x), identity). Without a downsample branch the identity leaves through conv1's autograd node when the worker-fusion
    layer offers that (`_gb_fork`): the skip gradient is then added inside conv1's data-gradient kernel, not by an add kernel."""
    fork = getattr(block.conv1, "_gb_fork", None)
    if block.downsample is None and fork is not None and FORK_SKIP[0]:
        return fork(x)
    split = getattr(block.conv1, "_gb_split", None)      # worker-fusion layer on its own kernels: the two gradients are summed by them
    if split is not None and x.requires_grad:
        a, b = split(x)
        return block.conv1(a), b
    return block.conv1(x), x


def _bottleneck_forward(self, x):
    y1, identity = _first_conv(self, x)
    out = fused_batch_norm(y1, self.bn1, relu=True)
    out = fused_batch_norm(self.conv2(out), self.bn2, relu=True)
    if self.downsample is not None:
        identity = fused_batch_norm(self.downsample[0](identity), self.downsample[1], relu=False)   # `identity` is x (or its fork view)
    return fused_batch_norm(self.conv3(out), self.bn3, residual=identity, relu=True)


def _basic_forward(self, x):
    y1, identity = _first_conv(self, x)
    out = fused_batch_norm(y1, self.bn1, relu=True)
    if self.downsample is not None:
        identity = fused_batch_norm(self.downsample[0](identity), self.downsample[1], relu=False)   # `identity` is x (or its fork view)
    return fused_batch_norm(self.conv2(out), self.bn2, residual=identity, relu=True)


def _resnet_forward(self, x):
    x = fused_batch_norm(self.conv1(x), self.bn1, relu=True)
    x = self.maxpool(x)
    x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
    return self.fc(torch.flatten(self.avgpool(x), 1))


def fuse_resnet_blocks(module):
    """Patch a torchvision ResNet (resnet18/34/50/101/152) in place; returns the number of patched blocks."""
    try:
        from torchvision.models.resnet import BasicBlock, Bottleneck, ResNet
    except ImportError:
        return 0
    patched = 0
    for m in module.modules():
        simple_downsample = m.downsample is None or (isinstance(m.downsample, torch.nn.Sequential) and len(m.downsample) == 2
                                                     and isinstance(m.downsample[1], torch.nn.BatchNorm2d)) if hasattr(m, "downsample") else False
        if isinstance(m, Bottleneck) and simple_downsample:
            m.forward = types.MethodType(_bottleneck_forward, m)
            patched += 1
        elif isinstance(m, BasicBlock) and simple_downsample:
            m.forward = types.MethodType(_basic_forward, m)
            patched += 1
    if isinstance(module, ResNet) and isinstance(module.bn1, torch.nn.BatchNorm2d):
        module.forward = types.MethodType(_resnet_forward, module)
    return patched
