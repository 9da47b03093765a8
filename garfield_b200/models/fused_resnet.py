"""Route the BatchNorm / ReLU / residual-add epilogues of torchvision-style ResNets through the fused
channels-last kernels (`ops.nn.fused_batch_norm`): per bottleneck block 3 BatchNorm + 3 ReLU + 1 add
launches become 3 launches forward and 3 backward. Parameters, buffers and `state_dict` keys are untouched —
only `forward` of the block (and the stem) is replaced, and any situation the fused kernel does not
cover (eval mode, NCHW tensors, CPU) falls back to the standard ops inside `fused_batch_norm`."""

import types

import torch

from ..ops.nn import fused_batch_norm

__all__ = ["fuse_resnet_blocks"]


def _bottleneck_forward(self, x):
    out = fused_batch_norm(self.conv1(x), self.bn1, relu=True)
    out = fused_batch_norm(self.conv2(out), self.bn2, relu=True)
    identity = x
    if self.downsample is not None:
        identity = fused_batch_norm(self.downsample[0](x), self.downsample[1], relu=False)
    return fused_batch_norm(self.conv3(out), self.bn3, residual=identity, relu=True)


def _basic_forward(self, x):
    out = fused_batch_norm(self.conv1(x), self.bn1, relu=True)
    identity = x
    if self.downsample is not None:
        identity = fused_batch_norm(self.downsample[0](x), self.downsample[1], relu=False)
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
