"""CIFAR-style model zoo (32×32 inputs): ResNet, PreActResNet, VGG, GoogLeNet, DenseNet,
ResNeXt-29, MobileNet(V2), DPN, ShuffleNet(V2), SENet, EfficientNet-B0, RegNet, PNASNet.

Same family list, constructor names and parameter counts as the reference zoo
(/root/reference/pytorch_impl/libs/garfieldpp/models/*.py); written as
compact config-driven builders rather than one file per family. All models take
`num_classes` and return logits.
"""

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

__all__ = [
    "ResNet18", "ResNet34", "ResNet50", "ResNet101", "ResNet152",
    "PreActResNet18", "PreActResNet34", "PreActResNet50", "PreActResNet101", "PreActResNet152",
    "VGG", "VGG11", "VGG13", "VGG16", "VGG19", "GoogLeNet",
    "DenseNet121", "DenseNet169", "DenseNet201", "DenseNet161", "densenet_cifar",
    "ResNeXt29_2x64d", "ResNeXt29_4x64d", "ResNeXt29_8x64d", "ResNeXt29_32x4d",
    "MobileNet", "MobileNetV2", "DPN26", "DPN92", "ShuffleNetG2", "ShuffleNetG3", "ShuffleNetV2",
    "SENet18", "EfficientNetB0", "RegNetX_200MF", "RegNetX_400MF", "RegNetY_400MF", "PNASNetA", "PNASNetB",
]


def conv_bn(cin, cout, k, stride=1, padding=0, groups=1):
    return nn.Conv2d(cin, cout, k, stride=stride, padding=padding, groups=groups, bias=False), nn.BatchNorm2d(cout)


def _projection(cin, cout, stride):
    """1×1 conv + BN shortcut when the shape changes, identity otherwise."""
    if stride != 1 or cin != cout:
        return nn.Sequential(*conv_bn(cin, cout, 1, stride=stride))
    return nn.Sequential()


# --------------------------------------------------------------------------------- ResNet ---

class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, in_planes, planes, stride=1):
        super().__init__()
        self.conv1, self.bn1 = conv_bn(in_planes, planes, 3, stride, 1)
        self.conv2, self.bn2 = conv_bn(planes, planes, 3, 1, 1)
        self.shortcut = _projection(in_planes, planes * self.expansion, stride)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return F.relu(out + self.shortcut(x))


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, in_planes, planes, stride=1):
        super().__init__()
        self.conv1, self.bn1 = conv_bn(in_planes, planes, 1)
        self.conv2, self.bn2 = conv_bn(planes, planes, 3, stride, 1)
        self.conv3, self.bn3 = conv_bn(planes, planes * self.expansion, 1)
        self.shortcut = _projection(in_planes, planes * self.expansion, stride)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = F.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        return F.relu(out + self.shortcut(x))


def _stack(block, in_planes, planes, count, stride):
    blocks = []
    for s in [stride] + [1] * (count - 1):
        blocks.append(block(in_planes, planes, s))
        in_planes = planes * block.expansion
    return nn.Sequential(*blocks), in_planes


class ResNet(nn.Module):
    def __init__(self, block, num_blocks, num_classes=10):
        super().__init__()
        self.conv1, self.bn1 = conv_bn(3, 64, 3, 1, 1)
        width = 64
        self.layer1, width = _stack(block, width, 64, num_blocks[0], 1)
        self.layer2, width = _stack(block, width, 128, num_blocks[1], 2)
        self.layer3, width = _stack(block, width, 256, num_blocks[2], 2)
        self.layer4, width = _stack(block, width, 512, num_blocks[3], 2)
        self.linear = nn.Linear(width, num_classes)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.layer4(self.layer3(self.layer2(self.layer1(out))))
        return self.linear(F.avg_pool2d(out, 4).flatten(1))


def ResNet18(num_classes=10): return ResNet(BasicBlock, [2, 2, 2, 2], num_classes)
def ResNet34(num_classes=10): return ResNet(BasicBlock, [3, 4, 6, 3], num_classes)
def ResNet50(num_classes=10): return ResNet(Bottleneck, [3, 4, 6, 3], num_classes)
def ResNet101(num_classes=10): return ResNet(Bottleneck, [3, 4, 23, 3], num_classes)
def ResNet152(num_classes=10): return ResNet(Bottleneck, [3, 8, 36, 3], num_classes)


# -------------------------------------------------------------------------- PreActResNet ---

class PreActBlock(nn.Module):
    expansion = 1

    def __init__(self, in_planes, planes, stride=1):
        super().__init__()
        self.bn1 = nn.BatchNorm2d(in_planes)
        self.conv1 = nn.Conv2d(in_planes, planes, 3, stride, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        if stride != 1 or in_planes != planes * self.expansion:
            self.shortcut = nn.Sequential(nn.Conv2d(in_planes, planes * self.expansion, 1, stride, bias=False))

    def forward(self, x):
        out = F.relu(self.bn1(x))
        skip = self.shortcut(out) if hasattr(self, "shortcut") else x
        out = self.conv1(out)
        out = self.conv2(F.relu(self.bn2(out)))
        return out + skip


class PreActBottleneck(nn.Module):
    expansion = 4

    def __init__(self, in_planes, planes, stride=1):
        super().__init__()
        self.bn1 = nn.BatchNorm2d(in_planes)
        self.conv1 = nn.Conv2d(in_planes, planes, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, planes * self.expansion, 1, bias=False)
        if stride != 1 or in_planes != planes * self.expansion:
            self.shortcut = nn.Sequential(nn.Conv2d(in_planes, planes * self.expansion, 1, stride, bias=False))

    def forward(self, x):
        out = F.relu(self.bn1(x))
        skip = self.shortcut(out) if hasattr(self, "shortcut") else x
        out = self.conv1(out)
        out = self.conv2(F.relu(self.bn2(out)))
        out = self.conv3(F.relu(self.bn3(out)))
        return out + skip


class PreActResNet(nn.Module):
    def __init__(self, block, num_blocks, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 64, 3, 1, 1, bias=False)
        width = 64
        self.layer1, width = _stack(block, width, 64, num_blocks[0], 1)
        self.layer2, width = _stack(block, width, 128, num_blocks[1], 2)
        self.layer3, width = _stack(block, width, 256, num_blocks[2], 2)
        self.layer4, width = _stack(block, width, 512, num_blocks[3], 2)
        self.linear = nn.Linear(width, num_classes)

    def forward(self, x):
        out = self.layer4(self.layer3(self.layer2(self.layer1(self.conv1(x)))))
        return self.linear(F.avg_pool2d(out, 4).flatten(1))


def PreActResNet18(num_classes=10): return PreActResNet(PreActBlock, [2, 2, 2, 2], num_classes)
def PreActResNet34(num_classes=10): return PreActResNet(PreActBlock, [3, 4, 6, 3], num_classes)
def PreActResNet50(num_classes=10): return PreActResNet(PreActBottleneck, [3, 4, 6, 3], num_classes)
def PreActResNet101(num_classes=10): return PreActResNet(PreActBottleneck, [3, 4, 23, 3], num_classes)
def PreActResNet152(num_classes=10): return PreActResNet(PreActBottleneck, [3, 8, 36, 3], num_classes)


# ------------------------------------------------------------------------------------ VGG ---

_VGG_PLANS = {
    "VGG11": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG13": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG16": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "VGG19": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"],
}


class VGG(nn.Module):
    def __init__(self, vgg_name, num_classes=10):
        super().__init__()
        layers, cin = [], 3
        for item in _VGG_PLANS[vgg_name]:
            if item == "M":
                layers.append(nn.MaxPool2d(2, 2))
            else:
                layers += [nn.Conv2d(cin, item, 3, padding=1), nn.BatchNorm2d(item), nn.ReLU(inplace=True)]
                cin = item
        layers.append(nn.AvgPool2d(1, 1))
        self.features = nn.Sequential(*layers)
        self.classifier = nn.Linear(512, num_classes)

    def forward(self, x):
        return self.classifier(self.features(x).flatten(1))


def VGG11(num_classes=10): return VGG("VGG11", num_classes)
def VGG13(num_classes=10): return VGG("VGG13", num_classes)
def VGG16(num_classes=10): return VGG("VGG16", num_classes)
def VGG19(num_classes=10): return VGG("VGG19", num_classes)


# ------------------------------------------------------------------------------ GoogLeNet ---

def _cbr(cin, cout, k, padding=0):
    return [nn.Conv2d(cin, cout, k, padding=padding), nn.BatchNorm2d(cout), nn.ReLU(True)]


class Inception(nn.Module):
    def __init__(self, in_planes, n1x1, n3x3red, n3x3, n5x5red, n5x5, pool_planes):
        super().__init__()
        self.b1 = nn.Sequential(*_cbr(in_planes, n1x1, 1))
        self.b2 = nn.Sequential(*_cbr(in_planes, n3x3red, 1), *_cbr(n3x3red, n3x3, 3, 1))
        self.b3 = nn.Sequential(*_cbr(in_planes, n5x5red, 1), *_cbr(n5x5red, n5x5, 3, 1), *_cbr(n5x5, n5x5, 3, 1))
        self.b4 = nn.Sequential(nn.MaxPool2d(3, stride=1, padding=1), *_cbr(in_planes, pool_planes, 1))

    def forward(self, x):
        return torch.cat([self.b1(x), self.b2(x), self.b3(x), self.b4(x)], 1)


class GoogLeNet(nn.Module):
    def __init__(self, num_classes=10):
        super().__init__()
        self.pre_layers = nn.Sequential(*_cbr(3, 192, 3, 1))
        self.a3 = Inception(192, 64, 96, 128, 16, 32, 32)
        self.b3 = Inception(256, 128, 128, 192, 32, 96, 64)
        self.maxpool = nn.MaxPool2d(3, stride=2, padding=1)
        self.a4 = Inception(480, 192, 96, 208, 16, 48, 64)
        self.b4 = Inception(512, 160, 112, 224, 24, 64, 64)
        self.c4 = Inception(512, 128, 128, 256, 24, 64, 64)
        self.d4 = Inception(512, 112, 144, 288, 32, 64, 64)
        self.e4 = Inception(528, 256, 160, 320, 32, 128, 128)
        self.a5 = Inception(832, 256, 160, 320, 32, 128, 128)
        self.b5 = Inception(832, 384, 192, 384, 48, 128, 128)
        self.avgpool = nn.AvgPool2d(8, stride=1)
        self.linear = nn.Linear(1024, num_classes)

    def forward(self, x):
        out = self.b3(self.a3(self.pre_layers(x)))
        out = self.maxpool(out)
        out = self.e4(self.d4(self.c4(self.b4(self.a4(out)))))
        out = self.maxpool(out)
        out = self.avgpool(self.b5(self.a5(out)))
        return self.linear(out.flatten(1))


# ------------------------------------------------------------------------------- DenseNet ---

class DenseLayer(nn.Module):
    def __init__(self, in_planes, growth_rate):
        super().__init__()
        self.bn1 = nn.BatchNorm2d(in_planes)
        self.conv1 = nn.Conv2d(in_planes, 4 * growth_rate, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(4 * growth_rate)
        self.conv2 = nn.Conv2d(4 * growth_rate, growth_rate, 3, padding=1, bias=False)

    def forward(self, x):
        out = self.conv1(F.relu(self.bn1(x)))
        out = self.conv2(F.relu(self.bn2(out)))
        return torch.cat([out, x], 1)


class Transition(nn.Module):
    def __init__(self, in_planes, out_planes):
        super().__init__()
        self.bn = nn.BatchNorm2d(in_planes)
        self.conv = nn.Conv2d(in_planes, out_planes, 1, bias=False)

    def forward(self, x):
        return F.avg_pool2d(self.conv(F.relu(self.bn(x))), 2)


class DenseNet(nn.Module):
    def __init__(self, nblocks, growth_rate=12, reduction=0.5, num_classes=10):
        super().__init__()
        planes = 2 * growth_rate
        self.conv1 = nn.Conv2d(3, planes, 3, padding=1, bias=False)
        for stage, count in enumerate(nblocks, start=1):
            layers = []
            for _ in range(count):
                layers.append(DenseLayer(planes, growth_rate))
                planes += growth_rate
            setattr(self, "dense%d" % stage, nn.Sequential(*layers))
            if stage < len(nblocks):
                out_planes = int(math.floor(planes * reduction))
                setattr(self, "trans%d" % stage, Transition(planes, out_planes))
                planes = out_planes
        self.bn = nn.BatchNorm2d(planes)
        self.linear = nn.Linear(planes, num_classes)

    def forward(self, x):
        out = self.conv1(x)
        out = self.trans1(self.dense1(out))
        out = self.trans2(self.dense2(out))
        out = self.trans3(self.dense3(out))
        out = self.dense4(out)
        out = F.avg_pool2d(F.relu(self.bn(out)), 4)
        return self.linear(out.flatten(1))


def DenseNet121(num_classes=10): return DenseNet([6, 12, 24, 16], 32, num_classes=num_classes)
def DenseNet169(num_classes=10): return DenseNet([6, 12, 32, 32], 32, num_classes=num_classes)
def DenseNet201(num_classes=10): return DenseNet([6, 12, 48, 32], 32, num_classes=num_classes)
def DenseNet161(num_classes=10): return DenseNet([6, 12, 36, 24], 48, num_classes=num_classes)
def densenet_cifar(num_classes=10): return DenseNet([6, 12, 24, 16], 12, num_classes=num_classes)


# -------------------------------------------------------------------------------- ResNeXt ---

class ResNeXtBlock(nn.Module):
    expansion = 2

    def __init__(self, in_planes, cardinality=32, bottleneck_width=4, stride=1):
        super().__init__()
        width = cardinality * bottleneck_width
        self.conv1, self.bn1 = conv_bn(in_planes, width, 1)
        self.conv2, self.bn2 = conv_bn(width, width, 3, stride, 1, groups=cardinality)
        self.conv3, self.bn3 = conv_bn(width, self.expansion * width, 1)
        self.shortcut = _projection(in_planes, self.expansion * width, stride)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = F.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        return F.relu(out + self.shortcut(x))


class ResNeXt(nn.Module):
    def __init__(self, num_blocks, cardinality, bottleneck_width, num_classes=10):
        super().__init__()
        self.conv1, self.bn1 = conv_bn(3, 64, 1)
        planes, width = 64, bottleneck_width
        for stage, (count, stride) in enumerate(zip(num_blocks, (1, 2, 2)), start=1):
            blocks = []
            for s in [stride] + [1] * (count - 1):
                blocks.append(ResNeXtBlock(planes, cardinality, width, s))
                planes = ResNeXtBlock.expansion * cardinality * width
            setattr(self, "layer%d" % stage, nn.Sequential(*blocks))
            width *= 2
        self.linear = nn.Linear(cardinality * bottleneck_width * 8, num_classes)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.layer3(self.layer2(self.layer1(out)))
        return self.linear(F.avg_pool2d(out, 8).flatten(1))


def ResNeXt29_2x64d(num_classes=10): return ResNeXt([3, 3, 3], 2, 64, num_classes)
def ResNeXt29_4x64d(num_classes=10): return ResNeXt([3, 3, 3], 4, 64, num_classes)
def ResNeXt29_8x64d(num_classes=10): return ResNeXt([3, 3, 3], 8, 64, num_classes)
def ResNeXt29_32x4d(num_classes=10): return ResNeXt([3, 3, 3], 32, 4, num_classes)


# ------------------------------------------------------------------------------ MobileNet ---

class DepthwiseSeparable(nn.Module):
    def __init__(self, in_planes, out_planes, stride=1):
        super().__init__()
        self.conv1, self.bn1 = conv_bn(in_planes, in_planes, 3, stride, 1, groups=in_planes)
        self.conv2, self.bn2 = conv_bn(in_planes, out_planes, 1)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        return F.relu(self.bn2(self.conv2(out)))


class MobileNet(nn.Module):
    plan = [64, (128, 2), 128, (256, 2), 256, (512, 2), 512, 512, 512, 512, 512, (1024, 2), 1024]

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1, self.bn1 = conv_bn(3, 32, 3, 1, 1)
        blocks, planes = [], 32
        for item in self.plan:
            out_planes, stride = (item, 1) if isinstance(item, int) else item
            blocks.append(DepthwiseSeparable(planes, out_planes, stride))
            planes = out_planes
        self.layers = nn.Sequential(*blocks)
        self.linear = nn.Linear(1024, num_classes)

    def forward(self, x):
        out = self.layers(F.relu(self.bn1(self.conv1(x))))
        return self.linear(F.avg_pool2d(out, 2).flatten(1))


class InvertedResidual(nn.Module):
    def __init__(self, in_planes, out_planes, expansion, stride):
        super().__init__()
        self.stride = stride
        planes = expansion * in_planes
        self.conv1, self.bn1 = conv_bn(in_planes, planes, 1)
        self.conv2, self.bn2 = conv_bn(planes, planes, 3, stride, 1, groups=planes)
        self.conv3, self.bn3 = conv_bn(planes, out_planes, 1)
        self.shortcut = nn.Sequential()
        if stride == 1 and in_planes != out_planes:
            self.shortcut = nn.Sequential(*conv_bn(in_planes, out_planes, 1))

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = F.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        return out + self.shortcut(x) if self.stride == 1 else out


class MobileNetV2(nn.Module):
    plan = [(1, 16, 1, 1), (6, 24, 2, 1), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1)]

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1, self.bn1 = conv_bn(3, 32, 3, 1, 1)
        blocks, planes = [], 32
        for expansion, out_planes, count, stride in self.plan:
            for s in [stride] + [1] * (count - 1):
                blocks.append(InvertedResidual(planes, out_planes, expansion, s))
                planes = out_planes
        self.layers = nn.Sequential(*blocks)
        self.conv2, self.bn2 = conv_bn(320, 1280, 1)
        self.linear = nn.Linear(1280, num_classes)

    def forward(self, x):
        out = self.layers(F.relu(self.bn1(self.conv1(x))))
        out = F.relu(self.bn2(self.conv2(out)))
        return self.linear(F.avg_pool2d(out, 4).flatten(1))


# ------------------------------------------------------------------------------------ DPN ---

class DualPathBlock(nn.Module):
    def __init__(self, last_planes, in_planes, out_planes, dense_depth, stride, first_layer):
        super().__init__()
        self.out_planes, self.dense_depth = out_planes, dense_depth
        self.conv1, self.bn1 = conv_bn(last_planes, in_planes, 1)
        self.conv2, self.bn2 = conv_bn(in_planes, in_planes, 3, stride, 1, groups=32)
        self.conv3, self.bn3 = conv_bn(in_planes, out_planes + dense_depth, 1)
        self.shortcut = nn.Sequential()
        if first_layer:
            self.shortcut = nn.Sequential(*conv_bn(last_planes, out_planes + dense_depth, 1, stride=stride))

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = F.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        x = self.shortcut(x)
        d = self.out_planes
        out = torch.cat([x[:, :d] + out[:, :d], x[:, d:], out[:, d:]], 1)
        return F.relu(out)


class DPN(nn.Module):
    def __init__(self, cfg, num_classes=10):
        super().__init__()
        in_planes, out_planes = cfg["in_planes"], cfg["out_planes"]
        num_blocks, dense_depth = cfg["num_blocks"], cfg["dense_depth"]
        self.conv1, self.bn1 = conv_bn(3, 64, 3, 1, 1)
        last = 64
        for stage in range(4):
            blocks = []
            for i, s in enumerate([1 if stage == 0 else 2] + [1] * (num_blocks[stage] - 1)):
                blocks.append(DualPathBlock(last, in_planes[stage], out_planes[stage], dense_depth[stage], s, i == 0))
                last = out_planes[stage] + (i + 2) * dense_depth[stage]
            setattr(self, "layer%d" % (stage + 1), nn.Sequential(*blocks))
        self.linear = nn.Linear(out_planes[3] + (num_blocks[3] + 1) * dense_depth[3], num_classes)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.layer4(self.layer3(self.layer2(self.layer1(out))))
        return self.linear(F.avg_pool2d(out, 4).flatten(1))


def DPN26(num_classes=10):
    return DPN(dict(in_planes=(96, 192, 384, 768), out_planes=(256, 512, 1024, 2048), num_blocks=(2, 2, 2, 2),
                    dense_depth=(16, 32, 24, 128)), num_classes)


def DPN92(num_classes=10):
    return DPN(dict(in_planes=(96, 192, 384, 768), out_planes=(256, 512, 1024, 2048), num_blocks=(3, 4, 20, 3),
                    dense_depth=(16, 32, 24, 128)), num_classes)


# ----------------------------------------------------------------------------- ShuffleNet ---

def channel_shuffle(x, groups):
    n, c, h, w = x.shape
    return x.view(n, groups, c // groups, h, w).transpose(1, 2).reshape(n, c, h, w)


class ShuffleUnit(nn.Module):
    def __init__(self, in_planes, out_planes, stride, groups):
        super().__init__()
        self.stride, self.groups = stride, groups
        mid = out_planes // 4
        g = 1 if in_planes == 24 else groups
        self.conv1, self.bn1 = conv_bn(in_planes, mid, 1, groups=g)
        self.shuffle_groups = g
        self.conv2, self.bn2 = conv_bn(mid, mid, 3, stride, 1, groups=mid)
        self.conv3, self.bn3 = conv_bn(mid, out_planes, 1, groups=groups)
        self.shortcut = nn.Sequential(nn.AvgPool2d(3, stride=2, padding=1)) if stride == 2 else nn.Sequential()

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = channel_shuffle(out, self.shuffle_groups)
        out = F.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        res = self.shortcut(x)
        return F.relu(torch.cat([out, res], 1)) if self.stride == 2 else F.relu(out + res)


class ShuffleNet(nn.Module):
    def __init__(self, cfg, num_classes=10):
        super().__init__()
        out_planes, num_blocks, groups = cfg["out_planes"], cfg["num_blocks"], cfg["groups"]
        self.conv1, self.bn1 = conv_bn(3, 24, 1)
        planes = 24
        for stage in range(3):
            blocks = []
            for i in range(num_blocks[stage]):
                stride = 2 if i == 0 else 1
                cat_planes = planes if i == 0 else 0
                blocks.append(ShuffleUnit(planes, out_planes[stage] - cat_planes, stride, groups))
                planes = out_planes[stage]
            setattr(self, "layer%d" % (stage + 1), nn.Sequential(*blocks))
        self.linear = nn.Linear(out_planes[2], num_classes)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.layer3(self.layer2(self.layer1(out)))
        return self.linear(F.avg_pool2d(out, 4).flatten(1))


def ShuffleNetG2(num_classes=10):
    return ShuffleNet(dict(out_planes=[200, 400, 800], num_blocks=[4, 8, 4], groups=2), num_classes)


def ShuffleNetG3(num_classes=10):
    return ShuffleNet(dict(out_planes=[240, 480, 960], num_blocks=[4, 8, 4], groups=3), num_classes)


class ShuffleV2Basic(nn.Module):
    def __init__(self, in_channels, split_ratio=0.5):
        super().__init__()
        self.split = int(in_channels * split_ratio)
        c = self.split
        self.conv1, self.bn1 = conv_bn(c, c, 1)
        self.conv2, self.bn2 = conv_bn(c, c, 3, 1, 1, groups=c)
        self.conv3, self.bn3 = conv_bn(c, c, 1)

    def forward(self, x):
        keep, work = x[:, :self.split], x[:, self.split:]
        out = F.relu(self.bn1(self.conv1(work)))
        out = self.bn2(self.conv2(out))
        out = F.relu(self.bn3(self.conv3(out)))
        return channel_shuffle(torch.cat([keep, out], 1), 2)


class ShuffleV2Down(nn.Module):
    def __init__(self, in_channels, out_channels):
        super().__init__()
        mid = out_channels // 2
        self.conv1, self.bn1 = conv_bn(in_channels, in_channels, 3, 2, 1, groups=in_channels)
        self.conv2, self.bn2 = conv_bn(in_channels, mid, 1)
        self.conv3, self.bn3 = conv_bn(in_channels, mid, 1)
        self.conv4, self.bn4 = conv_bn(mid, mid, 3, 2, 1, groups=mid)
        self.conv5, self.bn5 = conv_bn(mid, mid, 1)

    def forward(self, x):
        left = F.relu(self.bn2(self.conv2(self.bn1(self.conv1(x)))))
        right = F.relu(self.bn3(self.conv3(x)))
        right = self.bn4(self.conv4(right))
        right = F.relu(self.bn5(self.conv5(right)))
        return channel_shuffle(torch.cat([left, right], 1), 2)


_SHUFFLE_V2 = {0.5: ((48, 96, 192, 1024), (3, 7, 3)), 1: ((116, 232, 464, 1024), (3, 7, 3)),
               1.5: ((176, 352, 704, 1024), (3, 7, 3)), 2: ((224, 488, 976, 2048), (3, 7, 3))}


class ShuffleNetV2(nn.Module):
    def __init__(self, net_size=0.5, num_classes=10):
        super().__init__()
        out_channels, num_blocks = _SHUFFLE_V2[net_size]
        self.conv1, self.bn1 = conv_bn(3, 24, 3, 1, 1)
        channels = 24
        for stage in range(3):
            blocks = [ShuffleV2Down(channels, out_channels[stage])]
            blocks += [ShuffleV2Basic(out_channels[stage]) for _ in range(num_blocks[stage])]
            channels = out_channels[stage]
            setattr(self, "layer%d" % (stage + 1), nn.Sequential(*blocks))
        self.conv2, self.bn2 = conv_bn(out_channels[2], out_channels[3], 1)
        self.linear = nn.Linear(out_channels[3], num_classes)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.layer3(self.layer2(self.layer1(out)))
        out = F.relu(self.bn2(self.conv2(out)))
        return self.linear(F.avg_pool2d(out, 4).flatten(1))


# ---------------------------------------------------------------------------------- SENet ---

class SEPreActBlock(nn.Module):
    def __init__(self, in_planes, planes, stride=1):
        super().__init__()
        self.bn1 = nn.BatchNorm2d(in_planes)
        self.conv1 = nn.Conv2d(in_planes, planes, 3, stride, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        if stride != 1 or in_planes != planes:
            self.shortcut = nn.Sequential(nn.Conv2d(in_planes, planes, 1, stride, bias=False))
        self.fc1 = nn.Conv2d(planes, planes // 16, 1)
        self.fc2 = nn.Conv2d(planes // 16, planes, 1)

    def forward(self, x):
        out = F.relu(self.bn1(x))
        skip = self.shortcut(out) if hasattr(self, "shortcut") else x
        out = self.conv1(out)
        out = self.conv2(F.relu(self.bn2(out)))
        gate = F.avg_pool2d(out, out.size(2))
        gate = torch.sigmoid(self.fc2(F.relu(self.fc1(gate))))
        return out * gate + skip


class SENet(nn.Module):
    def __init__(self, num_blocks, num_classes=10):
        super().__init__()
        self.conv1, self.bn1 = conv_bn(3, 64, 3, 1, 1)
        planes = 64
        for stage, (width, stride) in enumerate(zip((64, 128, 256, 512), (1, 2, 2, 2)), start=1):
            blocks = []
            for s in [stride] + [1] * (num_blocks[stage - 1] - 1):
                blocks.append(SEPreActBlock(planes, width, s))
                planes = width
            setattr(self, "layer%d" % stage, nn.Sequential(*blocks))
        self.linear = nn.Linear(512, num_classes)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.layer4(self.layer3(self.layer2(self.layer1(out))))
        return self.linear(F.avg_pool2d(out, 4).flatten(1))


def SENet18(num_classes=10): return SENet([2, 2, 2, 2], num_classes)


# --------------------------------------------------------------------------- EfficientNet ---

def swish(x):
    return x * x.sigmoid()


def drop_connect(x, ratio):
    keep = 1.0 - ratio
    mask = torch.empty([x.shape[0], 1, 1, 1], dtype=x.dtype, device=x.device).bernoulli_(keep)
    return x.div(keep) * mask


class SwishSqueezeExcite(nn.Module):
    def __init__(self, in_channels, se_channels):
        super().__init__()
        self.se1 = nn.Conv2d(in_channels, se_channels, 1, bias=True)
        self.se2 = nn.Conv2d(se_channels, in_channels, 1, bias=True)

    def forward(self, x):
        gate = F.adaptive_avg_pool2d(x, (1, 1))
        return x * self.se2(swish(self.se1(gate))).sigmoid()


class MBConv(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size, stride, expand_ratio=1, se_ratio=0., drop_rate=0.):
        super().__init__()
        self.stride, self.drop_rate, self.expand_ratio = stride, drop_rate, expand_ratio
        channels = expand_ratio * in_channels
        self.conv1, self.bn1 = conv_bn(in_channels, channels, 1)
        self.conv2, self.bn2 = conv_bn(channels, channels, kernel_size, stride, 1 if kernel_size == 3 else 2, groups=channels)
        self.se = SwishSqueezeExcite(channels, int(in_channels * se_ratio))
        self.conv3, self.bn3 = conv_bn(channels, out_channels, 1)
        self.has_skip = stride == 1 and in_channels == out_channels

    def forward(self, x):
        out = x if self.expand_ratio == 1 else swish(self.bn1(self.conv1(x)))
        out = swish(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(self.se(out)))
        if self.has_skip:
            if self.training and self.drop_rate > 0:
                out = drop_connect(out, self.drop_rate)
            out = out + x
        return out


class EfficientNet(nn.Module):
    def __init__(self, cfg, num_classes=10):
        super().__init__()
        self.cfg = cfg
        self.conv1, self.bn1 = conv_bn(3, 32, 3, 1, 1)
        blocks, channels = [], 32
        total = sum(cfg["num_blocks"])
        index = 0
        for expansion, out_channels, count, kernel, stride in zip(cfg["expansion"], cfg["out_channels"], cfg["num_blocks"],
                                                                  cfg["kernel_size"], cfg["stride"]):
            for s in [stride] + [1] * (count - 1):
                rate = cfg["drop_connect_rate"] * index / total
                blocks.append(MBConv(channels, out_channels, kernel, s, expansion, se_ratio=0.25, drop_rate=rate))
                channels = out_channels
                index += 1
        self.layers = nn.Sequential(*blocks)
        self.linear = nn.Linear(cfg["out_channels"][-1], num_classes)

    def forward(self, x):
        out = self.layers(swish(self.bn1(self.conv1(x))))
        out = F.adaptive_avg_pool2d(out, 1).flatten(1)
        if self.training and self.cfg["dropout_rate"] > 0:
            out = F.dropout(out, p=self.cfg["dropout_rate"])
        return self.linear(out)


def EfficientNetB0(num_classes=10):
    return EfficientNet(dict(num_blocks=[1, 2, 2, 3, 3, 4, 1], expansion=[1, 6, 6, 6, 6, 6, 6],
                             out_channels=[16, 24, 40, 80, 112, 192, 320], kernel_size=[3, 3, 5, 3, 5, 5, 3],
                             stride=[1, 2, 2, 2, 1, 2, 1], dropout_rate=0.2, drop_connect_rate=0.2), num_classes)


# --------------------------------------------------------------------------------- RegNet ---

class SqueezeExcite(nn.Module):
    def __init__(self, in_planes, se_planes):
        super().__init__()
        self.se1 = nn.Conv2d(in_planes, se_planes, 1, bias=True)
        self.se2 = nn.Conv2d(se_planes, in_planes, 1, bias=True)

    def forward(self, x):
        gate = F.adaptive_avg_pool2d(x, (1, 1))
        return x * self.se2(F.relu(self.se1(gate))).sigmoid()


class RegNetBlock(nn.Module):
    def __init__(self, w_in, w_out, stride, group_width, bottleneck_ratio, se_ratio):
        super().__init__()
        w_b = int(round(w_out * bottleneck_ratio))
        self.conv1, self.bn1 = conv_bn(w_in, w_b, 1)
        self.conv2, self.bn2 = conv_bn(w_b, w_b, 3, stride, 1, groups=w_b // group_width)
        self.with_se = se_ratio > 0
        if self.with_se:
            self.se = SqueezeExcite(w_b, int(round(w_in * se_ratio)))
        self.conv3, self.bn3 = conv_bn(w_b, w_out, 1)
        self.shortcut = _projection(w_in, w_out, stride)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = F.relu(self.bn2(self.conv2(out)))
        if self.with_se:
            out = self.se(out)
        out = self.bn3(self.conv3(out))
        return F.relu(out + self.shortcut(x))


class RegNet(nn.Module):
    def __init__(self, cfg, num_classes=10):
        super().__init__()
        self.conv1, self.bn1 = conv_bn(3, 64, 3, 1, 1)
        planes = 64
        for stage in range(4):
            blocks = []
            for i in range(cfg["depths"][stage]):
                blocks.append(RegNetBlock(planes, cfg["widths"][stage], cfg["strides"][stage] if i == 0 else 1,
                                          cfg["group_width"], cfg["bottleneck_ratio"], cfg["se_ratio"]))
                planes = cfg["widths"][stage]
            setattr(self, "layer%d" % (stage + 1), nn.Sequential(*blocks))
        self.linear = nn.Linear(cfg["widths"][-1], num_classes)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.layer4(self.layer3(self.layer2(self.layer1(out))))
        return self.linear(F.adaptive_avg_pool2d(out, (1, 1)).flatten(1))


def RegNetX_200MF(num_classes=10):
    return RegNet(dict(depths=[1, 1, 4, 7], widths=[24, 56, 152, 368], strides=[1, 1, 2, 2], group_width=8,
                       bottleneck_ratio=1, se_ratio=0), num_classes)


def RegNetX_400MF(num_classes=10):
    return RegNet(dict(depths=[1, 2, 7, 12], widths=[32, 64, 160, 384], strides=[1, 1, 2, 2], group_width=16,
                       bottleneck_ratio=1, se_ratio=0), num_classes)


def RegNetY_400MF(num_classes=10):
    # same stage plan as X_400MF plus squeeze-excite (the reference's Y_400MF config, regnet.py:134-143)
    return RegNet(dict(depths=[1, 2, 7, 12], widths=[32, 64, 160, 384], strides=[1, 1, 2, 2], group_width=16,
                       bottleneck_ratio=1, se_ratio=0.25), num_classes)


# -------------------------------------------------------------------------------- PNASNet ---

class SepConv(nn.Module):
    def __init__(self, in_planes, out_planes, kernel_size, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(in_planes, out_planes, kernel_size, stride, padding=(kernel_size - 1) // 2, bias=False,
                               groups=in_planes)
        self.bn1 = nn.BatchNorm2d(out_planes)

    def forward(self, x):
        return self.bn1(self.conv1(x))


class CellA(nn.Module):
    def __init__(self, in_planes, out_planes, stride=1):
        super().__init__()
        self.stride = stride
        self.sep_conv1 = SepConv(in_planes, out_planes, 7, stride)
        if stride == 2:
            self.conv1, self.bn1 = conv_bn(in_planes, out_planes, 1)

    def forward(self, x):
        y1 = self.sep_conv1(x)
        y2 = F.max_pool2d(x, 3, stride=self.stride, padding=1)
        if self.stride == 2:
            y2 = self.bn1(self.conv1(y2))
        return F.relu(y1 + y2)


class CellB(nn.Module):
    def __init__(self, in_planes, out_planes, stride=1):
        super().__init__()
        self.stride = stride
        self.sep_conv1 = SepConv(in_planes, out_planes, 7, stride)
        self.sep_conv2 = SepConv(in_planes, out_planes, 3, stride)
        self.sep_conv3 = SepConv(in_planes, out_planes, 5, stride)
        if stride == 2:
            self.conv1, self.bn1 = conv_bn(in_planes, out_planes, 1)
        self.conv2, self.bn2 = conv_bn(2 * out_planes, out_planes, 1)

    def forward(self, x):
        y1, y2 = self.sep_conv1(x), self.sep_conv2(x)
        y3 = F.max_pool2d(x, 3, stride=self.stride, padding=1)
        if self.stride == 2:
            y3 = self.bn1(self.conv1(y3))
        y4 = self.sep_conv3(x)
        merged = torch.cat([F.relu(y1 + y2), F.relu(y3 + y4)], 1)
        return F.relu(self.bn2(self.conv2(merged)))


class PNASNet(nn.Module):
    def __init__(self, cell_type, num_cells, num_planes, num_classes=10):
        super().__init__()
        self.conv1, self.bn1 = conv_bn(3, num_planes, 3, 1, 1)
        planes = num_planes

        def cells(count, width):
            return nn.Sequential(*[cell_type(planes if i == 0 else width, width, 1) for i in range(count)])

        self.layer1 = cells(6 if num_cells is None else num_cells, num_planes)
        self.layer2 = cell_type(num_planes, num_planes * 2, 2)
        planes = num_planes * 2
        self.layer3 = cells(num_cells, num_planes * 2)
        self.layer4 = cell_type(num_planes * 2, num_planes * 4, 2)
        planes = num_planes * 4
        self.layer5 = cells(num_cells, num_planes * 4)
        self.linear = nn.Linear(num_planes * 4, num_classes)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.layer5(self.layer4(self.layer3(self.layer2(self.layer1(out)))))
        return self.linear(F.avg_pool2d(out, 8).flatten(1))


def PNASNetA(num_classes=10): return PNASNet(CellA, 6, 44, num_classes)
def PNASNetB(num_classes=10): return PNASNet(CellB, 6, 32, num_classes)
