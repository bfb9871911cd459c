"""CIFAR ResNets as torch modules, generated from two small tables.

Architecture (reference: pytorch_code/model_ops/resnet.py:14-113, the kuangliu CIFAR family): 3x3 stem to 64 channels,
four stages of widths 64 / 128 / 256 / 512 whose first unit has stride 1 / 2 / 2 / 2, 4x4 average pool, linear
classifier. Module names (``conv1 bn1 layer<s>.<i>.conv<j> bn<j> shortcut.0 shortcut.1 linear``) and therefore
``state_dict`` keys and ``parameters()`` order are those of the reference, so checkpoints interchange; the code that
produces them is table-driven instead of one hand-written class per block type.
"""
from functools import partial

import torch
import torch.nn as nn

STAGE_WIDTHS = (64, 128, 256, 512)
STAGE_STRIDES = (1, 2, 2, 2)
# residual path of a unit: (kernel, takes the unit's stride?, output width as a multiple of `planes`) per conv
RESIDUAL_PATHS = {
    "basic": ((3, True, 1), (3, False, 1)),
    "bottleneck": ((1, False, 1), (3, True, 1), (1, False, 4)),
}
# depth -> (unit kind, units per stage)
DEPTHS = {
    18: ("basic", (2, 2, 2, 2)),
    34: ("basic", (3, 4, 6, 3)),
    50: ("bottleneck", (3, 4, 6, 3)),
    101: ("bottleneck", (3, 4, 23, 3)),
    152: ("bottleneck", (3, 8, 36, 3)),
}


class ResidualUnit(nn.Module):
    """conv-bn(-relu) chain described by ``RESIDUAL_PATHS[kind]`` plus identity or 1x1-projection shortcut; ReLU after
    the sum. Sub-modules are registered as conv1, bn1, conv2, bn2, ... and ``shortcut`` (a Sequential, empty = identity)."""
    kind = "basic"
    expansion = 1

    def __init__(self, in_planes, planes, stride=1):
        super().__init__()
        width = in_planes
        path = RESIDUAL_PATHS[self.kind]
        for j, (kernel, strided, mult) in enumerate(path, start=1):
            out = planes * mult
            self.add_module(f"conv{j}", nn.Conv2d(width, out, kernel, stride if strided else 1, kernel // 2, bias=False))
            self.add_module(f"bn{j}", nn.BatchNorm2d(out))
            width = out
        self.depth = len(path)
        projection = stride != 1 or in_planes != width
        self.shortcut = nn.Sequential(*([nn.Conv2d(in_planes, width, 1, stride, bias=False), nn.BatchNorm2d(width)] if projection else []))

    def forward(self, x):
        y = x
        for j in range(1, self.depth + 1):
            y = getattr(self, f"bn{j}")(getattr(self, f"conv{j}")(y))
            if j < self.depth:
                y = torch.relu(y)
        return torch.relu(y + self.shortcut(x))


class BasicBlock(ResidualUnit):
    kind, expansion = "basic", 1


class Bottleneck(ResidualUnit):
    kind, expansion = "bottleneck", 4


def build_stages(owner: nn.Module, block, units_per_stage) -> int:
    """Registers ``layer1`` .. ``layer4`` on ``owner``; returns the channel count leaving the last stage."""
    width = STAGE_WIDTHS[0]
    for s, (planes, stride, n) in enumerate(zip(STAGE_WIDTHS, STAGE_STRIDES, units_per_stage), start=1):
        units = []
        for i in range(n):
            units.append(block(width, planes, stride if i == 0 else 1))
            width = planes * block.expansion
        setattr(owner, f"layer{s}", nn.Sequential(*units))
    return width


class ResNet(nn.Module):
    def __init__(self, block, num_blocks, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, STAGE_WIDTHS[0], 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(STAGE_WIDTHS[0])
        width = build_stages(self, block, num_blocks)
        self.linear = nn.Linear(width, num_classes)          # registered last: parameters() ends with the classifier

    def stages(self):
        return [getattr(self, f"layer{s}") for s in range(1, len(STAGE_WIDTHS) + 1)]

    def forward(self, x):
        y = torch.relu(self.bn1(self.conv1(x)))
        for stage in self.stages():
            y = stage(y)
        return self.linear(torch.flatten(nn.functional.avg_pool2d(y, 4), 1))

    def name(self):
        return "resnet"


def resnet(depth: int, num_classes: int = 10) -> ResNet:
    kind, units = DEPTHS[depth]
    return ResNet({"basic": BasicBlock, "bottleneck": Bottleneck}[kind], units, num_classes)


ResNet18 = partial(resnet, 18)
ResNet34 = partial(resnet, 34)
ResNet50 = partial(resnet, 50)
ResNet101 = partial(resnet, 101)
ResNet152 = partial(resnet, 152)
