"""Layer-split CIFAR ResNets: same parameters as :mod:`.resnet`, backward streams gradients per op.

Reference: pytorch_code/model_ops/resnet_split.py (BasicBlockSplit :17-87, ResNetSplit :119-440). The reference
leaves ``Bottleneck`` unsplit (:90-116) so ResNetSplit50+ cannot stream; here both block types are split.
The residual branch point is a single tape leaf consumed by ``conv1`` and by the shortcut, so its gradient is
the sum of both paths (reference: in-place ``out += self.shortcut(x)``, :81).
"""
import torch.nn as nn
import torch.nn.functional as F

from . import resnet as R
from .split import SplitNet, Step


def _block_steps(prefix, blk, in_key, out_key):
    """Expand one residual block into tape steps. Parameter order inside a block: conv1, bn1, conv2, bn2,
    [conv3, bn3,] shortcut conv, shortcut bn -- the steps are listed in that order so that the reversed
    walk emits gradients in reverse parameters() order."""
    k = lambda s: f"{prefix}.{s}"  # noqa: E731
    steps = [
        Step(k("conv1"), blk.conv1, [in_key], k("c1"), [blk.conv1]),
        Step(k("bn1"), blk.bn1, [k("c1")], k("b1"), [blk.bn1]),
        Step(k("relu1"), F.relu, [k("b1")], k("r1")),
        Step(k("conv2"), blk.conv2, [k("r1")], k("c2"), [blk.conv2]),
    ]
    if isinstance(blk, R.Bottleneck):
        steps += [
            Step(k("bn2"), blk.bn2, [k("c2")], k("b2"), [blk.bn2]),
            Step(k("relu2"), F.relu, [k("b2")], k("r2")),
            Step(k("conv3"), blk.conv3, [k("r2")], k("c3"), [blk.conv3]),
        ]
        last_bn, last_in = blk.bn3, k("c3")
    else:
        last_bn, last_in = blk.bn2, k("c2")
    if len(blk.shortcut) > 0:
        steps += [
            Step(k("bn_last"), last_bn, [last_in], k("bl"), [last_bn]),
            Step(k("sc_conv"), blk.shortcut[0], [in_key], k("sc"), [blk.shortcut[0]]),
            Step(k("sc_bn"), blk.shortcut[1], [k("sc")], k("sb"), [blk.shortcut[1]]),
            Step(k("add_relu"), lambda a, b: F.relu(a + b), [k("bl"), k("sb")], out_key),
        ]
    else:
        steps += [
            Step(k("bn_last"), last_bn, [last_in], k("bl"), [last_bn]),
            Step(k("add_relu"), lambda a, b: F.relu(a + b), [k("bl"), in_key], out_key),
        ]
    return steps


BasicBlockSplit = R.BasicBlock   # blocks are plain modules; the split happens in the tape
Bottleneck = R.Bottleneck


class ResNetSplit(SplitNet):
    def __init__(self, block, num_blocks, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, R.STAGE_WIDTHS[0], 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(R.STAGE_WIDTHS[0])
        width = R.build_stages(self, block, num_blocks)
        self.linear = nn.Linear(width, num_classes)
        steps = [
            Step("conv1", self.conv1, ["x"], "stem.c", [self.conv1]),
            Step("bn1", self.bn1, ["stem.c"], "stem.b", [self.bn1]),
            Step("relu", F.relu, ["stem.b"], "stem.r"),
        ]
        cur = "stem.r"
        for si, stage in enumerate([self.layer1, self.layer2, self.layer3, self.layer4], start=1):
            for bi, blk in enumerate(stage):
                out_key = f"layer{si}.{bi}.out"
                steps += _block_steps(f"layer{si}.{bi}", blk, cur, out_key)
                cur = out_key
        steps += [
            Step("avgpool", lambda t: F.avg_pool2d(t, 4).flatten(1), [cur], "pooled"),
            Step("linear", self.linear, ["pooled"], "out", [self.linear]),
        ]
        self.set_steps(steps)

    def name(self):
        return "resnet"


def ResNetSplit18():
    return ResNetSplit(R.BasicBlock, [2, 2, 2, 2])


def ResNetSplit34():
    return ResNetSplit(R.BasicBlock, [3, 4, 6, 3])


def ResNetSplit50():
    return ResNetSplit(R.Bottleneck, [3, 4, 6, 3])


def ResNetSplit101():
    return ResNetSplit(R.Bottleneck, [3, 4, 23, 3])


def ResNetSplit152():
    return ResNetSplit(R.Bottleneck, [3, 8, 36, 3])
