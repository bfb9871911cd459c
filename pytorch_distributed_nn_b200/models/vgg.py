"""CIFAR VGG-11/13/16/19 with BatchNorm (north-star addition: the reference snapshot ships no VGG).
3x3 convs (with bias) + BN + ReLU, 2x2 max pools, one linear classifier on the 512-d 1x1 feature map."""
import torch.nn as nn

from .split import SplitNet, Step

CFG = {
    "VGG11": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG13": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "VGG16": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "VGG19": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"],
}


def _features(cfg):
    layers, cin = [], 3
    for v in cfg:
        if v == "M":
            layers.append(nn.MaxPool2d(2, 2))
        else:
            layers += [nn.Conv2d(cin, v, 3, padding=1), nn.BatchNorm2d(v), nn.ReLU(inplace=False)]
            cin = v
    return nn.Sequential(*layers)


class VGG(nn.Module):
    def __init__(self, vgg_name="VGG11", num_classes=10):
        super().__init__()
        self.features = _features(CFG[vgg_name])
        self.classifier = nn.Linear(512, num_classes)

    def forward(self, x):
        return self.classifier(self.features(x).flatten(1))

    def name(self):
        return "vgg"


class VGGSplit(SplitNet):
    """Layer-split VGG: one tape step per module of ``features`` plus the classifier."""

    def __init__(self, vgg_name="VGG11", num_classes=10):
        super().__init__()
        self.features = _features(CFG[vgg_name])
        self.classifier = nn.Linear(512, num_classes)
        steps, cur = [], "x"
        for i, m in enumerate(self.features):
            owns = [m] if any(True for _ in m.parameters()) else []
            steps.append(Step(f"features.{i}", m, [cur], f"f{i}", owns))
            cur = f"f{i}"
        steps.append(Step("classifier", lambda t: self.classifier(t.flatten(1)), [cur], "out", [self.classifier]))
        self.set_steps(steps)

    def name(self):
        return "vgg"


def VGG11():
    return VGG("VGG11")


def VGG13():
    return VGG("VGG13")


def VGG16():
    return VGG("VGG16")


def VGG19():
    return VGG("VGG19")
