"""Model zoo (torch modules). ``build_model(name)`` resolves the reference's ``--network`` names."""
from .lenet import LeNet, LeNetSplit
from .mlp import MLP, MLPSplit, DIMS_CPP, DIMS_DEEP
from .resnet import BasicBlock, Bottleneck, ResNet, ResNet18, ResNet34, ResNet50, ResNet101, ResNet152
from .resnet_split import (ResNetSplit, ResNetSplit18, ResNetSplit34, ResNetSplit50, ResNetSplit101,
                           ResNetSplit152)
from .vgg import VGG, VGG11, VGG13, VGG16, VGG19, VGGSplit

_PLAIN = {
    "LeNet": LeNet, "ResNet": ResNet18, "Resnet": ResNet18, "ResNet18": ResNet18, "ResNet34": ResNet34,
    "ResNet50": ResNet50, "ResNet101": ResNet101, "ResNet152": ResNet152,
    "VGG11": VGG11, "VGG13": VGG13, "VGG16": VGG16, "VGG19": VGG19, "FC": MLP, "MLP": MLP,
    "FCDeep": lambda: MLP(DIMS_DEEP),          # the 28-layer net of pure_py_code/distributed_nn.py:41-96 and the TF study
}
_SPLIT = {
    "LeNet": LeNetSplit, "ResNet": ResNetSplit18, "Resnet": ResNetSplit18, "ResNet18": ResNetSplit18,
    "ResNet34": ResNetSplit34, "ResNet50": ResNetSplit50, "ResNet101": ResNetSplit101, "ResNet152": ResNetSplit152,
    "VGG11": lambda: VGGSplit("VGG11"), "VGG13": lambda: VGGSplit("VGG13"), "VGG16": lambda: VGGSplit("VGG16"),
    "VGG19": lambda: VGGSplit("VGG19"), "FC": MLPSplit, "MLP": MLPSplit, "FCDeep": lambda: MLPSplit(DIMS_DEEP),
}


def build_model(name: str, split: bool = False):
    table = _SPLIT if split else _PLAIN
    if name not in table:
        raise ValueError(f"unknown --network {name!r}; choose from {sorted(table)}")
    return table[name]()


def input_shape(network: str, dataset: str = None):
    if network in ("LeNet", "FC", "MLP", "FCDeep") or (dataset or "").upper() == "MNIST":
        return (1, 28, 28)
    return (3, 32, 32)
