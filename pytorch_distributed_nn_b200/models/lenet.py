"""LeNet for 1x28x28 inputs (torch modules: CPU / gloo path and the numerical oracle of the B200 engine).

Architecture as in the reference (pytorch_code/model_ops/lenet.py:12-33): conv(1->20,5) . maxpool2 . relu .
conv(20->50,5) . maxpool2 . relu . fc(800->500) . fc(500->10) -- pooling *before* ReLU and no activation
between the two fully connected layers. ``LeNetSplit`` is the layer-split variant whose ``backward`` emits
each parameter gradient as soon as it exists (reference lenet.py:36-225); it is built on the generic
:class:`~pytorch_distributed_nn_b200.models.split.SplitNet`.
"""
import torch.nn as nn
import torch.nn.functional as F

from .split import SplitNet, Step


class LeNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 20, 5, 1)
        self.conv2 = nn.Conv2d(20, 50, 5, 1)
        self.fc1 = nn.Linear(4 * 4 * 50, 500)
        self.fc2 = nn.Linear(500, 10)

    def forward(self, x):
        x = F.relu(F.max_pool2d(self.conv1(x), 2, 2))
        x = F.relu(F.max_pool2d(self.conv2(x), 2, 2))
        x = x.flatten(1)
        return self.fc2(self.fc1(x))

    def name(self):
        return "lenet"


class LeNetSplit(SplitNet):
    """Same weights and math as :class:`LeNet`; forward records one tape entry per op (8 ops, as the
    reference's ``self.output`` list) so that backward can stream gradients in reverse parameter order."""

    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 20, 5, 1)
        self.conv2 = nn.Conv2d(20, 50, 5, 1)
        self.fc1 = nn.Linear(4 * 4 * 50, 500)
        self.fc2 = nn.Linear(500, 10)
        pool = lambda t: F.max_pool2d(t, 2, 2)  # noqa: E731
        self.set_steps([
            Step("conv1", self.conv1, ["x"], "c1", [self.conv1]),
            Step("pool1", pool, ["c1"], "p1"),
            Step("relu1", F.relu, ["p1"], "r1"),
            Step("conv2", self.conv2, ["r1"], "c2", [self.conv2]),
            Step("pool2", pool, ["c2"], "p2"),
            Step("relu2", F.relu, ["p2"], "r2"),
            Step("fc1", lambda t: self.fc1(t.flatten(1)), ["r2"], "f1", [self.fc1]),
            Step("fc2", self.fc2, ["f1"], "out", [self.fc2]),
        ])

    def name(self):
        return "lenet"
