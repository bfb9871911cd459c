"""Fully connected sigmoid MLPs of the reference's numpy and C++ stacks, as torch modules.

* ``FC_NN``-style 8-layer net 784-500-500-800-800-200-100-100-10 (MPI_code/src/distributed_nn.cpp:35-47)
* 28-layer net 784-500-500-800-(800)x22-200-100-10 (pure_py_code/distributed_nn.py:41-96,
  distributed_TF/src/mnist.py:150-412)
Sigmoid between layers, softmax cross-entropy on the last (pure_py_code/nn/nn_layer.py:113-145).
"""
import torch
import torch.nn as nn

from .split import SplitNet, Step

DIMS_CPP = [784, 500, 500, 800, 800, 200, 100, 100, 10]
DIMS_DEEP = [784, 500, 500, 800] + [800] * 22 + [200, 100, 10]


class MLP(nn.Module):
    def __init__(self, dims=DIMS_CPP):
        super().__init__()
        self.dims = list(dims)
        self.layers = nn.ModuleList(nn.Linear(dims[i], dims[i + 1]) for i in range(len(dims) - 1))

    def forward(self, x):
        x = x.flatten(1)
        for i, fc in enumerate(self.layers):
            x = fc(x)
            if i < len(self.layers) - 1:
                x = torch.sigmoid(x)
        return x

    def name(self):
        return "mlp"


class MLPSplit(SplitNet):
    def __init__(self, dims=DIMS_CPP):
        super().__init__()
        self.dims = list(dims)
        self.layers = nn.ModuleList(nn.Linear(dims[i], dims[i + 1]) for i in range(len(dims) - 1))
        steps, cur = [], "x"
        n = len(self.layers)
        for i, fc in enumerate(self.layers):
            if i == 0:
                steps.append(Step(f"fc{i}", lambda t, fc=fc: fc(t.flatten(1)), [cur], f"z{i}", [fc]))
            else:
                steps.append(Step(f"fc{i}", fc, [cur], f"z{i}", [fc]))
            cur = f"z{i}"
            if i < n - 1:
                steps.append(Step(f"sig{i}", torch.sigmoid, [cur], f"a{i}"))
                cur = f"a{i}"
        self.set_steps(steps)

    def name(self):
        return "mlp"
