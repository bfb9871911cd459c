"""``backward(variables, grad_variables)`` entry point (reference pytorch_code/distributed_functions/distributed_backward.py).

The reference vendors torch-0.2's ``torch.autograd.backward`` as a hook point for distributing backward and never
modifies or calls it. The hook point this framework actually uses is the layer-split tape (``models/split.py``) on
CPU and the engine's explicit layer backward on GPU; this module keeps the import path alive and forwards to
``torch.autograd.backward`` so that code written against the reference's name still runs.
"""
import torch


def backward(variables, grad_variables=None, retain_graph=None, create_graph=False):
    return torch.autograd.backward(variables, grad_variables, retain_graph=retain_graph, create_graph=create_graph)
