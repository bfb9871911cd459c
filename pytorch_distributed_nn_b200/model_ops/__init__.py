"""Reference-compatible import path: ``model_ops.lenet``, ``model_ops.resnet``, ``model_ops.resnet_split``, ``model_ops.vgg``."""
from ..models import lenet, resnet, resnet_split, vgg, mlp  # noqa: F401
from ..models import *  # noqa: F401,F403
