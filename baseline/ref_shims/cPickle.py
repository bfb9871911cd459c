"""Python-2 module name the reference's CIFAR reader imports (cifar10/cifar10.py:11); same API as pickle."""
from pickle import *          # noqa: F401,F403
from pickle import load, loads, dump, dumps  # noqa: F401
