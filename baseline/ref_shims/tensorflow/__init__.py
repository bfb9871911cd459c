"""Import stub for the few TensorFlow names the reference's *data readers* touch at import time
(cifar10/cifar10.py:13-26, mnist/mnist.py:14-17): the flags object, the dtype handles and the
``Datasets`` tuple. No TensorFlow computation exists on the reference's PyTorch path."""
import collections as _c
import types as _t

from .python.framework import dtypes as _dtypes

float32, int64, uint8 = _dtypes.float32, _dtypes.int64, _dtypes.uint8


class _Flags(object):
    class _Values(object):
        pass

    def __init__(self):
        self.FLAGS = self._Values()

    def _define(self, name, default, doc=""):
        setattr(self.FLAGS, name, default)

    DEFINE_integer = DEFINE_float = DEFINE_string = DEFINE_boolean = DEFINE_bool = _define


app = _t.SimpleNamespace(flags=_Flags(), run=lambda main=None, argv=None: main(argv) if main else None)
flags = app.flags


def placeholder(*a, **k):
    raise NotImplementedError("tensorflow stub: no graph execution on the reference's PyTorch path")
