"""dtype handles compared by identity in DataSet.__init__ (cifar10.py:55-57, mnist.py:54-57)."""


class DType(object):
    def __init__(self, name):
        self.name = name

    @property
    def base_dtype(self):
        return self

    def __repr__(self):
        return "tf." + self.name


float32, float64, uint8, int32, int64 = (DType(n) for n in ("float32", "float64", "uint8", "int32", "int64"))


def as_dtype(d):
    if isinstance(d, DType):
        return d
    raise TypeError("tensorflow stub: unknown dtype %r" % (d,))
