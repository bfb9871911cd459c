"""Environment shims that let the UNMODIFIED reference (Python 2 / torch <= 0.3 era) run under Python 3.12 and
torch 2.11. Nothing under ``baseline/_ref`` is edited; everything here changes the interpreter the reference is
run in, before ``runpy`` starts its script. ``apply(device, autocast)`` installs:

always
  * ``builtins.xrange / unicode / basestring / long``               (cifar10.py:104, my_data_loader.py:14)
  * ``torch.autograd.stochastic_function``                           (distributed_backward.py:6, removed in torch 0.4)
  * ``tensor[0]`` on a 0-dim tensor answers ``tensor.item()``        (``loss.data[0]``, distributed_worker.py:163)
  * ``y.backward(g)`` reshapes ``g`` when only its shape (not its size) differs (lenet.py:196)
  * ``state_dict()`` lists no BatchNorm ``num_batches_tracked`` keys   (state_dict walk, distributed_worker.py:252-264)
  * ``pandas.isnull(tensor_or_None)`` answers ``x is None``          (resnet_split.py:244; torch <= 0.3 Variables were
                                                                     opaque objects to pandas, tensors are array-likes)
  * ``Tensor.numpy()`` detaches first                                (``layer.data.numpy()`` etc. are unaffected; a
                                                                     requires_grad view would otherwise raise)
device = "cuda:N"   (the reference's own scripts never move anything to a GPU: --no-cuda is parsed and ignored,
                     distributed_nn.py:52. This mode gives it the B200 anyway so that the comparison is not
                     "GPU vs CPU"; ``device="cpu"`` runs the stock path exactly as written.)
  * ``torch.set_default_device(device)``: its ``nn.Module``s are built on the GPU
  * ``torch.autograd.Variable(t)`` moves ``t`` to the device (the reference wraps every input batch, label batch
    and inter-layer activation in ``Variable(...)``), keeps its legacy meaning otherwise (detached, shares storage)
  * ``Tensor.numpy()`` on a GPU tensor copies to the host first (fp32), i.e. the device -> host read a GPU user of
    the reference would have had to write
autocast = "bf16"   wraps the run in ``torch.autocast("cuda", torch.bfloat16)`` so that cuDNN / cuBLAS use the
                     tensor cores in the same compute dtype as our arm
"""
import builtins
import sys
import types


def apply(device="cpu", autocast="none"):
    import torch

    # ---- Python 2 names ------------------------------------------------------------------------------
    for name, val in (("xrange", range), ("unicode", str), ("basestring", str), ("long", int)):
        if not hasattr(builtins, name):
            setattr(builtins, name, val)

    # ---- modules removed from torch ------------------------------------------------------------------
    if "torch.autograd.stochastic_function" not in sys.modules:
        m = types.ModuleType("torch.autograd.stochastic_function")

        class StochasticFunction(torch.autograd.Function):
            pass

        m.StochasticFunction = StochasticFunction
        sys.modules["torch.autograd.stochastic_function"] = m
        torch.autograd.stochastic_function = m

    # ---- 0-dim indexing (loss.data[0]) ------------------------------------------------------------------
    _getitem = torch.Tensor.__getitem__

    def getitem(self, idx):
        if self.dim() == 0 and isinstance(idx, int) and idx == 0:
            return self.item()
        return _getitem(self, idx)

    torch.Tensor.__getitem__ = getitem

    # ---- y.backward(g) with g of equal size but different shape (lenet.py:196 feeds a [B, 800] gradient to a
    #      [B, 50, 4, 4] output; torch <= 0.3 only compared element counts) ------------------------------------
    _backward = torch.Tensor.backward

    def backward(self, gradient=None, *a, **k):
        if isinstance(gradient, torch.Tensor) and gradient.shape != self.shape and gradient.numel() == self.numel():
            gradient = gradient.reshape(self.shape)
        return _backward(self, gradient, *a, **k)

    torch.Tensor.backward = backward

    # ---- state_dict() without the ``num_batches_tracked`` keys (buffer added in torch 0.4.1): the reference walks
    #      state_dict() and treats every key that is not running_mean / running_var as a parameter
    #      (distributed_worker.py:252-264, sync_replicas_master_nn.py:283-296) ---------------------------------
    _state_dict = torch.nn.Module.state_dict

    def state_dict(self, *a, **k):
        sd = _state_dict(self, *a, **k)
        for key in [key for key in sd if key.endswith("num_batches_tracked")]:
            del sd[key]
        return sd

    torch.nn.Module.state_dict = state_dict

    # ---- pandas.isnull on tensors ---------------------------------------------------------------------
    try:
        import pandas as pd
        _isnull = pd.isnull

        def isnull(obj):
            if obj is None:
                return True
            if isinstance(obj, torch.Tensor):
                return False
            return _isnull(obj)

        pd.isnull = isnull
        pd.isna = isnull
    except ImportError:
        pass

    # ---- Tensor.numpy ---------------------------------------------------------------------------------
    _numpy = torch.Tensor.numpy

    def numpy(self, *a, **k):
        t = self.detach()
        if t.is_cuda:
            t = t.float().cpu() if t.dtype in (torch.bfloat16, torch.float16) else t.cpu()
        elif t.dtype in (torch.bfloat16, torch.float16):
            t = t.float()
        return _numpy(t, *a, **k)

    torch.Tensor.numpy = numpy

    # ---- GPU placement --------------------------------------------------------------------------------
    if device != "cpu":
        dev = torch.device(device)
        torch.cuda.set_device(dev)
        torch.set_default_device(dev)

        class _VariableMeta(type(torch.Tensor)):
            def __instancecheck__(cls, other):
                return isinstance(other, torch.Tensor)

            def __call__(cls, data=None, requires_grad=False, volatile=False, **kw):
                if data is None:
                    data = torch.empty(0)
                t = data.detach()
                if t.device != dev:
                    t = t.to(dev, non_blocking=True)
                if requires_grad:
                    t = t.requires_grad_(True) if t is not data else t.detach().requires_grad_(True)
                return t

        class Variable(torch.Tensor, metaclass=_VariableMeta):
            pass

        torch.autograd.Variable = Variable
        import torch.autograd.variable as _v
        _v.Variable = Variable

    ctx = None
    if autocast == "bf16" and device != "cpu":
        ctx = torch.autocast("cuda", dtype=torch.bfloat16)
    elif autocast == "bf16":
        ctx = torch.autocast("cpu", dtype=torch.bfloat16)
    return ctx
