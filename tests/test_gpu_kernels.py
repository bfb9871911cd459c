"""GPU numerics: every sm_100a kernel against a plain PyTorch fp32 reference of the same op
(the comparisons live in bench/kernel_check.py so that the same code is the gpurun diagnostic)."""
import importlib.util
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("kernel_check", os.path.join(ROOT, "bench", "kernel_check.py"))
kc = importlib.util.module_from_spec(spec)
spec.loader.exec_module(kc)

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", sorted(kc.CASES))
def test_kernel_case(name):
    import torch
    torch.manual_seed(0)
    assert kc.CASES[name](), f"kernel check {name} reported a mismatch (see captured stdout)"
    torch.cuda.synchronize()


def test_library_is_loaded_and_counts_launches():
    import torch
    from pytorch_distributed_nn_b200.ops import _lib
    from pytorch_distributed_nn_b200.ops import functional as F
    _lib.reset_launch_count()
    x = torch.randn(1024, device="cuda").bfloat16()
    F.relu(x)
    assert _lib.launch_count() == 1
    with open("/proc/self/maps") as f:
        assert "libpdnn_b200.so" in f.read()


def test_errors_are_loud():
    import torch
    from pytorch_distributed_nn_b200.ops import _lib
    from pytorch_distributed_nn_b200.ops import functional as F
    x = torch.randn(2, 8, 8, 48, device="cuda").bfloat16()      # C not a multiple of 64
    w = torch.randn(64, 3, 3, 48, device="cuda").bfloat16()
    with pytest.raises(_lib.PdnnError):
        F.conv_fprop(x, w, pad=1)
