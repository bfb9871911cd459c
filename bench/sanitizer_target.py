#!/usr/bin/env python3
"""Small-shape workload for compute-sanitizer (memcheck / racecheck / synccheck / initcheck): every kernel family of the
library runs at least once in the configuration the engine uses it in.

    compute-sanitizer --tool memcheck python bench/sanitizer_target.py resnet18     # see tools/run_sanitizer.sh

cases: resnet18 (tcgen05 GEMM / implicit-GEMM convs, halo + stem + strided kernels, BatchNorm, pools, loss, the whole
parameter-server step incl. push / master buckets / weight fetch), vgg11, lenet (fp32 SIMT path), kernels (direct calls
of a GEMM, the delay kernels, the allocation-free BatchNorm)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def ps_steps(network, batch, steps=2, **kw):
    from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer
    ps = B200ParameterServer(network, batch, 0, 1, lr=0.01, momentum=0.5, **kw)
    chw = ps.net.in_chw
    x = torch.rand(batch, *chw, device="cuda") * (255 if chw[0] == 3 else 1)
    y = torch.randint(0, 10, (batch,), device="cuda")
    for _ in range(steps):
        loss = ps.train_step(x, y)
    torch.cuda.synchronize()
    r = ps.last_result()
    assert r[3] == steps and float(loss) == float(loss), (r, float(loss))
    ps.close()


def kernels():
    from pytorch_distributed_nn_b200.ops import _lib
    from pytorch_distributed_nn_b200.ops import functional as F
    from pytorch_distributed_nn_b200.parallel.delays import DeviceDelay
    a = torch.randn(256, 192, device="cuda").bfloat16()
    b = torch.randn(128, 192, device="cuda").bfloat16()
    out = F.gemm(a, b, M=256, N=128, K=192)
    ref = a.float() @ b.float().t()
    assert float((out.float() - ref).abs().max() / ref.abs().max()) < 2e-2
    x = torch.randn(64, 8, 8, 64, device="cuda").bfloat16()
    g, bt = torch.ones(64, device="cuda"), torch.zeros(64, device="cuda")
    y = torch.empty_like(x)
    mi = torch.empty(2, 64, device="cuda")
    _lib.call("pdnn_bn_fwd_noalloc", x, g, bt, None, y, mi, None, None, 64 * 64, 64, 1e-5, 0.1, 1, _lib.stream_ptr())
    stepw = torch.ones(1, device="cuda", dtype=torch.int32)
    dd = DeviceDelay("uniform:lo=0.01:hi=0.02:per=layer", seed=3)
    dd.log = torch.zeros(4, device="cuda", dtype=torch.int32)
    for layer in range(3):
        dd.launch(0, layer, stepw, None, _lib.stream_ptr())
    _lib.call("pdnn_delay", 10_000, None, None, _lib.stream_ptr())
    torch.cuda.synchronize()


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "resnet18"
    if which == "resnet18":
        ps_steps("ResNet18", 32)
    elif which == "vgg11":
        ps_steps("VGG11", 32)
    elif which == "lenet":
        ps_steps("LeNet", 32, grad_dtype="fp32")
    elif which == "kernels":
        kernels()
    else:
        raise SystemExit("unknown case " + which)
    print("SANITIZER_TARGET_OK", which)
