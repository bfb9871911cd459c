"""The allocation-free BatchNorm entry point, called through the C ABI (collected last on purpose).

``pdnn_bn_fwd_noalloc`` exists for users of libpdnn_b200.so who cannot provide the [2][C] statistics buffer that the
regular pair ``pdnn_bn_fwd_stats`` + ``pdnn_bn_fwd_apply`` needs: plain-C hosts without a device allocator, processes
whose HBM is already fully reserved by another runtime's pool (a raw cudaMalloc fails there), and step loops that must
not risk the device-wide stall an allocation can bring. The statistics then sit in a fixed array inside the library.
Here it is compared with the regular pair on two calls in a row -- the second with fewer channels and other data, which
only comes out right if the shared array is cleared per call -- and it must refuse channel counts beyond its limit.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

_P, _I, _L, _F = C.c_void_p, C.c_int, C.c_longlong, C.c_float


def _entry_points():
    from pytorch_distributed_nn_b200.ops import _lib
    h = _lib.lib()
    fn = h.pdnn_bn_fwd_noalloc
    fn.argtypes = [_P, _P, _P, _P, _P, _P, _P, _P, _L, _I, _F, _F, _I, _P]
    fn.restype = C.c_int
    stats = h.pdnn_bn_fwd_stats
    stats.argtypes = [_P, _P, _L, _I, _P]
    stats.restype = C.c_int
    apply_ = h.pdnn_bn_fwd_apply
    apply_.argtypes = [_P, _P, _P, _P, _P, _P, _P, _P, _P, _L, _I, _F, _F, _I, _P]
    apply_.restype = C.c_int
    return h, fn, stats, apply_


def _run_both(shape, seed, relu, with_residual):
    h, noalloc, stats_fn, apply_fn = _entry_points()
    g = torch.Generator(device="cuda").manual_seed(seed)
    c = shape[-1]
    m = 1
    for d in shape[:-1]:
        m *= d
    x = (torch.randn(shape, device="cuda", generator=g) * 1.7 + 0.3).bfloat16()
    res = torch.randn(shape, device="cuda", generator=g).bfloat16() if with_residual else None
    gamma = torch.rand(c, device="cuda", generator=g) + 0.5
    beta = torch.randn(c, device="cuda", generator=g)
    st = torch.cuda.current_stream().cuda_stream
    rp = res.data_ptr() if res is not None else None
    # existing entry points, caller-allocated scratch
    stats = torch.zeros(2, c, device="cuda")
    y_ref = torch.empty_like(x)
    mi_ref = torch.empty(2, c, device="cuda")
    rm_ref, rv_ref = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
    assert stats_fn(x.data_ptr(), stats.data_ptr(), m, c, st) == 0, h.pdnn_last_error()
    assert apply_fn(x.data_ptr(), stats.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rp, y_ref.data_ptr(), mi_ref.data_ptr(),
                    rm_ref.data_ptr(), rv_ref.data_ptr(), m, c, 1e-5, 0.1, int(relu), st) == 0, h.pdnn_last_error()
    # allocation-free entry point
    y = torch.empty_like(x)
    mi = torch.empty(2, c, device="cuda")
    rm, rv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
    rc = noalloc(x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rp, y.data_ptr(), mi.data_ptr(), rm.data_ptr(), rv.data_ptr(),
                 m, c, 1e-5, 0.1, int(relu), st)
    assert rc == 0, h.pdnn_last_error()
    torch.cuda.synchronize()

    def rel(a, b):
        return float((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-6))

    assert rel(y, y_ref) <= 1.5e-2            # bf16 output tolerance of the kernel tests
    assert rel(mi, mi_ref) <= 2e-3 and rel(rm, rm_ref) <= 2e-3 and rel(rv, rv_ref) <= 2e-3
    # and both agree with plain PyTorch fp32
    xr = x.float().reshape(m, c)
    out = torch.nn.functional.batch_norm(xr, None, None, gamma, beta, training=True)
    if res is not None:
        out = out + res.float().reshape(m, c)
    if relu:
        out = torch.relu(out)
    assert rel(y.reshape(m, c), out) <= 1.5e-2


def test_noalloc_bn_matches_regular_entry_points_twice_in_a_row():
    _run_both((8, 16, 16, 256), seed=1, relu=True, with_residual=True)
    # second call: fewer channels, different data -> stale sums in the shared array would show up here
    _run_both((4, 8, 8, 64), seed=2, relu=False, with_residual=False)


def test_noalloc_bn_rejects_sizes_beyond_its_scratch():
    h, noalloc, _, _ = _entry_points()
    c = 4096
    x = torch.zeros(16, c, device="cuda", dtype=torch.bfloat16)
    y = torch.empty_like(x)
    mi = torch.empty(2, c, device="cuda")
    g = torch.ones(c, device="cuda")
    b = torch.zeros(c, device="cuda")
    rc = noalloc(x.data_ptr(), g.data_ptr(), b.data_ptr(), None, y.data_ptr(), mi.data_ptr(), None, None, 16, c, 1e-5, 0.1, 0,
                 torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    assert b"exceeds the built-in scratch" in h.pdnn_last_error()
