"""GPU: out-of-bounds / skipped-element detection without compute-sanitizer (closed on this pool, profiles/sanitizer/).

Every kernel family writes into an output that lives in the middle of a larger allocation whose surroundings hold a
sentinel; afterwards the sentinels must be untouched and the output must equal the fp32 PyTorch reference. Shapes are
small and odd on purpose: partial M tiles (TMA store boxes clipped at the tensor edge), N tails that leave the vector
paths, 1x1 / strided / stem / halo forms, batch sizes that are not multiples of the tile."""
import pytest
import torch

pytestmark = pytest.mark.gpu
PAD = 4096          # sentinel elements on each side
BF16, F32 = torch.bfloat16, torch.float32


class Guarded:
    """A tensor of ``shape`` carved out of a sentinel-filled buffer (256-byte aligned start)."""

    def __init__(self, shape, dtype, sentinel):
        n = 1
        for s in shape:
            n *= s
        self.buf = torch.full((n + 2 * PAD,), sentinel, device="cuda", dtype=dtype)
        self.t = self.buf[PAD: PAD + n].view(shape)
        self.sentinel, self.n = sentinel, n

    def intact(self):
        lo, hi = self.buf[:PAD].float(), self.buf[PAD + self.n:].float()
        return bool((lo == self.sentinel).all()) and bool((hi == self.sentinel).all())


def rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12))


def conv_ref(x, w, pad, stride=1):
    y = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), padding=pad, stride=stride)
    return y.permute(0, 2, 3, 1)


@pytest.mark.parametrize("M,N,K", [(200, 64, 96), (130, 192, 64), (64, 40, 72), (1000, 256, 128), (33, 8, 24)])
def test_gemm_outputs_stay_inside_their_rows(M, N, K):
    from pytorch_distributed_nn_b200.ops import functional as F
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    bias = torch.randn(N, device="cuda")
    g = Guarded((M, N), BF16, 77.0)
    F.gemm(a, b, M=M, N=N, K=K, out=g.t, bias=bias, relu=True)
    torch.cuda.synchronize()
    assert g.intact()
    assert rel(g.t, torch.relu(a.float() @ b.float().t() + bias)) < 1.5e-2
    g32 = Guarded((M, N), F32, 77.0)
    g32.t.zero_()
    F.gemm(a, b, M=M, N=N, K=K, out=g32.t, accumulate=True)
    torch.cuda.synchronize()
    assert g32.intact() and rel(g32.t, a.float() @ b.float().t()) < 1.5e-2


@pytest.mark.parametrize("n,h,c,k", [(2, 32, 64, 64), (4, 16, 64, 128), (8, 8, 128, 256), (32, 4, 256, 512), (64, 2, 512, 512)])
def test_conv_fprop_dgrad_wgrad_outputs_are_guarded(n, h, c, k):
    from pytorch_distributed_nn_b200.ops import functional as F
    x = torch.randn(n, h, h, c, device="cuda").bfloat16()
    w = (torch.randn(k, 3, 3, c, device="cuda") / (3 * c ** 0.5)).bfloat16()
    res = torch.randn(n, h, h, k, device="cuda").bfloat16()
    y = Guarded((n, h, h, k), BF16, 99.0)
    F.conv_fprop(x, w, pad=1, out=y.t, residual=res, relu=True)
    torch.cuda.synchronize()
    assert y.intact() and rel(y.t, torch.relu(conv_ref(x, w, 1) + res.float())) < 1.5e-2
    if F.conv_halo_supported(tuple(x.shape), k, 3, 3, 1):
        yh = Guarded((n, h, h, k), BF16, 99.0)
        stats = Guarded((2, k), F32, 99.0)
        stats.t.zero_()
        F.conv_halo_fprop(x, w, out=yh.t, bn_stats=stats.t)
        torch.cuda.synchronize()
        ref = conv_ref(x, w, 1)
        assert yh.intact() and stats.intact() and rel(yh.t, ref) < 1.5e-2
        assert rel(stats.t[0], yh.t.float().sum(dim=(0, 1, 2))) < 2e-2
    dy = torch.randn(n, h, h, k, device="cuda").bfloat16()
    dw = Guarded((k, 3, 3, c), F32, 55.0)
    dw.t.zero_()
    F.conv_wgrad(x, dy, R=3, S=3, pad=1, out=dw.t)
    torch.cuda.synchronize()
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(False)
    ref_dw = torch.nn.grad.conv2d_weight(xr, (k, c, 3, 3), dy.float().permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    assert dw.intact() and rel(dw.t, ref_dw) < 2e-2
    if F.conv_halo_wgrad_supported(tuple(x.shape), k, 3, 3, 1):
        dwh = Guarded((k, 3, 3, c), F32, 55.0)
        dwh.t.zero_()
        F.conv_halo_wgrad(x, dy, out=dwh.t)
        torch.cuda.synchronize()
        assert dwh.intact() and rel(dwh.t, ref_dw) < 2e-2


@pytest.mark.parametrize("n,h,c,k,r", [(4, 16, 64, 128, 3), (4, 16, 64, 128, 1), (16, 8, 128, 256, 3)])
def test_strided_conv_family_is_guarded(n, h, c, k, r):
    from pytorch_distributed_nn_b200.ops import functional as F
    pad = 1 if r == 3 else 0
    x = torch.randn(n, h, h, c, device="cuda").bfloat16()
    w = (torch.randn(k, r, r, c, device="cuda") / (r * c ** 0.5)).bfloat16()
    oh = h // 2
    y = Guarded((n, oh, oh, k), BF16, 99.0)
    F.conv_fprop_strided(x, w, pad=pad, stride=2, out=y.t)
    torch.cuda.synchronize()
    assert y.intact() and rel(y.t, conv_ref(x, w, pad, 2)) < 1.5e-2
    dy = torch.randn(n, oh, oh, k, device="cuda").bfloat16()
    dw = Guarded((k, r, r, c), F32, 55.0)
    dw.t.zero_()
    F.conv_wgrad_strided(x, dy, R=r, S=r, pad=pad, stride=2, out=dw.t)
    wt = F.wgt_flip_transpose(w)
    dx = Guarded((n, h, h, c), BF16, 99.0)
    F.conv_dgrad_s2(dy, wt, R=r, S=r, pad=pad, out=dx.t)
    torch.cuda.synchronize()
    ref_dw = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (k, c, r, r), dy.float().permute(0, 3, 1, 2), padding=pad, stride=2)
    ref_dx = torch.nn.grad.conv2d_input((n, c, h, h), w.float().permute(0, 3, 1, 2), dy.float().permute(0, 3, 1, 2), padding=pad, stride=2)
    assert dw.intact() and rel(dw.t, ref_dw.permute(0, 2, 3, 1)) < 2e-2
    assert dx.intact() and rel(dx.t, ref_dx.permute(0, 2, 3, 1)) < 2e-2


@pytest.mark.parametrize("n,h,c,res", [(8, 16, 64, False), (6, 8, 128, True), (32, 2, 512, True), (3, 4, 24, False)])
def test_batchnorm_and_pool_passes_are_guarded(n, h, c, res):
    from pytorch_distributed_nn_b200.ops import functional as F
    x = torch.randn(n, h, h, c, device="cuda").bfloat16()
    r = torch.randn(n, h, h, c, device="cuda").bfloat16() if res else None
    gamma, beta = torch.rand(c, device="cuda") + 0.5, torch.randn(c, device="cuda") * 0.1
    y = Guarded((n, h, h, c), BF16, 99.0)
    stats = Guarded((2, c), F32, 99.0)
    stats.t.zero_()
    mi = Guarded((2, c), F32, 99.0)
    F.bn_forward(x, gamma, beta, relu=True, residual=r, stats=stats.t, mean_invstd=mi.t, out=y.t)
    xf = x.float()
    mean, var = xf.mean(dim=(0, 1, 2)), xf.var(dim=(0, 1, 2), unbiased=False)
    ref = (xf - mean) / torch.sqrt(var + 1e-5) * gamma + beta
    ref = torch.relu(ref + (r.float() if res else 0))
    dy = torch.randn(n, h, h, c, device="cuda").bfloat16()
    dx = Guarded((n, h, h, c), BF16, 99.0)
    dres = Guarded((n, h, h, c), BF16, 99.0)
    sums = Guarded((2, c), F32, 99.0)
    sums.t.zero_()
    dg, db = Guarded((c,), F32, 55.0), Guarded((c,), F32, 55.0)
    dg.t.zero_()
    db.t.zero_()
    F.bn_backward(x, y.t, dy, gamma, mi.t, relu=True, want_dresidual=res, sums=sums.t, dx=dx.t, dresidual=dres.t if res else None,
                  dgamma=dg.t, dbeta=db.t, beta=beta)
    torch.cuda.synchronize()
    for g in (y, stats, mi, dx, sums, dg, db) + ((dres,) if res else ()):
        assert g.intact()
    assert rel(y.t, ref) < 1.5e-2
    if c % 8 == 0 and h % 2 == 0:
        p = Guarded((n, h // 2, h // 2, c), BF16, 99.0)
        a = Guarded((n, c), BF16, 99.0)
        F.maxpool2(x, out=p.t)
        F.avgpool(x, out=a.t)
        torch.cuda.synchronize()
        assert p.intact() and a.intact() and rel(a.t, xf.mean(dim=(1, 2))) < 1e-2
