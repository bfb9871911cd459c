"""Shared helpers: compare the B200 engine with the torch fp32 models.

bf16 networks at random init are chaotic in the forward direction (a 0.2% rounding difference flips ReLU masks of
near-zero activations and gradients are heavily cancelling sums), so backward is verified *teacher-forced*: the
torch model's forward VALUES are replaced by the engine's stored activations at every conv / linear / block output,
which makes both backward passes differentiate the same forward state.
"""
import torch

from pytorch_distributed_nn_b200.engine import layers as L
from pytorch_distributed_nn_b200.engine.network import build_network
from pytorch_distributed_nn_b200.models import build_model, input_shape


def make_pair(name, batch, seed=1):
    net = build_network(name, batch)
    net.init_parameters(seed=seed)
    ref = build_model(name).cuda().float()
    sd = net.state_dict()
    if net.dtype == torch.bfloat16:
        sd = {k: (v.bfloat16().float() if v.is_floating_point() else v) for k, v in sd.items()}
    res = ref.load_state_dict(sd, strict=False)
    assert not res.missing_keys, res
    ref.train()
    return net, ref


def make_batch(name, batch, bf16):
    x = torch.randn(batch, *input_shape(name), device="cuda")
    if bf16:
        x = x.bfloat16().float()
    y = torch.randint(0, 10, (batch,), device="cuda")
    return x, y


def teacher_force(net, ref):
    eng = {}
    for layer in net._all_layers():
        if isinstance(layer, (L.Conv2d, L.Linear)):
            eng[layer.name] = layer.y
    for blk in net.layers:
        if isinstance(blk, (L.BasicBlock, L.Bottleneck)):
            eng[blk.name] = (blk.bn3 if isinstance(blk, L.Bottleneck) else blk.bn2).y

    def force(nm):
        def f(mod, inp, out):
            e = eng[nm].float()
            e = e.permute(0, 3, 1, 2) if e.dim() == 4 else e[:, : out.shape[1]]
            return out + (e - out).detach()
        return f

    handles = [mod.register_forward_hook(force(nm)) for nm, mod in ref.named_modules() if nm in eng]
    return handles


def l2_rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12))


def grad_report(net, ref):
    grads = net.export_grads()
    gmax = max(float(p.grad.norm()) for p in ref.parameters())
    out = {}
    for k, p in ref.named_parameters():
        if float(p.grad.norm()) < 1e-5 * gmax:      # analytically zero (conv bias in front of BN)
            continue
        g = grads[k]
        out[k] = (float(torch.nn.functional.cosine_similarity(g.flatten(), p.grad.flatten(), dim=0)), l2_rel(g, p.grad))
    return out
