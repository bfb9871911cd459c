#!/usr/bin/env python3
"""Layer-by-layer comparison of the B200 engine against the torch fp32 model (debug aid).

    python bench/engine_debug.py ResNet50 [batch]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
from pytorch_distributed_nn_b200.engine.network import build_network  # noqa: E402
from pytorch_distributed_nn_b200.engine import layers as L  # noqa: E402
from pytorch_distributed_nn_b200.models import build_model, input_shape  # noqa: E402


def emulate_bf16_storage(model):
    """Make the fp32 oracle round activations to bf16 at the points where the engine stores them (conv / linear /
    BN outputs and block outputs). Without this, ReLU masks of near-zero activations flip between the two runs and
    the comparison measures that chaos instead of kernel correctness."""
    from pytorch_distributed_nn_b200.models import resnet as R

    def rnd(mod, inp, out):
        return out.bfloat16().float()   # the cast's backward also rounds the gradient to bf16, as the engine does

    for m in model.modules():
        if isinstance(m, (torch.nn.Conv2d, torch.nn.Linear, torch.nn.BatchNorm2d, R.BasicBlock, R.Bottleneck)):
            m.register_forward_hook(rnd)


def main():
    name = sys.argv[1]
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    torch.manual_seed(0)
    net = build_network(name, batch)
    net.init_parameters(seed=1)
    ref = build_model(name).cuda().float()
    sd = {k: (v.bfloat16().float() if v.is_floating_point() and net.dtype == torch.bfloat16 else v) for k, v in net.state_dict().items()}
    ref.load_state_dict(sd, strict=False)
    ref.train()
    if net.dtype == torch.bfloat16 and os.environ.get("PDNN_NO_EMULATE") != "1":
        emulate_bf16_storage(ref)
    acts = {}

    def hook(nm):
        def f(mod, inp, out):
            acts[nm] = out.detach()
        return f

    for nm, mod in ref.named_modules():
        if isinstance(mod, (torch.nn.Conv2d, torch.nn.BatchNorm2d, torch.nn.Linear)):
            mod.register_forward_hook(hook(nm))
    x = torch.randn(batch, *input_shape(name), device="cuda").bfloat16().float()
    y = torch.randint(0, 10, (batch,), device="cuda")
    loss = net.train_step(x, y)
    if os.environ.get("PDNN_FORCE") == "1":
        # teacher-forced oracle: torch's forward VALUES are replaced by the engine's stored activations at every
        # conv / linear / block output, so both backward passes differentiate the same forward state (same ReLU
        # masks, same BN statistics). Isolates the backward kernels from bf16 forward chaos.
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

        for nm, mod in ref.named_modules():
            if nm in eng:
                mod.register_forward_hook(force(nm))
    out = ref(x)
    rl = torch.nn.functional.cross_entropy(out, y)
    rl.backward()
    print("loss", float(loss), float(rl))
    for layer in net._all_layers():
        if isinstance(layer, (L.Conv2d, L.Linear)) and layer.name in acts:
            a = acts[layer.name]
            mine = layer.y.float()
            if a.dim() == 4:
                a = a.permute(0, 2, 3, 1)
            else:
                mine = mine[:, : a.shape[1]]
            err = float((mine - a).abs().max() / a.abs().max())
            mode = getattr(layer, "mode", "linear")
            print(f"FWD {layer.name:28s} {mode:9s} rel={err:.3e} {'<<<' if err > 3e-2 else ''}")
    grads = net.export_grads()
    for k, p in ref.named_parameters():
        g, r = grads[k], p.grad
        cos = float(torch.nn.functional.cosine_similarity(g.flatten(), r.flatten(), dim=0))
        rel = float((g - r).abs().max() / r.abs().max().clamp_min(1e-12))
        flag = "<<<" if cos < 0.99 else ""
        print(f"GRAD {k:34s} cos={cos:.5f} rel={rel:.3e} |ref|={float(r.norm()):.3e} |got|={float(g.norm()):.3e} {flag}")


if __name__ == "__main__":
    main()
