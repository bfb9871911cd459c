#!/usr/bin/env python3
"""End-to-end numerics of the B200 engine (forward, loss, every parameter gradient) against the torch fp32
models with identical weights, plus step timings against torch eager (bf16 autocast, channels_last).

    python bench/engine_check.py [LeNet ResNet18 VGG11 MLP ResNet50] [--time]
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False

from pytorch_distributed_nn_b200.engine.network import build_network  # noqa: E402
from pytorch_distributed_nn_b200.models import build_model, input_shape  # noqa: E402


def check(name, batch):
    torch.manual_seed(0)
    net = build_network(name, batch)
    net.init_parameters(seed=1)
    ref = build_model(name).cuda().float()
    sd = net.state_dict()
    if net.dtype == torch.bfloat16:   # the engine computes with bf16 weights: give the oracle the same values
        sd = {k: (v.bfloat16().float() if v.is_floating_point() else v) for k, v in sd.items()}
    missing = ref.load_state_dict(sd, strict=False)
    assert not missing.missing_keys, missing
    ref.train()
    if net.dtype == torch.bfloat16:
        emulate_bf16_storage(ref)
    shp = input_shape(name)
    x = torch.randn(batch, *shp, device="cuda")
    if net.dtype == torch.bfloat16:
        x = x.bfloat16().float()
    y = torch.randint(0, 10, (batch,), device="cuda")
    xin = x.flatten(1) if name in ("MLP", "FC") else x
    loss = net.train_step(xin, y)
    torch.cuda.synchronize()
    out = ref(x)
    rloss = torch.nn.functional.cross_entropy(out, y)
    rloss.backward()
    logits = net.logits[:, :10].float()
    e_logits = float((logits - out).abs().max() / out.abs().max())
    e_loss = abs(float(loss) - float(rloss)) / abs(float(rloss))
    grads = net.export_grads()
    worst = (0.0, None)
    cos_min = 1.0
    gmax = max(float(p.grad.norm()) for p in ref.parameters())
    for k, p in ref.named_parameters():
        g = grads[k]
        r = p.grad
        if float(r.norm()) < 1e-5 * gmax:     # analytically-zero gradients (conv bias in front of BN): pure noise
            continue
        err = float((g - r).abs().max() / r.abs().max().clamp_min(1e-8))
        cos = float(torch.nn.functional.cosine_similarity(g.flatten(), r.flatten(), dim=0))
        cos_min = min(cos_min, cos)
        if err > worst[0]:
            worst = (err, k)
    tol = 1e-3 if net.dtype == torch.float32 else 6e-2
    ok = e_logits < tol and e_loss < tol and cos_min > (0.9999 if net.dtype == torch.float32 else 0.995)
    print(f"ENGINE {name} batch={batch} {'ok' if ok else 'FAIL'} logits_rel={e_logits:.3e} loss_rel={e_loss:.3e} "
          f"worst_grad_rel={worst[0]:.3e} ({worst[1]}) min_grad_cos={cos_min:.5f} loss={float(loss):.4f}", flush=True)
    return ok


def timeit(name, batch, iters=10):
    from pytorch_distributed_nn_b200.ops import _lib
    net = build_network(name, batch)
    net.init_parameters(seed=1)
    shp = input_shape(name)
    x = torch.randn(batch, *shp, device="cuda")
    y = torch.randint(0, 10, (batch,), device="cuda")
    xin = x.flatten(1) if name in ("MLP", "FC") else x
    for _ in range(3):
        net.train_step(xin, y)
    torch.cuda.synchronize()
    _lib.reset_launch_count()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    s.record()
    for _ in range(iters):
        net.train_step(xin, y)
    e.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / iters
    launches = _lib.launch_count() / iters
    # torch eager reference: bf16 autocast + channels_last (cuDNN / cuBLAS)
    ref = build_model(name).cuda()
    if name not in ("MLP", "FC", "LeNet"):
        ref = ref.to(memory_format=torch.channels_last)
        xr = x.contiguous(memory_format=torch.channels_last)
    else:
        xr = x
    opt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.5)

    def ref_step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=name not in ("LeNet",)):
            out = ref(xr)
            loss = torch.nn.functional.cross_entropy(out.float(), y)
        loss.backward()

    for _ in range(3):
        ref_step()
    torch.cuda.synchronize()
    s.record()
    for _ in range(iters):
        ref_step()
    e.record()
    torch.cuda.synchronize()
    rms = s.elapsed_time(e) / iters
    print(f"STEP {name} batch={batch} ours_ms={ms:.3f} ({batch / ms * 1e3:.0f} img/s, {launches:.0f} launches, "
          f"host_issue_ms={(t1 - t0) / iters * 1e3:.3f}) torch_eager_bf16_ms={rms:.3f} ({batch / rms * 1e3:.0f} img/s)", flush=True)


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
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    names = args or ["LeNet", "MLP", "ResNet18", "VGG11"]
    ok = True
    for n in names:
        try:
            ok &= check(n, 64 if n != "LeNet" else 32)
        except Exception as ex:  # keep going: one failing model must not hide the others
            import traceback
            traceback.print_exc()
            print(f"ENGINE {n} EXCEPTION {ex}", flush=True)
            ok = False
    if "--time" in sys.argv:
        for n in names:
            for b in ((128, 1024) if n not in ("LeNet", "MLP") else (128,)):
                try:
                    timeit(n, b)
                except Exception as ex:
                    import traceback
                    traceback.print_exc()
                    print(f"STEP {n} EXCEPTION {ex}", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
