import sys, os
sys.path.insert(0, os.getcwd())
import torch
from pytorch_distributed_nn_b200.engine.network import build_network
name = sys.argv[1]; batch = int(sys.argv[2])
net = build_network(name, batch); net.init_parameters(1)
x = torch.randn(batch, 3, 32, 32, device="cuda"); y = torch.randint(0, 10, (batch,), device="cuda")
net.zero_grad()
h = net.stage_input(x); torch.cuda.synchronize()
for l in net.layers:
    h = l.forward(h); torch.cuda.synchronize(); print("fwd ok", l.name, getattr(l, "mode", ""), getattr(l, "halo_f", ""), tuple(h.shape), flush=True)
net.logits = h
net.loss_and_grad(y); torch.cuda.synchronize(); print("loss", float(net.loss), flush=True)
dy = net.dlogits
for l in reversed(net.layers):
    dy = l.backward(dy); torch.cuda.synchronize(); print("bwd ok", l.name, getattr(l, "mode", ""), flush=True)
