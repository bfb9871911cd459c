import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("PDNN_PS_TIMEOUT_S", "3")
import torch
from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer, M_ARRIVAL, W_READY, W_CLOSED, W_STEP
net = sys.argv[1] if len(sys.argv) > 1 else "ResNet18"
mode = sys.argv[2] if len(sys.argv) > 2 else "fit"
B = 64
ps = B200ParameterServer(net, B, 0, 1, lr=0.05, momentum=0.5, seed=1)
chw = ps.net.in_chw
xs = [torch.rand(B, *chw) * 255 for _ in range(4)]
ys = [torch.randint(0, 10, (B,)) for _ in range(4)]
class L:
    i = 0
    def next_batch(self, b):
        self.i += 1
        return xs[self.i % 4], ys[self.i % 4]
try:
    if mode == "fit":
        out = ps.fit(L(), 8, log_interval=4, on_log=lambda i, v: print("log", i, v, ps.last_result(), flush=True))
    else:
        for i in range(8):
            ps.train_step(xs[i % 4].cuda(), ys[i % 4].cuda())
        torch.cuda.synchronize()
        out = ps.last_result()
    print("OK", out)
except Exception as e:
    torch.cuda.synchronize()
    f = ps.flags.cpu().tolist()
    print("FAILED", e)
    print("ready/closed/step", f[W_READY], f[W_CLOSED], f[W_STEP], "arrival", f[M_ARRIVAL:M_ARRIVAL + 8], "result", ps.result.cpu().tolist()[:16])
