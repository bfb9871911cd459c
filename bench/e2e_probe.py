#!/usr/bin/env python3
"""Where the end-to-end loop (B200ParameterServer.fit) spends its time beyond the device-resident step: the same 40 steps
with (a) resident inputs, (b) fit without loss read-back, (c) fit with the loss read every step. 1 GPU."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer  # noqa: E402

B, K = 1024, 40
ps = B200ParameterServer("ResNet18", B, 0, 1, lr=0.01, momentum=0.5)
hx = [torch.empty(B, 3, 32, 32).uniform_(0, 255).pin_memory() for _ in range(4)]
hy = [torch.randint(0, 10, (B,)).pin_memory() for _ in range(4)]
dx, dy = [t.cuda() for t in hx], [t.cuda() for t in hy]


class L:
    i = 0

    def next_batch(self, b):
        self.i += 1
        return hx[self.i % 4], hy[self.i % 4]


for _ in range(2):
    ps.train_step(dx[0], dy[0])
ps.enable_cuda_graph(dx[0], dy[0])


def timed(fn):
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    fn()
    ps.join()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / K


def resident():
    for i in range(K):
        ps.train_step(dx[i % 4], dy[i % 4])


for rep in range(2):
    print("resident          %.4f ms" % timed(resident))
    print("fit, no loss read %.4f ms" % timed(lambda: ps.fit(L(), ps.step + K - 1, log_interval=1 << 30, graph_after=0)))
    print("fit, loss / step  %.4f ms" % timed(lambda: ps.fit(L(), ps.step + K - 1, log_interval=1 << 30, graph_after=0, read_loss_every_step=True)))
ps.close()
