#!/usr/bin/env python3
"""Run a few parameter-server steps; the last one sits between cudaProfilerStart/Stop so that
``ncu --profile-from-start off`` captures exactly one steady-state step.

    python bench/profile_step.py [network] [batch] [steps_profiled]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer  # noqa: E402


def main():
    network = sys.argv[1] if len(sys.argv) > 1 else "ResNet18"
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    nprof = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    torch.cuda.set_device(0)
    ps = B200ParameterServer(network, batch, 0, 1, lr=0.01, momentum=0.5)
    chw = ps.net.in_chw
    x = torch.empty(batch, *chw, device="cuda").uniform_(0, 255)
    y = torch.randint(0, 10, (batch,), device="cuda")
    for _ in range(3):
        ps.train_step(x, y)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    for _ in range(nprof):
        ps.train_step(x, y)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("profiled ok", float(ps.net.loss))
    ps.close()


if __name__ == "__main__":
    main()
