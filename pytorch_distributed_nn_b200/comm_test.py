"""Two-rank communication smoke test (reference pytorch_code/comm_test/comm_test.py): rank 1 runs ONE LeNet
forward/backward and sends every parameter gradient (tag = 88 + index, deepest first); rank 0 receives them through
the master's receiver thread and prints the shapes. Run:

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 -m pytorch_distributed_nn_b200.comm_test
"""
import queue

import torch


def main():
    from .models import build_model
    from .parallel.comm import TorchDistCommunicator
    from .parallel.launch import init_distributed
    rank, world, _ = init_distributed("gloo")
    assert world == 2, "comm_test runs with exactly 2 ranks"
    comm = TorchDistCommunicator("gloo", wire_dtype=torch.float64)        # the reference ships float64
    torch.manual_seed(0)
    net = build_model("LeNet", split=True)
    shapes = [tuple(p.shape) for p in net.parameters()]
    received = {}
    if rank == 0:
        bufs = [[torch.zeros(p.numel(), dtype=torch.float64)] for p in net.parameters()]
        q = queue.Queue()
        comm.start_grad_receivers([1], bufs, q)
        while len(received) < len(shapes):
            item = q.get(timeout=60)
            if item[1] >= 0:
                received[item[1]] = item[3].view(shapes[item[1]])
                print("master received grad of tensor", item[1], "shape", tuple(received[item[1]].shape), "step", item[2])
        comm.join_grad_receivers(10)
    else:
        comm.cur_step = 1
        out = net(torch.randn(8, 1, 28, 28))
        torch.nn.functional.cross_entropy(out, torch.randint(0, 10, (8,))).backward()
        reqs = net.backward(out.grad, comm, [])
        comm.wait_all(reqs)
        comm.wait_all([comm.send_grad_eos()])
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()
    return received


if __name__ == "__main__":
    main()
