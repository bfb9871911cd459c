"""CPU multi-process tests (gloo, world_size 3): BASELINE config 1 -- LeNet on MNIST-shape data, master + 2 workers."""
import json
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.slow


def _run(extra, tmp_path, nprocs=3, timeout=240):
    from pytorch_distributed_nn_b200.parallel.launch import spawn_local
    args = [os.path.join(ROOT, "distributed_nn.py"), "--network", "LeNet", "--dataset", "MNIST", "--synthetic", "--no-cuda",
            "--batch-size", "16", "--log-interval", "1", "--train-dir", str(tmp_path)] + extra
    outs = spawn_local(args, nprocs, timeout=timeout)
    for rc, o in outs:
        assert rc == 0, o[-3000:]
    return outs


def _master_log(tmp_path):
    with open(os.path.join(str(tmp_path), "master_rank0.jsonl")) as f:
        return [json.loads(l) for l in f if '"step"' in l]


def test_all_n_training_equals_single_process_replay(tmp_path):
    """Master + 2 workers, wait-for-all: the checkpointed weights equal a single-process replay that averages the two
    workers' gradients (same init seed, same per-worker data streams) -- i.e. the wire protocol loses nothing."""
    steps, lr, mu = 4, 0.05, 0.5
    _run(["--max-steps", str(steps), "--straggler-policy", "all", "--lr", str(lr), "--momentum", str(mu), "--eval-freq", "4",
          "--grad-dtype", "fp32", "--comm-type", "Bcast"], tmp_path)
    from pytorch_distributed_nn_b200 import mnist
    from pytorch_distributed_nn_b200.models import build_model
    from pytorch_distributed_nn_b200.utils import checkpoint as ckpt
    sd, step, _ = ckpt.load_checkpoint(ckpt.latest_checkpoint(str(tmp_path))[0])
    assert step == steps
    torch.manual_seed(1)                                   # --seed default: the master's init
    net = build_model("LeNet", split=True)
    data = [mnist.read_data_sets("MNIST_data", reshape=True, synthetic=True, worker_id=r, n_workers=3).train for r in (1, 2)]
    params = list(net.parameters())
    mom = [torch.zeros_like(p) for p in params]
    for _ in range(steps):
        gsum = [torch.zeros_like(p) for p in params]
        for d in data:
            x, y = d.next_batch(16)
            net.zero_grad()
            out = net(torch.from_numpy(x).float())
            loss = torch.nn.functional.cross_entropy(out, torch.from_numpy(y))
            loss.backward()
            net.backward(out.grad, None, [])
            for g, p in zip(gsum, params):
                g += p.grad
        with torch.no_grad():
            for p, g, m in zip(params, gsum, mom):
                m.mul_(mu).add_(g / 2)
                p.add_(m, alpha=-lr)
    for (k, v), p in zip(net.named_parameters(), params):
        assert torch.allclose(sd[k], p.detach(), rtol=1e-5, atol=1e-6), k
    log = _master_log(tmp_path)
    assert [r["selected"] for r in log] == [[0, 1]] * steps and all(r["killed"] == [] for r in log)


def test_k_of_n_kills_the_injected_straggler(tmp_path):
    """k = 1 of 2 with a 400 ms injected delay on rank 2: every step the master selects worker 0, names worker 1 in
    the kill list, and rank 2 abandons its backward (no host round trip besides the kill message itself)."""
    outs = _run(["--max-steps", "4", "--num-aggregate", "1", "--inject-delay", "const:ranks=2:ms=400", "--comm-type", "Async"], tmp_path)
    log = _master_log(tmp_path)
    assert len(log) == 4
    # step 1 is dominated by process start-up skew; from step 2 on the injected delay decides
    assert all(r["selected"] == [0] and r["killed"] == [1] for r in log[1:])
    with open(os.path.join(str(tmp_path), "worker_rank2.jsonl")) as f:
        w2 = [json.loads(l) for l in f]
    assert sum(1 for r in w2 if r["killed"]) >= 2
    with open(os.path.join(str(tmp_path), "worker_rank1.jsonl")) as f:
        assert not any(json.loads(l)["killed"] for l in list(f)[1:])


def test_evaluator_rank_and_interval_policy(tmp_path):
    """4 ranks: master, 2 workers, evaluator (C++ reference role). Interval policy: the master closes each step at the
    deadline with whatever arrived (TF reference's interval mode)."""
    _run(["--max-steps", "3", "--straggler-policy", "interval", "--interval-ms", "150", "--inject-delay", "const:ranks=2:ms=600",
          "--evaluator"], tmp_path, nprocs=4)
    log = _master_log(tmp_path)
    assert all(r["selected"] == [0] for r in log[1:])
    rows = open(os.path.join(str(tmp_path), "time_loss_out_SyncReplicas")).read().strip().splitlines()
    assert len(rows) == 3 and all(len(r.split()) == 4 for r in rows)


def test_backup_dist_example_two_ranks_stay_in_sync():
    """pytorch_code/backup/dist_nn_test.py counterpart on gloo: all-reduce-averaged replicas end bit-equal (asserted
    inside the script) and rank 0 logs the loss."""
    from pytorch_distributed_nn_b200.parallel.launch import spawn_local
    outs = spawn_local(["-m", "pytorch_distributed_nn_b200.backup.dist_nn_test", "--steps", "6", "--batch-size", "16"], 2, timeout=180,
                       env={"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")})
    for rc, o in outs:
        assert rc == 0, o[-2000:]
    assert "Train step 5 world 2" in outs[0][1]


def test_comm_test_two_ranks_deliver_all_eight_lenet_gradients():
    """pytorch_code/comm_test/comm_test.py counterpart: rank 1 streams the eight LeNet parameter gradients (float64 wire),
    rank 0 receives every one of them with its shape."""
    from pytorch_distributed_nn_b200.parallel.launch import spawn_local
    outs = spawn_local(["-m", "pytorch_distributed_nn_b200.comm_test"], 2, timeout=180,
                       env={"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")})
    for rc, o in outs:
        assert rc == 0, o[-2000:]
    got = [l for l in outs[0][1].splitlines() if l.startswith("master received grad of tensor")]
    assert len(got) == 8
    assert any("shape (500, 800)" in l for l in got) and any("shape (20, 1, 5, 5)" in l for l in got)


_DDP_SNIPPET = r"""
import torch, torch.distributed as dist
from pytorch_distributed_nn_b200.parallel.launch import init_distributed
from pytorch_distributed_nn_b200.data_parallel_dist import DistributedDataParallel
from pytorch_distributed_nn_b200.models import build_model
rank, world, _ = init_distributed("gloo")
torch.manual_seed(0)
net = DistributedDataParallel(build_model("LeNet"))
g = torch.Generator().manual_seed(100 + rank)
x = torch.randn(4, 1, 28, 28, generator=g)
net(x).square().mean().backward()
local = [p.grad.clone() for p in net.module.parameters()]
net.finish_backward()
flat_local = torch.cat([t.flatten() for t in local])
gathered = [torch.empty_like(flat_local) for _ in range(world)]
dist.all_gather(gathered, flat_local)
want = sum(gathered) / world
got = torch.cat([p.grad.flatten() for p in net.module.parameters()])
assert torch.allclose(got, want, atol=1e-6), float((got - want).abs().max())
print("DDP_OK rank", rank, "buckets", len(net.buckets))
dist.barrier(); dist.destroy_process_group()
"""


def test_ddp_comparator_two_ranks_average_the_bucketed_gradients():
    """Vendored-DDP counterpart (pytorch_code/data_parallel_dist/data_parallel_dist.py:70-83, 238-264): bucketed all-reduce
    leaves every replica with the mean of the local gradients."""
    from pytorch_distributed_nn_b200.parallel.launch import spawn_local
    outs = spawn_local(["-c", _DDP_SNIPPET], 2, timeout=180, env={"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")})
    for rc, o in outs:
        assert rc == 0, o[-2000:]
        assert "DDP_OK" in o


def test_resume_continues_weights_step_number_and_momentum(tmp_path):
    """``--resume model_step_<n>`` on the message backend: the job continues at step n+1 from the checkpointed weights AND the
    momentum of its sidecar (TF Supervisor restores everything, distributed_train.py:215-223). Checked against a
    single-process replay in which the two workers' data streams start again (they do: workers are new processes) and the
    momentum is carried over; a replay that starts the momentum at zero must NOT match."""
    lr, mu = 0.05, 0.9
    common = ["--straggler-policy", "all", "--lr", str(lr), "--momentum", str(mu), "--grad-dtype", "fp32", "--comm-type", "Bcast"]
    a, b = tmp_path / "a", tmp_path / "b"
    _run(["--max-steps", "2", "--eval-freq", "2"] + common, a)
    from pytorch_distributed_nn_b200.utils import checkpoint as ckpt
    first = ckpt.latest_checkpoint(str(a))
    assert first is not None and first[1] == 2
    _run(["--max-steps", "4", "--eval-freq", "4", "--resume", first[0]] + common, b)
    sd, step, _ = ckpt.load_checkpoint(ckpt.latest_checkpoint(str(b))[0])
    assert step == 4
    assert [r["step"] for r in _master_log(b)] == [3, 4]
    from pytorch_distributed_nn_b200 import mnist
    from pytorch_distributed_nn_b200.models import build_model

    def replay(carry_momentum):
        torch.manual_seed(1)
        net = build_model("LeNet", split=True)
        params = list(net.parameters())
        mom = [torch.zeros_like(p) for p in params]
        for leg in range(2):                                # two jobs of two steps; each job's workers read their data from the start
            data = [mnist.read_data_sets("MNIST_data", reshape=True, synthetic=True, worker_id=r, n_workers=3).train for r in (1, 2)]
            if leg == 1 and not carry_momentum:
                mom = [torch.zeros_like(p) for p in params]
            for _ in range(2):
                gsum = [torch.zeros_like(p) for p in params]
                for d in data:
                    x, y = d.next_batch(16)
                    net.zero_grad()
                    out = net(torch.from_numpy(x).float())
                    torch.nn.functional.cross_entropy(out, torch.from_numpy(y)).backward()
                    net.backward(out.grad, None, [])
                    for g, p in zip(gsum, params):
                        g += p.grad
                with torch.no_grad():
                    for p, g, m in zip(params, gsum, mom):
                        m.mul_(mu).add_(g / 2)
                        p.add_(m, alpha=-lr)
        return {k: p.detach().clone() for (k, _), p in zip(net.named_parameters(), params)}

    good, bad = replay(True), replay(False)
    for k in good:
        assert torch.allclose(sd[k], good[k], rtol=1e-5, atol=1e-6), k
    assert any(not torch.allclose(sd[k], bad[k], rtol=1e-3, atol=1e-4) for k in bad)
