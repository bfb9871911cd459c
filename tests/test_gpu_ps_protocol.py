"""GPU: the device parameter-server protocol beyond the single master step -- interval deadline, device-sampled
delays, the public training loop (fit), resume of weights AND momentum, and the multi-GPU command line."""
import ctypes as C
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_interval_policy_closes_a_step_at_the_deadline_with_what_arrived():
    """Interval policy on the device (reference: distributed_TF/src/timeout_manager.py, cfg/50_workers/*_interval): only
    worker 1 of 3 has delivered; k = 3 can never be reached, the 20 ms deadline closes the step with that one set."""
    from pytorch_distributed_nn_b200.ops import _lib
    from test_gpu_engine import _master_args
    n, nw, step, stride = 40_000, 3, 4, 8
    dev = "cuda"
    grads = [torch.randn(n, device=dev) for _ in range(nw)]
    flags = torch.zeros(nw, stride, device=dev, dtype=torch.int32)
    flags[1, 0] = step
    land = [[torch.zeros(n, device=dev, dtype=torch.bfloat16) for _ in range(nw)] for _ in range(2)]
    ready = torch.zeros(nw, device=dev, dtype=torch.int32)
    closed = torch.zeros(nw, device=dev, dtype=torch.int32)
    w0 = torch.randn(n, device=dev)
    master_w, mom = w0.clone(), torch.zeros(n, device=dev)
    ctl = torch.zeros(256, device=dev, dtype=torch.int32)
    result = torch.zeros(80, device=dev, dtype=torch.int32)
    args = _master_args(nw, n, 3, grads, flags, stride, land, ready, closed, master_w, mom, ctl, step, [(0, n)], lr=0.1, mu=0.0)
    args.deadline_ns = 20_000_000
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    _lib.call("pdnn_ps_master_step", C.byref(args), result.data_ptr(), _lib.stream_ptr())
    e.record()
    torch.cuda.synchronize()
    r = [int(v) for v in result.tolist()]
    assert tuple(r[:4]) == (0b010, 0b101, 1, step)
    assert 19.0 < s.elapsed_time(e) < 200.0 and r[4] >= 19_000
    assert closed.tolist() == [step, 0, step] and ready.tolist() == [step + 1] * 3
    assert torch.allclose(master_w, w0 - 0.1 * grads[1], rtol=1e-5, atol=1e-6)      # mean over the ONE set that made it


def test_device_sampled_delays_match_the_host_replica():
    from pytorch_distributed_nn_b200.parallel.delays import DeviceDelay
    stepw = torch.zeros(1, device="cuda", dtype=torch.int32)
    st = torch.cuda.current_stream().cuda_stream
    for spec in ("uniform:lo=0.1:hi=0.4:seed=9", "const:ranks=2:ms=0.3", "exp:mean=0.05:p=0.5:seed=3", "pareto:xm=0.01:alpha=2.5:seed=4",
                 "lognormal:mu=-3:sigma=0.5:seed=5"):
        dd = DeviceDelay(spec + ":per=layer", seed=1)
        dd.log = torch.zeros(8, device="cuda", dtype=torch.int32)
        for step in (1, 2, 77):
            stepw.fill_(step)
            for worker in (0, 1):
                for layer in range(3):
                    dd.launch(worker, layer, stepw, None, st)
                torch.cuda.synchronize()
                got = dd.log[:3].tolist()
                want = [dd.sample_ns(worker, step, layer) for layer in range(3)]
                for g, w in zip(got, want):
                    assert abs(g - w) <= max(2, int(2e-3 * w)), (spec, step, worker, got, want)
    dd = DeviceDelay("const:ranks=2:ms=0.3")
    assert dd.sample_ns(1, 5) == 300_000 and dd.sample_ns(0, 5) == 0          # reference spelling: rank 2 = worker 1
    # per=step (default): only the hook of the first backward layer launches; the wait really takes that long
    dd = DeviceDelay("const:ms=15")
    stepw.fill_(1)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for layer in range(4):
        dd.launch(0, layer, stepw, None, st)
    e.record()
    torch.cuda.synchronize()
    assert 14.5 < s.elapsed_time(e) < 30.0


def _close(a, b, tol=2e-3):
    """Relative L2 distance: two runs of the same fp32 training differ by the order of their split-K atomics, and a max-pool
    tie or ReLU gate that flips on a 1e-7 difference moves single weights by far more than an element-wise tolerance."""
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12)) < tol


class _Batches:
    """Host batches handed out like utils.data datasets do (next_batch contract)."""

    def __init__(self, xs, ys):
        self.xs, self.ys, self.i = xs, ys, 0

    def next_batch(self, batch_size):
        j = self.i % len(self.xs)
        self.i += 1
        return self.xs[j], self.ys[j]


def _lenet_batches(n, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(32, 1, 28, 28, generator=g) for _ in range(n)], [torch.randint(0, 10, (32,), generator=g) for _ in range(n)]


def test_fit_loop_with_pinned_staging_and_graph_equals_eager_steps():
    """The loop the roles run (pinned double-buffered staging, CUDA graph after two eager steps) reaches the same master
    weights as feeding the same batches step by step."""
    from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer
    xs, ys = _lenet_batches(7)
    a = B200ParameterServer("LeNet", 32, 0, 1, lr=0.05, momentum=0.5, seed=3, grad_dtype="fp32")
    for x, y in zip(xs, ys):
        a.train_step(x.cuda(), y.cuda())
    torch.cuda.synchronize()
    wa, ma = a.net.w32.clone(), a.momentum.clone()
    a.close()
    b = B200ParameterServer("LeNet", 32, 0, 1, lr=0.05, momentum=0.5, seed=3, grad_dtype="fp32")
    seen = []
    losses = b.fit(_Batches(xs, ys), 7, log_interval=3, on_log=lambda i, v: seen.append(i), read_loss_every_step=True)
    assert b._graph is not None and b.step == 8 and b.last_result() == (1, 0, 1, 7)
    assert [i for i, _ in losses] == list(range(1, 8)) and all(v == v for _, v in losses) and seen == [1, 3, 6, 7]
    assert _close(b.net.w32, wa) and _close(b.momentum, ma)
    b.close()


def test_resume_restores_weights_and_momentum(tmp_path):
    """2N steps == N steps, checkpoint, resume, N steps -- weights AND momentum (the TF reference's Supervisor restores the
    whole optimizer state, distributed_TF/src/distributed_train.py:215-223)."""
    from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer
    from pytorch_distributed_nn_b200.utils import checkpoint as ckpt
    xs, ys = _lenet_batches(6, seed=1)
    kw = dict(lr=0.05, momentum=0.9, seed=5, grad_dtype="fp32")
    a = B200ParameterServer("LeNet", 32, 0, 1, **kw)
    for x, y in zip(xs, ys):
        a.train_step(x.cuda(), y.cuda())
    torch.cuda.synchronize()
    wa, ma = a.net.w32.clone(), a.momentum.clone()
    a.close()
    b = B200ParameterServer("LeNet", 32, 0, 1, **kw)
    for x, y in zip(xs[:3], ys[:3]):
        b.train_step(x.cuda(), y.cuda())
    torch.cuda.synchronize()
    path, _ = ckpt.save_checkpoint(str(tmp_path), 3, b.net.state_dict(), b.momentum)
    b.close()
    sd, step, mom = ckpt.load_checkpoint(path)
    assert step == 3 and mom is not None and float(mom.abs().max()) > 0
    c = B200ParameterServer("LeNet", 32, 0, 1, init_state_dict=sd, init_momentum=mom, **kw)
    for x, y in zip(xs[3:], ys[3:]):
        c.train_step(x.cuda(), y.cuda())
    torch.cuda.synchronize()
    assert _close(c.net.w32, wa) and _close(c.momentum, ma)
    # without the sidecar the continuation differs: the test would not notice a dropped momentum otherwise
    d = B200ParameterServer("LeNet", 32, 0, 1, init_state_dict=sd, **kw)
    for x, y in zip(xs[3:], ys[3:]):
        d.train_step(x.cuda(), y.cuda())
    torch.cuda.synchronize()
    assert not _close(d.net.w32, wa) and not _close(d.momentum, ma, 5e-2)      # a dropped momentum is far beyond the noise
    c.close()
    d.close()


def _torchrun(nproc, args, port, timeout=900):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(port)] + args
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("policy", ["kofn", "interval"])
def test_cli_two_gpus_k_less_than_n_learns_from_the_selected_worker(tmp_path, policy):
    """The public command line on 2 GPUs with k = 1 of n = 2 (the configuration that silently averaged zeros when the
    workers were not told k): the master's weights must move, every logged step must name exactly one selected and one
    killed worker (k-of-n) / at least one selected (interval), and the checkpoint must differ from the initial weights."""
    d = str(tmp_path)
    extra = ["--num-aggregate", "1"] if policy == "kofn" else ["--straggler-policy", "interval", "--interval-ms", "2.0", "--num-aggregate", "2"]
    r = _torchrun(2, ["distributed_nn.py", "--network", "LeNet", "--dataset", "MNIST", "--synthetic", "--batch-size", "64", "--backend", "b200",
                      "--colocate-worker", "--max-steps", "8", "--log-interval", "1", "--eval-freq", "8", "--train-dir", d, "--lr", "0.05",
                      "--grad-dtype", "fp32", "--inject-delay", "const:ranks=2:ms=6"] + extra, 29581 if policy == "kofn" else 29582)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    rows = [json.loads(l) for l in open(os.path.join(d, "master_rank0.jsonl")) if '"selected"' in l]
    assert len(rows) == 8
    for row in rows:
        if policy == "kofn":
            assert row["selected"] == [0] and row["killed"] == [1], row        # worker 1 carries the injected 6 ms delay
        else:
            assert 0 in row["selected"], row                                     # whoever beat the 2 ms deadline; worker 0 always does
    sd = torch.load(os.path.join(d, "model_step_8"))
    from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer
    init = B200ParameterServer("LeNet", 32, 0, 1, seed=1, grad_dtype="fp32")
    sd0 = init.net.state_dict()
    init.close()
    moved = max(float((sd[k] - sd0[k].cpu()).abs().max()) for k in sd0)
    assert moved > 1e-4, "the master's weights did not move: gradients never reached the slots it reads"
    assert all(bool(torch.isfinite(v).all()) for v in sd.values())


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_cli_evaluator_rank_on_the_device_backend(tmp_path):
    """--evaluator on the b200 backend: the last rank receives every broadcast and writes time_loss_out_<scheme> rows
    (reference: MPI_code/src/distributed/evaluator_nn.h:29-60)."""
    d = str(tmp_path)
    r = _torchrun(2, ["distributed_nn.py", "--network", "LeNet", "--dataset", "MNIST", "--synthetic", "--batch-size", "64", "--backend", "b200",
                      "--colocate-worker", "--evaluator", "--max-steps", "12", "--eval-freq", "4", "--log-interval", "4", "--train-dir", d,
                      "--lr", "0.05", "--test-batch-size", "256"], 29583)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    path = os.path.join(d, "time_loss_out_SyncReplicasWithBackup1_1")
    rows = [l.split() for l in open(path) if not l.startswith("#")]
    assert len(rows) >= 2 and int(rows[-1][0]) == 12 and all(len(x) == 4 and float(x[2]) == float(x[2]) for x in rows)
    steps = [int(x[0]) for x in rows]
    assert steps == sorted(steps)


def test_abort_guard_covers_the_layer_kernels():
    """A killed worker skips not only its GEMMs / convs but also the BatchNorm, pooling and element-wise passes (about a
    third of a ResNet step): with the guard closed the kernels leave their outputs untouched."""
    from pytorch_distributed_nn_b200.ops import functional as F
    x = torch.randn(256, 8, 8, 64, device="cuda").bfloat16()
    closed = torch.tensor([5], device="cuda", dtype=torch.int32)
    stepw = torch.tensor([5], device="cuda", dtype=torch.int32)
    y = torch.full_like(x, 7.0)
    pooled = torch.full((256, 4, 4, 64), 7.0, device="cuda", dtype=torch.bfloat16)
    avg = torch.full((256, 64), 7.0, device="cuda", dtype=torch.bfloat16)
    F.set_abort_guard(closed, stepw)
    try:
        F.relu(x, out=y)
        F.maxpool2(x, out=pooled)
        F.avgpool(x, out=avg)
        F.add(x, x, out=y)
        torch.cuda.synchronize()
        assert float(y.float().min()) == 7.0 and float(pooled.float().min()) == 7.0 and float(avg.float().min()) == 7.0
        stepw.fill_(6)                      # the worker moved on to a step that is still open
        F.relu(x, out=y)
        F.maxpool2(x, out=pooled)
        torch.cuda.synchronize()
        assert torch.equal(y, torch.relu(x)) and torch.equal(pooled, torch.nn.functional.max_pool2d(x.permute(0, 3, 1, 2).float(), 2).permute(0, 2, 3, 1).bfloat16())
    finally:
        F.set_abort_guard(None, None)
    F.relu(x, out=y)
    torch.cuda.synchronize()
    assert torch.equal(y, torch.relu(x))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_master_only_checkpoint_carries_trained_batchnorm_statistics(tmp_path):
    """Reference topology (rank 0 = parameter server only): the checkpoint's BatchNorm running statistics are the first
    worker's, not the initial zeros / ones a rank that never runs forward would have."""
    d = str(tmp_path)
    r = _torchrun(2, ["distributed_nn.py", "--network", "VGG11", "--dataset", "Cifar10", "--synthetic", "--batch-size", "64", "--backend", "b200",
                      "--max-steps", "6", "--eval-freq", "6", "--log-interval", "3", "--train-dir", d, "--lr", "0.01"], 29584)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    sd = torch.load(os.path.join(d, "model_step_6"))
    means = [v for k, v in sd.items() if k.endswith("running_mean")]
    vars_ = [v for k, v in sd.items() if k.endswith("running_var")]
    assert means and all(float(m.abs().max()) > 0 for m in means) and any(float((v - 1).abs().max()) > 1e-3 for v in vars_)
    assert all(int(v) == 6 for k, v in sd.items() if k.endswith("num_batches_tracked"))
