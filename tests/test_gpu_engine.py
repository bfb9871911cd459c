"""GPU: the B200 engine (forward, loss, all parameter gradients) vs the torch fp32 models; state_dict round trip;
the parameter-server master kernel vs the k-of-n oracle."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fp32_reference():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)


@pytest.mark.parametrize("name,batch", [("LeNet", 32), ("MLP", 64)])
def test_fp32_networks_match_torch_exactly(name, batch):
    from _oracle import grad_report, l2_rel, make_batch, make_pair
    net, ref = make_pair(name, batch)
    x, y = make_batch(name, batch, False)
    loss = net.train_step(x.flatten(1) if name == "MLP" else x, y)
    out = ref(x)
    rl = torch.nn.functional.cross_entropy(out, y)
    rl.backward()
    assert l2_rel(net.logits[:, :10], out) < 1e-5
    assert abs(float(loss) - float(rl)) < 1e-5
    for k, (cos, rel) in grad_report(net, ref).items():
        assert cos > 0.99999 and rel < 1e-4, (k, cos, rel)


@pytest.mark.parametrize("name,batch", [("ResNet18", 64), ("VGG11", 64), ("ResNet50", 32)])
def test_bf16_networks_forward_and_teacher_forced_backward(name, batch):
    from _oracle import grad_report, l2_rel, make_batch, make_pair, teacher_force
    net, ref = make_pair(name, batch)
    x, y = make_batch(name, batch, True)
    loss = net.train_step(x, y)
    teacher_force(net, ref)
    out = ref(x)
    rl = torch.nn.functional.cross_entropy(out, y)
    rl.backward()
    assert l2_rel(net.logits[:, :10], out) < 2e-2
    assert abs(float(loss) - float(rl)) / float(rl) < 5e-3
    rep = grad_report(net, ref)
    bad = {k: v for k, v in rep.items() if v[0] < 0.995}
    assert not bad, bad


def test_state_dict_roundtrip_matches_torch_layout():
    from pytorch_distributed_nn_b200.engine.network import build_network
    from pytorch_distributed_nn_b200.models import build_model
    for name in ("LeNet", "ResNet18", "VGG11"):
        net = build_network(name, 32)
        net.init_parameters(seed=3)
        sd = net.state_dict()
        ref = build_model(name)
        rsd = ref.state_dict()
        assert list(sd.keys()) == list(rsd.keys()), name
        for k in sd:
            assert tuple(sd[k].shape) == tuple(rsd[k].shape), (name, k)
        net2 = build_network(name, 32)
        net2.load_state_dict(sd)
        assert torch.equal(net2.w32, net.w32)


def test_training_reduces_loss_through_parameter_server():
    from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer
    ps = B200ParameterServer("ResNet18", 32, 0, 1, lr=0.05, momentum=0.5)
    x = torch.empty(32, 3, 32, 32, device="cuda").uniform_(0, 255)
    y = torch.randint(0, 10, (32,), device="cuda")
    losses = [float(ps.train_step(x, y)) for _ in range(6)]
    assert losses[-1] < 0.5 * losses[0], losses
    assert ps.last_result() == (1, 0, 1, 6)
    ps.close()


def test_ps_single_worker_update_equals_momentum_sgd():
    from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer
    ps = B200ParameterServer("LeNet", 32, 0, 1, lr=0.1, momentum=0.9, grad_dtype="fp32")
    net = ps.net
    x = torch.randn(32, 1, 28, 28, device="cuda")
    y = torch.randint(0, 10, (32,), device="cuda")
    w = net.w32.clone()
    m = torch.zeros_like(w)
    for _ in range(3):
        ps.train_step(x, y)
        torch.cuda.synchronize()
        g = net.g.clone()
        m = 0.9 * m + g
        w = w - 0.1 * m
        assert torch.allclose(net.w32, w, rtol=1e-5, atol=1e-6)
        # fp32 network: the broadcast copy (landing buffer of the NEXT step's parity) == master copy
        assert torch.equal(ps.land[ps.step & 1][: net.total], net.w32)
    ps.close()


def _master_args(n_workers, n, k, slots, flags, stride, land, ready, closed, master_w, mom, ctl, step, buckets, avg_mode=0,
                 bf16_slots=False, lr=0.1, mu=0.5):
    from pytorch_distributed_nn_b200.ops import _lib
    m = _lib.PSMaster()
    for w in range(n_workers):
        m.slots[w] = slots[w].data_ptr()
        for par in range(2):
            m.recv_weights[par][w] = land[par][w].data_ptr()
        m.recv_ready[w] = ready.data_ptr() + 4 * w
        m.worker_closed[w] = closed.data_ptr() + 4 * w
    m.arrival_flags = flags.data_ptr()
    m.arrival_stride = stride
    m.master_w = master_w.data_ptr()
    m.momentum = mom.data_ptr()
    m.control = ctl.data_ptr()
    m.n, m.n_workers, m.n_receivers, m.k = n, n_workers, n_workers, k
    m.slot_is_bf16 = int(bf16_slots)
    m.weights_are_bf16 = 1
    m.avg_mode = avg_mode
    m.lr, m.mu = lr, mu
    m.step = step
    m.deadline_ns = 0
    m.max_ctas = 0
    m.n_buckets = len(buckets)
    for b, (lo, hi) in enumerate(buckets):
        m.bucket_lo[b], m.bucket_hi[b] = lo, hi
    return m


@pytest.mark.parametrize("avg_mode", ["k", "n"])
def test_master_kernel_k_of_n_matches_oracle(avg_mode):
    """All ranks' data on one GPU (the multi-rank path is emulated over every worker's slot), two buckets, k = 3 of 4,
    step 7: bucket 0 (arrives first, the deep half of the model) has workers 0, 2, 3, bucket 1 has workers 0, 1, 3.
    Each bucket is averaged over ITS first k arrivals (reference: per-layer accumulation in arrival order,
    sync_replicas_master_nn.py:148-166); the kill set is who missed the last bucket (:172-187)."""
    from pytorch_distributed_nn_b200.ops import _lib
    from pytorch_distributed_nn_b200.parallel.oracle import kofn_update
    from pytorch_distributed_nn_b200.parallel.protocol import AVG_MODES
    n, nw, k, step, stride = 100_000, 4, 3, 7, 8
    buckets = [(60_000, 100_000), (0, 60_000)]
    dev = "cuda"
    grads = [torch.randn(n, device=dev) for _ in range(nw)]
    flags = torch.zeros(nw, stride, device=dev, dtype=torch.int32)
    flags[:, 0] = torch.tensor([7, 6, 7, 7])
    flags[:, 1] = torch.tensor([7, 7, 3, 7])
    land = [[torch.zeros(n, device=dev, dtype=torch.bfloat16) for _ in range(nw)] for _ in range(2)]
    ready = torch.zeros(nw, device=dev, dtype=torch.int32)
    closed = torch.zeros(nw, device=dev, dtype=torch.int32)
    w0 = torch.randn(n, device=dev)
    m0 = torch.randn(n, device=dev) * 0.1
    master_w, mom = w0.clone(), m0.clone()
    ctl = torch.zeros(256, device=dev, dtype=torch.int32)
    result = torch.zeros(80, device=dev, dtype=torch.int32)
    args = _master_args(nw, n, k, grads, flags, stride, land, ready, closed, master_w, mom, ctl, step, buckets, AVG_MODES[avg_mode])
    _lib.call("pdnn_ps_master_step", C.byref(args), result.data_ptr(), _lib.stream_ptr())
    torch.cuda.synchronize()
    r = [int(v) for v in result.tolist()]
    assert tuple(r[:4]) == (0b1011, 0b0100, 3, 7) and r[7] == 2 and r[8:10] == [0b1101, 0b1011] and r[6] == 3
    assert closed.tolist() == [0, 0, 7, 0]                 # only the worker that missed the last bucket abandons step 7
    assert ready.tolist() == [8, 8, 8, 8]                  # everybody is released into step 8
    for (lo, hi), who in zip(buckets, ([0, 2, 3], [0, 1, 3])):
        w_ref, m_ref, selected, _ = kofn_update(w0[lo:hi].cpu(), m0[lo:hi].cpu(), {i: g[lo:hi].cpu() for i, g in enumerate(grads)}, who, k, nw,
                                                lr=0.1, mu=0.5, avg_mode=avg_mode)
        assert selected == who
        assert torch.allclose(master_w[lo:hi].cpu(), w_ref, rtol=1e-5, atol=1e-6)
        assert torch.allclose(mom[lo:hi].cpu(), m_ref, rtol=1e-5, atol=1e-6)
    for w in range(nw):
        assert torch.equal(land[0][w], master_w.bfloat16())        # weights of step 8 land in parity 0
        assert float(land[1][w].float().abs().max()) == 0.0        # the buffer step 7 is still being read from is untouched
    # a second step through the same control block (counters were reset): step 8, everybody present
    flags[:] = 8
    args.step = 8
    _lib.call("pdnn_ps_master_step", C.byref(args), result.data_ptr(), _lib.stream_ptr())
    torch.cuda.synchronize()
    r = [int(v) for v in result.tolist()]
    assert tuple(r[:4]) == (0b0111, 0b1000, 3, 8) and r[6] == 6 and ready.tolist() == [9, 9, 9, 9]
    assert torch.equal(land[1][0], master_w.bfloat16())


def test_push_abort_in_mid_grid_keeps_the_handshake_consistent():
    """A kill can land while a push grid is running: some CTAs copy, some abort. The completion counter must come back to
    zero and the arrival flag must stay down for that launch -- and the next, un-aborted push must publish normally."""
    from pytorch_distributed_nn_b200.ops import _lib
    n = 4_000_000
    g = torch.randn(n, device="cuda")
    slot = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
    flag = torch.zeros(1, device="cuda", dtype=torch.int32)
    stepw = torch.tensor([5], device="cuda", dtype=torch.int32)
    for trial in range(20):
        closed = torch.tensor([4], device="cuda", dtype=torch.int32)
        side = torch.cuda.Stream()
        _lib.call("pdnn_ps_push", g, slot, n, 1, flag, 5, closed, stepw, 8, _lib.stream_ptr())      # 8 CTAs: a long grid
        with torch.cuda.stream(side):
            closed.fill_(5)                                                                          # the kill, racing with the grid
        torch.cuda.synchronize()
        flag_after_race = int(flag)
        assert flag_after_race in (0, 5)
        flag.zero_()
        closed.fill_(4)
        torch.cuda.synchronize()
        _lib.call("pdnn_ps_push", g, slot, n, 1, flag, 5, closed, stepw, 8, _lib.stream_ptr())
        torch.cuda.synchronize()
        assert int(flag) == 5 and torch.equal(slot, g.bfloat16()), trial     # a stuck counter would publish early or never
        flag.zero_()


def test_push_kernel_casts_and_flags():
    from pytorch_distributed_nn_b200.ops import _lib
    n = 50_000
    g = torch.randn(n, device="cuda")
    slot = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
    flag = torch.zeros(1, device="cuda", dtype=torch.int32)
    _lib.call("pdnn_ps_push", g, slot, n, 1, flag, 5, None, None, 0, _lib.stream_ptr())
    torch.cuda.synchronize()
    assert int(flag) == 5 and torch.equal(slot, g.bfloat16())
    # a worker whose step was closed drops its push: slot and flag stay untouched
    closed = torch.tensor([6], device="cuda", dtype=torch.int32)
    stepw = torch.tensor([6], device="cuda", dtype=torch.int32)
    slot2 = torch.zeros_like(slot)
    _lib.call("pdnn_ps_push", g, slot2, n, 1, flag, 6, closed, stepw, 0, _lib.stream_ptr())
    torch.cuda.synchronize()
    assert int(flag) == 5 and float(slot2.float().abs().max()) == 0.0


def test_delay_kernel_busy_waits_and_aborts():
    from pytorch_distributed_nn_b200.ops import _lib
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    _lib.call("pdnn_delay", 20_000_000, None, None, _lib.stream_ptr())
    e.record()
    torch.cuda.synchronize()
    assert 19.0 < s.elapsed_time(e) < 40.0
    closed = torch.tensor([3], device="cuda", dtype=torch.int32)
    stepw = torch.tensor([3], device="cuda", dtype=torch.int32)
    s.record()
    _lib.call("pdnn_delay", 200_000_000, closed, stepw, _lib.stream_ptr())
    e.record()
    torch.cuda.synchronize()
    assert s.elapsed_time(e) < 20.0


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_multi_gpu_parameter_server_checks():
    """torchrun with one rank per GPU (2): all-n (incl. NVLS multicast / in-switch reduce when the fabric offers it),
    bf16 wire, k-of-n with an injected device-side delay -- each compared with the oracle (bench/ps_multigpu_check.py)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29577", os.path.join(root, "bench", "ps_multigpu_check.py"), "--network", "LeNet"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("PSCHECK") >= 3 and "FAIL" not in r.stdout


def test_cuda_graph_replay_matches_eager_steps():
    """Same init, same batches: 5 eager steps vs 2 eager + 1 capture warm-up + 2 graph-replayed steps give the same master
    weights (LeNet: fp32 path, so the only run-to-run difference is the order of split-K atomics)."""
    from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer
    x = [torch.randn(32, 1, 28, 28, device="cuda") for _ in range(5)]
    y = [torch.randint(0, 10, (32,), device="cuda") for _ in range(5)]
    a = B200ParameterServer("LeNet", 32, 0, 1, lr=0.05, momentum=0.5, seed=3, grad_dtype="fp32")
    for i in range(5):
        a.train_step(x[i], y[i])
    torch.cuda.synchronize()
    wa = a.net.w32.clone()
    a.close()
    b = B200ParameterServer("LeNet", 32, 0, 1, lr=0.05, momentum=0.5, seed=3, grad_dtype="fp32")
    b.train_step(x[0], y[0])
    b.train_step(x[1], y[1])
    b.enable_cuda_graph(x[2], y[2])        # runs one real step on its example batch (step 3) and captures
    from pytorch_distributed_nn_b200.ops import _lib
    b.train_step(x[3], y[3])
    _lib.reset_launch_count()
    loss = b.train_step(x[4], y[4])
    torch.cuda.synchronize()
    # a replay runs the captured library kernels without passing the host entry points: the counter still sees them
    assert b.graph_kernel_launches > 10 and _lib.launch_count() == b.graph_kernel_launches
    assert b.last_result() == (1, 0, 1, 5) and float(loss) == float(loss)
    assert float((b.net.w32 - wa).norm() / wa.norm()) < 2e-3
    b.close()


def test_poisoned_slots_leave_training_unchanged_and_end_up_nan():
    """Race-detection mode of the master kernel: every summed slot is overwritten with NaN before the workers are
    released. Training must not notice (each step's gradients are rewritten before they are summed), and the slot must be
    all-NaN after a step."""
    from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer
    x = [torch.randn(32, 1, 28, 28, device="cuda") for _ in range(3)]
    y = [torch.randint(0, 10, (32,), device="cuda") for _ in range(3)]
    out = []
    for poison in (False, True):
        ps = B200ParameterServer("LeNet", 32, 0, 1, lr=0.05, momentum=0.5, seed=3, grad_dtype="fp32", poison_slots=poison)
        for i in range(3):
            ps.train_step(x[i], y[i])
        torch.cuda.synchronize()
        assert ps.last_result()[3] == 3
        if poison:
            assert bool(torch.isnan(ps.local_slot(0)[: ps.net.total]).all())
        out.append(ps.net.w32.clone())
        ps.close()
    assert bool(torch.isfinite(out[1]).all()) and float((out[0] - out[1]).norm() / out[0].norm()) < 2e-3
