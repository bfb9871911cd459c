"""GPU: the native (C++/CUDA) fully-connected parameter server -- roles, stale-drop, short-circuit, evaluator output.

Which gradients win a backup-worker step depends on thread timing, so nothing here asserts on a loss curve: the
master logs every (step, worker, layer, batch) it accumulated and an fp32/fp64 PyTorch oracle replays exactly that
set (reference semantics: MPI_code/src/distributed/sync_replicas_master_nn.h:71-128)."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _problem(dims, n, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(n, dims[0], generator=g)
    y = torch.randint(0, dims[-1], (n,), generator=g)
    w0 = []
    for l in range(len(dims) - 1):
        w = torch.randn(dims[l] + 1, dims[l + 1], generator=g) * 0.1
        w[-1].zero_()
        w0.append(w)
    return (x, y), w0


def _max_rel(a, b):
    return max(float((p - q).abs().max() / q.abs().max().clamp_min(1e-12)) for p, q in zip(a, b))


def test_selftest_trains():
    from pytorch_distributed_nn_b200.parallel.native_fc import selftest
    first, last = selftest(200)          # one worker, 784-100-10, lr 0.5: deterministic and far from the noise floor
    assert last < first


def test_backup_workers_match_oracle_for_the_logged_selection(tmp_path):
    from pytorch_distributed_nn_b200.parallel.native_fc import oracle_replay, run_native_fc_ps
    dims, batch, lr, iters, n_workers, k = [784, 256, 128, 10], 64, 0.3, 40, 4, 2
    train, w0 = _problem(dims, 2048, 3)
    out = str(tmp_path)
    r = run_native_fc_ps(dims=dims, batch=batch, lr=lr, n_workers=n_workers, n_to_collect=k, iters=iters, shortcircuit=True,
                         timeline=True, delay_workers=(3,), delay_us_per_layer=3000, out_dir=out, train=train, init_weights=w0,
                         want_weights=True, want_log=True)
    L = len(dims) - 1
    assert r["steps"] == iters and r["scheme"] == "SyncReplicasWithBackup2_4_shortcircuit"
    log = r["accept_log"]
    assert len(log) == r["gradients_accepted"]
    for step in range(1, iters + 1):       # every layer of every step got at least k and at most n distinct workers
        for l in range(L):
            ws = [w for s, w, ll, _ in log if s == step and ll == l]
            assert k <= len(ws) <= n_workers and len(set(ws)) == len(ws), (step, l, ws)
    assert r["steps_abandoned"] > 0                      # the delayed worker notices newer steps and gives up
    assert 2.0 <= r["avg_gradients_per_step"] <= 4.0
    want = oracle_replay(dims, batch, lr, w0, train, log, iters)
    assert _max_rel(r["final_weights"], want) < 2e-4
    assert float((r["final_weights"][0] - w0[0]).abs().max()) > 0
    rows = open(os.path.join(out, "time_loss_out_" + r["scheme"])).read().strip().splitlines()
    assert len(rows) >= 1 and all(len(x.split()) == 4 for x in rows)      # how many snapshots it catches in between is timing
    assert int(rows[-1].split()[0]) == iters + 1         # the evaluator always scores the final model
    tl = open(os.path.join(out, "timeline_out_" + r["scheme"])).read().strip().splitlines()
    assert sum(1 for x in tl if x.endswith(" 1")) == iters  # one master mark per step


def test_host_protocol_still_matches_its_oracle(tmp_path):
    """The thread + condition-variable transport (device_protocol=False) stays available and correct."""
    from pytorch_distributed_nn_b200.parallel.native_fc import oracle_replay, run_native_fc_ps
    dims, batch, lr, iters = [784, 64, 10], 32, 0.3, 10
    train, w0 = _problem(dims, 512, 9)
    r = run_native_fc_ps(dims=dims, batch=batch, lr=lr, n_workers=3, n_to_collect=2, iters=iters, shortcircuit=True, out_dir=str(tmp_path),
                         train=train, init_weights=w0, want_weights=True, want_log=True, device_protocol=False)
    assert r["steps"] == iters
    want = oracle_replay(dims, batch, lr, w0, train, r["accept_log"], iters)
    assert _max_rel(r["final_weights"], want) < 2e-4


def test_all_workers_collected_without_shortcircuit(tmp_path):
    from pytorch_distributed_nn_b200.parallel.native_fc import oracle_replay, run_native_fc_ps
    dims, batch, lr, iters = [784, 100, 10], 32, 0.3, 15
    train, w0 = _problem(dims, 1024, 5)
    r = run_native_fc_ps(dims=dims, batch=batch, lr=lr, n_workers=3, n_to_collect=3, iters=iters, shortcircuit=False,
                         out_dir=str(tmp_path), train=train, init_weights=w0, want_weights=True, want_log=True)
    assert r["steps"] == iters and r["steps_abandoned"] == 0
    assert abs(r["avg_gradients_per_step"] - 3.0) < 1e-6
    want = oracle_replay(dims, batch, lr, w0, train, r["accept_log"], iters)
    assert _max_rel(r["final_weights"], want) < 2e-4
    # all-of-n training on a fixed problem is deterministic enough to assert that it learns
    assert r["final_loss"] < r["first_loss"]
