"""GPU: the native (C++/CUDA) fully-connected parameter server -- roles, stale-drop, short-circuit, evaluator output."""
import os

import pytest

pytestmark = pytest.mark.gpu


def test_selftest_trains():
    from pytorch_distributed_nn_b200.parallel.native_fc import selftest
    first, last = selftest(60)
    assert last < first


def test_backup_workers_stale_drop_and_shortcircuit(tmp_path):
    from pytorch_distributed_nn_b200.parallel.native_fc import run_native_fc_ps
    out = str(tmp_path)
    r = run_native_fc_ps(dims=[784, 256, 128, 10], batch=64, lr=0.3, n_workers=4, n_to_collect=2, iters=40, shortcircuit=True,
                         timeline=True, delay_workers=(3,), delay_us_per_layer=3000, synthetic_examples=2048, out_dir=out)
    assert r["steps"] == 40 and r["scheme"] == "SyncReplicasWithBackup2_4_shortcircuit"
    assert r["final_loss"] < r["first_loss"]
    assert r["steps_abandoned"] > 0                      # the delayed worker notices newer steps and gives up
    assert 2.0 <= r["avg_gradients_per_step"] <= 4.0     # at least n_to_collect per layer, at most n
    rows = open(os.path.join(out, "time_loss_out_" + r["scheme"])).read().strip().splitlines()
    assert len(rows) >= 2 and all(len(x.split()) == 4 for x in rows)
    tl = open(os.path.join(out, "timeline_out_" + r["scheme"])).read().strip().splitlines()
    assert sum(1 for x in tl if x.endswith(" 1")) == 40  # one master mark per step


def test_all_workers_collected_without_shortcircuit(tmp_path):
    from pytorch_distributed_nn_b200.parallel.native_fc import run_native_fc_ps
    r = run_native_fc_ps(dims=[784, 100, 10], batch=32, lr=0.3, n_workers=3, n_to_collect=3, iters=15, shortcircuit=False,
                         synthetic_examples=1024, out_dir=str(tmp_path))
    assert r["steps"] == 15 and r["steps_abandoned"] == 0
    assert abs(r["avg_gradients_per_step"] - 3.0) < 1e-6
