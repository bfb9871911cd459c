"""CPU: tooling -- cluster config / host files, experiment configs, straggler-study driver (dry run), plot scripts."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cluster_cfg_interpolation_and_host_files(tmp_path):
    sys.path.insert(0, ROOT)
    from tools.cluster import get_hosts, load_cfg
    cfg_path = tmp_path / "c.cfg"
    cfg_path.write_text('{"name": "t", "hosts": ["10.0.0.1", "10.0.0.2"], "gpus_per_host": 8, "train_dir": "/tmp/%(name)s_x"}')
    cfg = load_cfg(str(cfg_path))
    assert cfg["train_dir"] == "/tmp/t_x"
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        get_hosts(cfg)
        assert open("hosts_address").read().splitlines() == ["10.0.0.1 slots=8", "10.0.0.2 slots=8"]
        assert open("hosts_alias").read().split() == ["deeplearning-worker0", "deeplearning-worker1"]
    finally:
        os.chdir(cwd)


def test_experiment_configs_load_and_map_to_commands():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench", "straggler_study.py"), "--dry-run"] +
                       [os.path.join(ROOT, "configs", d, f) for d in ("k_sweep", "interval_sweep", "time_cdf")
                        for f in sorted(os.listdir(os.path.join(ROOT, "configs", d)))],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines() if l.startswith("#")]
    assert len(lines) == 18
    assert any("--num-aggregate 6" in l for l in lines) and any("--interval-ms 5" in l for l in lines)
    assert any("--worker-times-cdf" in l and "pareto" in l for l in lines)


def test_plot_scripts_summarise_reference_file_formats(tmp_path):
    sys.path.insert(0, ROOT)
    from tools.plot_time_loss import main as plot_main
    from tools.visualize_timeline import summarize
    out = tmp_path / "outfiles"
    out.mkdir()
    (out / "time_loss_out_SyncReplicasWithBackup2_6").write_text("1 10.0 2.30 0.90\n2 20.0 2.10 0.80\n")
    series = plot_main(str(out), str(tmp_path / "p.png"))
    assert series["SyncReplicasWithBackup2_6"][-1][2] == 2.10
    tl = tmp_path / "timeline"
    tl.write_text("0.0 1 1\n1.0 1 0\n2.0 1 0\n5.0 2 1\n6.0 2 0\n")
    steps, avg = summarize(str(tl))
    assert abs(avg - 1.5) < 1e-9 and len(steps[1]["arrivals"]) == 2


def test_backup_hook_walkthrough_order_and_scaling():
    """pytorch_code/backup/back_ward_hook_test.py counterpart: the tensor hook fires between the two layers' backward
    and scales what the lower layer receives by 0.1."""
    from pytorch_distributed_nn_b200.backup.back_ward_hook_test import run
    events, ratio = run(log=lambda *_: None)
    assert events.count("function backward") == 2 and "module hook" in events and "tensor hook" in events
    assert events.index("tensor hook") < len(events) - 1 and events[0] == "function backward"
    assert abs(ratio - 0.1) < 1e-5


def test_backup_dist_example_single_process_loss_falls():
    from pytorch_distributed_nn_b200.backup.dist_nn_test import train
    losses = train(steps=60, batch_size=64, lr=0.05, log=lambda *_: None)
    assert sum(losses[-10:]) / 10 < sum(losses[:10]) / 10


def test_cluster_download_verbs(tmp_path):
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("cluster_tool", os.path.join(os.path.dirname(os.path.dirname(__file__)), "tools", "cluster.py"))
    ct = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ct)
    src = tmp_path / "remote" / "outfiles"
    src.mkdir(parents=True)
    (src / "time_loss_out_x").write_text("1 2 3 4\n")
    cfg = ct.Cfg({"hosts": ["127.0.0.1"], "repo_dir": str(tmp_path), "ssh_user": "nobody"})
    got = ct.download(cfg, str(src), str(tmp_path / "local"), recursive=True)
    assert len(got) == 1 and os.path.exists(os.path.join(got[0], "time_loss_out_x"))
    got = ct.download(cfg, str(src / "time_loss_out_x"), str(tmp_path / "local2"), recursive=False)
    assert len(got) == 1 and open(got[0]).read() == "1 2 3 4\n"


def test_roofline_tool_groups_a_launch_list():
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(__file__))
    spec = importlib.util.spec_from_file_location("roofline_tool", os.path.join(root, "tools", "roofline_from_launches.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    t, w = m.main([os.path.join(root, "profiles", "launches_resnet18_b1024_v10.csv"), "--network", "ResNet18", "--batch", "1024"])
    assert t["batchnorm"] > 1000 and t["fprop+dgrad"] > t["wgrad"] > 1000        # microseconds
    assert abs(w["fprop"] / 1e12 - 1.137) < 0.01 and w["params"] > 11_000_000


def test_data_prepare_synthesizes_real_file_formats_once(tmp_path):
    """data_prepare writes the datasets in their on-disk formats; the package's readers then parse those files (the same
    path real data takes) and a second run only verifies them."""
    from pytorch_distributed_nn_b200.utils import data_prepare
    from pytorch_distributed_nn_b200.utils.data import read_cifar10, read_mnist
    md, cd = str(tmp_path / "m"), str(tmp_path / "c")
    assert "absent" in data_prepare.prepare_mnist(md)["status"]
    out = data_prepare.main(["--mnist-dir", md, "--cifar-dir", cd, "--synthesize"])
    assert [r["status"] for r in out] == ["synthesized", "synthesized"]
    assert out[0]["train"] == 60000 and out[0]["test"] == 10000 and out[1]["train"] == 50000 and out[1]["test"] == 10000
    again = data_prepare.main(["--mnist-dir", md, "--cifar-dir", cd])
    assert [r["status"] for r in again] == ["found", "found"]
    a, b = read_cifar10(cd, worker_id=0), read_cifar10(cd, worker_id=0)
    xa, ya = a.train.next_batch(16)
    xb, yb = b.train.next_batch(16)
    assert xa.shape == (16, 3, 32, 32) and (xa == xb).all() and (ya == yb).all()      # every rank reads the same files
    assert read_mnist(md).train.images.shape[1:] == (1, 28, 28)


def test_native_fc_driver_builds_and_parses_its_flags():
    """build/distributed_nn_native (counterpart of MPI_code's `make distributed`): links against the in-tree library,
    documents its flags and refuses unknown ones. Running a job needs a GPU (tests/test_gpu_native_fc.py)."""
    import importlib.util
    import shutil
    import pytest
    if shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"):
        pytest.skip("no nvcc")
    spec = importlib.util.spec_from_file_location("pdnn_build", os.path.join(ROOT, "csrc", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    if not os.path.exists(mod.LIB):
        pytest.skip("libpdnn_b200.so not built")
    exe = mod.build_native_driver()
    r = subprocess.run([exe, "--help"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "--collect" in r.stdout and "--host-protocol" in r.stdout
    r = subprocess.run([exe, "--no-such-flag"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "unknown argument" in r.stderr


def test_plot_study_reads_a_train_dir_and_a_sweep_file(tmp_path):
    """tools/plot_study.py (analysis half of distributed_TF/tools/benchmark.py:60-263): the five figures' series and the
    compute-time statistics (all samples / slowest worker per iteration) from the JSONL files a run leaves behind."""
    import json
    sys.path.insert(0, ROOT)
    from tools import plot_study
    run = tmp_path / "run"
    run.mkdir()
    with open(run / "master_rank0.jsonl", "w") as f:
        for s in range(1, 11):
            f.write(json.dumps({"step": s, "step_ms": 10.0 + s, "ts": 100.0 + s, "rank": 0}) + "\n")
        f.write(json.dumps({"event": "summary", "steps": 10, "ts": 111.0}) + "\n")
    for rank, slow in ((1, 1.0), (2, 3.0)):
        with open(run / f"worker_rank{rank}.jsonl", "w") as f:
            for s in range(1, 11):
                f.write(json.dumps({"step": s, "loss": 2.3 / s, "total_s": 0.01 * slow * s, "ts": 100.0 + s - 0.5, "rank": rank}) + "\n")
    with open(run / "time_loss_out_test", "w") as f:
        f.write("5 5000.0 1.2 0.40\n10 10000.0 0.6 0.25\n")
    study = tmp_path / "study.json"
    study.write_text(json.dumps([{"name": "c1", "policy": "kofn", "k": 3, "interval_ms": 0.0, "inject_delay": "none", "images_per_sec": 1000.0,
                                  "step_ms": {"p50": 1.0, "p99": 2.0}, "sets_per_step": 3.0}, {"name": "broken", "error": "x"}]))
    rep = plot_study.main([f"a={run}", "--out", str(tmp_path / "figs"), "--study", str(study)])
    st = rep["a"]["compute_times_ms"]
    assert rep["a"]["steps"] == 10 and st["all"]["n"] == 20 and abs(st["all"]["max"] - 300.0) < 1e-6
    assert abs(st["slowest_worker_per_iteration"]["p50"] - 150.0) < 1e-6          # worker 2 is always the slowest: 30 ms x step
    assert [r["name"] for r in rep["study"]] == ["c1"]
    assert len(rep["figures"]) == 5 and all(os.path.exists(p) for p in rep["figures"])
    if rep["figures"][0].endswith(".csv"):
        rows = open(os.path.join(tmp_path, "figs", "time_precision.csv")).read().splitlines()
        assert rows[1:] == ["a,5,60", "a,10,75"]


def test_chrome_trace_from_a_train_dir(tmp_path):
    """tools/chrome_trace.py (the TF study's --timeline_logging, distributed_train.py:284-286): spans of the worker phases and of
    the master's step / device timeline, one file for the run and one per step."""
    import json
    sys.path.insert(0, ROOT)
    from tools import chrome_trace
    run = tmp_path / "run"
    run.mkdir()
    with open(run / "master_rank0.jsonl", "w") as f:
        for s in (1, 2):
            f.write(json.dumps({"step": s, "step_ms": 8.0, "ts": 50.0 + 0.01 * s, "selected": [0, 1], "killed": [2],
                                "device_timeline_us": {"wait_for_kth_arrival": 5000, "reduce_update_broadcast": 15, "bucket_ready_us": [1000, 3000, 5000]}}) + "\n")
    with open(run / "worker_rank1.jsonl", "w") as f:
        for s in (1, 2):
            f.write(json.dumps({"step": s, "loss": 2.0, "killed": s == 2, "total_s": 0.007, "fetch_s": 0.001, "forward_s": 0.002, "backward_s": 0.004,
                                "ts": 50.0 + 0.01 * s - 0.001}) + "\n")
    files = chrome_trace.main([str(run), "-o", str(tmp_path / "trace.json"), "--per-step"])
    ev = json.load(open(files[0]))["traceEvents"]
    names = [e["name"] for e in ev]
    assert names.count("step") == 2 and names.count("forward") == 2 and "backward + push (killed)" in names
    assert sum(1 for n in names if n.startswith("bucket ")) == 6
    step1 = [e for e in ev if e["ph"] == "X" and e["args"]["step"] == 1]
    m = {e["name"]: e for e in step1}
    # the tail ends where the step ends; the last bucket's instant sits at the end of the wait span
    assert abs((m["reduce + update + broadcast tail"]["ts"] + 15) - (m["step"]["ts"] + m["step"]["dur"])) < 1e-3
    w = m["wait for k-th arrival (last bucket)"]
    last_bucket = [e for e in ev if e["name"] == "bucket 2 has k arrivals" and e["args"]["step"] == 1][0]
    assert abs(last_bucket["ts"] - (w["ts"] + w["dur"])) < 1e-3
    assert m["forward"]["ts"] == pytest_approx(m["fetch weights"]["ts"] + 1000.0)
    assert len(files) == 3 and all(os.path.exists(p) for p in files)
    per = json.load(open(files[1]))["traceEvents"]
    assert {e["args"]["step"] for e in per if e["ph"] != "M"} == {1}


def pytest_approx(v):
    import pytest
    return pytest.approx(v, abs=1e-2)
