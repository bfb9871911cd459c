"""CPU unit tests: k-of-n oracle, delay models, checkpoints, data readers, layer-split autograd, flags."""
import argparse
import os

import numpy as np
import pytest
import torch


def test_select_first_k_counts_complete_sets_in_arrival_order():
    from pytorch_distributed_nn_b200.parallel.oracle import select_first_k
    assert select_first_k([3, 1, 1, 0, 2], 2, 4) == ([1, 3], [0, 2])
    assert select_first_k([0, 1, 2], 3, 3) == ([0, 1, 2], [])
    assert select_first_k([2], 1, 3) == ([2], [0, 1])


@pytest.mark.parametrize("avg_mode,scale", [("k", 1 / 2), ("n", 1 / 4), ("received", 1 / 2)])
def test_kofn_update_matches_hand_computation(avg_mode, scale):
    from pytorch_distributed_nn_b200.parallel.oracle import kofn_update
    torch.manual_seed(0)
    w, m = torch.randn(50), torch.randn(50) * 0.1
    grads = {i: torch.randn(50) for i in range(4)}
    new_w, new_m, sel, killed = kofn_update(w, m, grads, [2, 0, 3, 1], k=2, n_workers=4, lr=0.1, mu=0.9, avg_mode=avg_mode)
    assert sel == [0, 2] and killed == [1, 3]
    g = (grads[0] + grads[2]) * scale
    assert torch.allclose(new_m, 0.9 * m + g)
    assert torch.allclose(new_w, w - 0.1 * (0.9 * m + g))


def test_momentum_zero_reproduces_reference_plain_sgd():
    """``--avg-mode=n --momentum 0`` is the python reference's update: param -= lr * sum/n (sync_replicas_master_nn.py:212-219)."""
    from pytorch_distributed_nn_b200.parallel.oracle import kofn_update
    from pytorch_distributed_nn_b200.sync_replicas_master_nn import update_params_dist_version
    w = torch.randn(20)
    grads = {0: torch.randn(20), 1: torch.randn(20)}
    new_w, _, _, _ = kofn_update(w, torch.zeros(20), grads, [0, 1], 2, 2, lr=0.01, mu=0.0, avg_mode="n")
    ref = update_params_dist_version(w.clone(), (grads[0] + grads[1]) / 2, 0.01)
    assert torch.equal(new_w, ref)


def test_delay_specs_are_deterministic_and_respect_rank_subsets():
    from pytorch_distributed_nn_b200.parallel.delays import parse_delay_spec, summarize_ms
    f = parse_delay_spec("const:ranks=1,2,3:ms=500:scale=0.01")
    assert f(0, 1) == f(1, 5) == f(2, 9) == 5_000_000            # workers 0..2 == reference ranks 1..3
    assert f(3, 1) == 0
    u = parse_delay_spec("uniform:lo=300:hi=1200:scale=0.001")
    vals = [u(w, s) for w in range(4) for s in range(50)]
    assert all(300_000 <= v <= 1_200_000 for v in vals)
    assert vals == [parse_delay_spec("uniform:lo=300:hi=1200:scale=0.001")(w, s) for w in range(4) for s in range(50)]
    for spec in ("exp:shift=1:mean=2", "lognormal:mu=0:sigma=1", "pareto:xm=1:alpha=1.5", "exp:mean=1:p=0.25"):
        g = parse_delay_spec(spec)
        xs = [g(1, s) / 1e6 for s in range(400)]
        st = summarize_ms(xs)
        assert st["p99"] >= st["p50"] >= 0 and st["max"] >= st["p99"]
    sparse = [parse_delay_spec("exp:mean=1:p=0.25")(0, s) for s in range(400)]
    assert 0.1 < sum(1 for v in sparse if v > 0) / 400 < 0.4


def test_checkpoint_layout_is_a_plain_state_dict(tmp_path):
    from pytorch_distributed_nn_b200.models import build_model
    from pytorch_distributed_nn_b200.utils import checkpoint as ckpt
    net = build_model("LeNet")
    path, _ = ckpt.save_checkpoint(str(tmp_path), 40, net.state_dict(), momentum=torch.ones(7))
    assert os.path.basename(path) == "model_step_40"
    ckpt.save_checkpoint(str(tmp_path), 120, net.state_dict())
    assert ckpt.latest_checkpoint(str(tmp_path))[1] == 120
    raw = torch.load(path)                                           # nothing but the model's own keys
    assert list(raw.keys()) == list(net.state_dict().keys())
    fresh = build_model("LeNet")
    fresh.load_state_dict(raw)
    sd, step, mom = ckpt.load_checkpoint(path)
    assert step == 40 and torch.equal(mom, torch.ones(7))
    for k, v in net.state_dict().items():
        assert torch.equal(sd[k], v)


def test_dataset_next_batch_epochs_and_per_worker_shuffle():
    from pytorch_distributed_nn_b200 import cifar10, mnist
    from pytorch_distributed_nn_b200.datasets import Cifar10Dataset, MNISTDataset
    a = mnist.read_data_sets("does-not-exist", reshape=True, worker_id=1)
    b = mnist.read_data_sets("does-not-exist", reshape=True, worker_id=2)
    xa, ya = a.train.next_batch(64)
    xb, yb = b.train.next_batch(64)
    assert xa.shape == (64, 1, 28, 28) and xa.min() >= -0.5 and xa.max() <= 0.5
    assert not np.array_equal(ya, yb)                                # workers do not see identical batches (§2.11-6)
    n = a.train.num_examples
    for _ in range(n // 64 + 1):
        a.train.next_batch(64)
    assert a.train.epochs_completed == 1
    c = cifar10.read_data_sets(padding_size=2, reshape=True)
    x, y = Cifar10Dataset(c.train, None).next_batch(8)
    assert tuple(x.shape) == (8, 3, 32, 32) or tuple(x.shape) == (8, 3, 36, 36)
    assert x.dtype == torch.float32 and y.dtype == torch.int64 and float(x.max()) > 1.0
    assert len(MNISTDataset(a.train, None)) == n


def test_mnist_idx_reader_parses_big_endian_files(tmp_path):
    import gzip
    import struct
    from pytorch_distributed_nn_b200.utils.data import extract_data, extract_labels
    imgs = (np.arange(3 * 28 * 28) % 256).astype(np.uint8)
    with gzip.open(tmp_path / "img.gz", "wb") as f:
        f.write(struct.pack(">IIII", 2051, 3, 28, 28) + imgs.tobytes())
    with open(tmp_path / "lab", "wb") as f:
        f.write(struct.pack(">II", 2049, 3) + bytes([7, 1, 9]))
    x = extract_data(str(tmp_path / "img.gz"), 3)
    assert x.shape == (3, 28, 28, 1) and abs(float(x[0, 0, 1, 0]) - (1 - 127.5) / 255) < 1e-6
    assert extract_labels(str(tmp_path / "lab"), 3).tolist() == [7, 1, 9]


@pytest.mark.parametrize("name", ["LeNet", "ResNet18", "VGG11", "MLP", "FCDeep"])
def test_split_backward_equals_autograd_and_streams_in_reverse_order(name):
    from pytorch_distributed_nn_b200.models import build_model, input_shape
    torch.manual_seed(0)
    plain, split = build_model(name), build_model(name, split=True)
    split.load_state_dict(plain.state_dict())
    x = torch.randn(4, *input_shape(name))
    y = torch.randint(0, 10, (4,))
    torch.nn.functional.cross_entropy(plain(x), y).backward()
    out = split(x)
    torch.nn.functional.cross_entropy(out, y).backward()
    sent = []

    class Comm:
        def send_grad(self, idx, g):
            sent.append((idx, g.clone()))

    split.backward(out.grad, Comm(), [])
    params = list(plain.parameters())
    assert [i for i, _ in sent] == list(range(len(params) - 1, -1, -1))      # strictly reverse parameters() order
    assert split.fetch_init_channel_index == len(params)
    for idx, g in sent:
        assert torch.allclose(g, params[idx].grad, atol=1e-6)


def test_split_backward_signal_kill_stops_sending():
    from pytorch_distributed_nn_b200.models import build_model
    split = build_model("LeNet", split=True)
    out = split(torch.randn(2, 1, 28, 28))
    out.sum().backward()
    sent = []

    class Comm:
        calls = 0

        def send_grad(self, idx, g):
            sent.append(idx)

        def should_abort(self, step):
            Comm.calls += 1
            return Comm.calls > 3          # killed before the 4th op of the reversed walk

    reqs, killed = split.backward_signal_kill(out.grad, Comm(), [], cur_step=5)
    assert killed and sent == [7, 6, 5, 4]      # fc2 (bias, weight), fc1 (bias, weight); conv grads never leave


def test_accuracy_and_flags():
    from pytorch_distributed_nn_b200.nn_ops import accuracy
    from pytorch_distributed_nn_b200.utils.config import add_fit_args, kwargs_from_args
    out = torch.tensor([[0.1, 0.9, 0.0], [0.8, 0.1, 0.1], [0.2, 0.3, 0.5]])
    p1, p2 = accuracy(out, torch.tensor([1, 2, 2]), topk=(1, 2))
    assert abs(float(p1) - 200 / 3) < 1e-4 and abs(float(p2) - 200 / 3) < 1e-4
    args = add_fit_args(argparse.ArgumentParser()).parse_args([])
    assert (args.batch_size, args.test_batch_size, args.epochs, args.lr, args.momentum, args.seed, args.log_interval) == \
        (128, 1000, 100, 0.01, 0.5, 1, 10)
    assert (args.network, args.dataset, args.comm_type, args.num_aggregate, args.no_cuda) == ("LeNet", "MNIST", "Bcast", 5, False)
    kw = kwargs_from_args(args, master=True)
    for key in ("batch_size", "learning_rate", "max_epochs", "momentum", "network", "comm_method", "kill_threshold"):
        assert key in kw
    # workers get k and the deadline as well: on the device backend every rank derives its data path from them
    kw_w = kwargs_from_args(args, master=False)
    assert kw_w["kill_threshold"] == 5 and "interval_ms" in kw_w and "straggler_policy" in kw_w


def test_engine_layout_matches_torch_parameter_order_on_meta_device():
    from pytorch_distributed_nn_b200.engine.network import build_network
    from pytorch_distributed_nn_b200.models import build_model
    for name, count in (("LeNet", 8), ("ResNet18", 62), ("ResNet50", 161), ("VGG11", 34), ("FC", 16), ("FCDeep", 56)):
        net = build_network(name, 128, device="meta")
        ref = build_model(name)
        assert [p.name for p in net.param_list] == [k for k, _ in ref.named_parameters()]
        assert [p.torch_shape for p in net.param_list] == [tuple(v.shape) for v in ref.parameters()]
        assert len(net.param_list) == count
        spans = sorted((b.start, b.end) for b in net.buckets)
        assert spans[0][0] == 0 and spans[-1][1] == net.total
        assert all(spans[i][1] == spans[i + 1][0] for i in range(len(spans) - 1))
        assert net.buckets[0].end == net.total                      # bucket 0 = deepest layers, closes first


def test_reference_import_paths_exist():
    """The reference's module names stay importable (model_ops, nn_ops, datasets, mnist, cifar10, data_loader_ops,
    distributed_functions, data_parallel_dist, the four role modules)."""
    import importlib
    for mod in ("model_ops.lenet", "model_ops.resnet", "model_ops.resnet_split", "model_ops.vgg", "nn_ops", "datasets", "mnist", "cifar10",
                "data_loader_ops.my_data_loader", "distributed_functions.distributed_backward", "data_parallel_dist.data_parallel_dist",
                "sync_replicas_master_nn", "sync_replicas_master_nn_normal", "distributed_worker", "distributed_worker_normal",
                "distributed_nn", "single_machine", "evaluator", "comm_test"):
        importlib.import_module("pytorch_distributed_nn_b200." + mod)
    from pytorch_distributed_nn_b200.model_ops.resnet_split import ResNetSplit18, ResNetSplit50  # noqa: F401
    from pytorch_distributed_nn_b200.sync_replicas_master_nn import GradientAccumulator, SyncReplicasMaster_NN  # noqa: F401
    from pytorch_distributed_nn_b200.distributed_worker import DistributedWorker, ModelBuffer  # noqa: F401
    from pytorch_distributed_nn_b200.distributed_functions.distributed_backward import backward
    import torch
    x = torch.ones(3, requires_grad=True)
    backward([(x * 2).sum()])
    assert torch.equal(x.grad, torch.full((3,), 2.0))


def test_ddp_comparator_single_process_is_identity():
    import torch
    from pytorch_distributed_nn_b200.data_parallel_dist import DistributedDataParallel
    from pytorch_distributed_nn_b200.models import build_model
    net = DistributedDataParallel(build_model("LeNet"))
    out = net(torch.randn(2, 1, 28, 28))
    out.sum().backward()
    net.finish_backward()
    assert len(net.buckets) >= 1 and all(p.grad is not None for p in net.module.parameters())


def test_cifar10_pickle_reader_keeps_channel_planes(tmp_path):
    """The python pickles store each image as 3 planes of 1024 bytes (CHW). The reference reshapes an NHWC array to NCHW
    without transposing (cifar10/cifar10.py:67-70); here plane c of image i must come back as [i, c]."""
    import pickle
    import numpy as np
    from pytorch_distributed_nn_b200.utils.data import read_cifar10
    root = tmp_path / "cifar-10-batches-py"
    root.mkdir()
    rng = np.random.RandomState(0)

    def dump(name, n, base):
        data = rng.randint(0, 256, size=(n, 3072)).astype(np.uint8)
        data[:, :1024] = base          # red plane constant: easy to recognise after loading
        with open(root / name, "wb") as f:
            pickle.dump({"data": data, "labels": [int(i % 10) for i in range(n)]}, f)
        return data
    for i in range(1, 6):
        dump(f"data_batch_{i}", 4, 10 * i)
    test = dump("test_batch", 3, 77)
    ds = read_cifar10(str(tmp_path))
    assert ds.train.images.shape == (20, 3, 32, 32) and ds.validation.images.shape == (3, 3, 32, 32)
    assert float(ds.validation.images[:, 0].min()) == 77.0 and float(ds.validation.images[:, 0].max()) == 77.0
    # the reader shuffles: every loaded image must be one of the pickled rows, plane for plane
    for img in ds.validation.images:
        assert any((img.reshape(-1) == row.astype(np.float32)).all() for row in test)
    nhwc = read_cifar10(str(tmp_path), reshape=False)
    assert nhwc.validation.images.shape == (3, 32, 32, 3) and float(nhwc.validation.images[..., 0].max()) == 77.0
    padded = read_cifar10(str(tmp_path), padding_size=2)
    assert padded.train.images.shape == (20, 3, 36, 36) and float(padded.train.images[:, :, 0, :].max()) == 0.0


def test_single_machine_cpu_trains_lenet_and_writes_a_state_dict(tmp_path):
    """pytorch_code/single_machine.py counterpart on the CPU path (torch modules + optim.SGD, nn_ops/__init__.py:46-85)."""
    from pytorch_distributed_nn_b200 import single_machine
    from pytorch_distributed_nn_b200.utils.checkpoint import latest_checkpoint, load_checkpoint
    hist = single_machine.main(["--network", "LeNet", "--dataset", "MNIST", "--synthetic", "--no-cuda", "--batch-size", "32",
                                "--max-steps", "30", "--lr", "0.05", "--log-interval", "1", "--train-dir", str(tmp_path)])
    assert [h["step"] for h in hist] == list(range(1, 31))
    assert all(h["loss"] == h["loss"] and h["loss"] < 10 for h in hist)            # finite, sane
    path, step = latest_checkpoint(str(tmp_path))
    sd = load_checkpoint(path)[0]
    assert step == 30 and "conv1.weight" in sd and sd["fc2.bias"].shape == (10,)


def test_polling_evaluator_restores_the_newest_checkpoint(tmp_path, capsys):
    """distributed_TF/src/nn_eval.py counterpart: the stand-alone evaluator restores whatever checkpoint is newest in the
    directory and prints `Num examples ... Precision @ 1 ... Loss ...` (nn_eval.py:49-111)."""
    import torch
    from pytorch_distributed_nn_b200.evaluator import evaluate_checkpoints
    from pytorch_distributed_nn_b200.models import build_model
    from pytorch_distributed_nn_b200.utils.checkpoint import save_checkpoint
    from pytorch_distributed_nn_b200.utils.data import read_mnist
    torch.manual_seed(0)
    net = build_model("LeNet")
    save_checkpoint(str(tmp_path), 10, net.state_dict())
    with torch.no_grad():
        for p in net.parameters():
            p.mul_(0.5)
    save_checkpoint(str(tmp_path), 20, net.state_dict())
    data = read_mnist(synthetic=True).validation
    res = evaluate_checkpoints(str(tmp_path), "LeNet", data, run_once=True)
    assert len(res) == 1 and res[0][0] == 20 and 0.0 <= res[0][2] <= 1.0 and res[0][1] == res[0][1]
    assert "Precision @ 1" in capsys.readouterr().out
    # nothing new: a second poll evaluates nothing
    assert evaluate_checkpoints(str(tmp_path), "LeNet", data, eval_interval_secs=0.01, max_polls=2)[0][0] == 20
    # the same thing as a process with nn_eval.py's flags
    import json
    import subprocess
    import sys
    out = tmp_path / "eval.jsonl"
    r = subprocess.run([sys.executable, "-m", "pytorch_distributed_nn_b200.evaluator", "--checkpoint-dir", str(tmp_path), "--network", "LeNet",
                        "--dataset", "MNIST", "--synthetic", "--run-once", "--out", str(out)], capture_output=True, text=True, timeout=300,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Precision @ 1" in r.stdout and json.loads(out.read_text().splitlines()[0])["step"] == 20


def test_philox_known_answers_and_device_delay_replica():
    """The host replica of the device-side delay draw: Philox4x32-10 known-answer vectors (Random123 kat_vectors) and the
    properties a study relies on (deterministic in seed / worker / step / layer, rank subsets in the reference's spelling)."""
    from pytorch_distributed_nn_b200.parallel.delays import philox4x32_10
    assert philox4x32_10((0, 0, 0, 0), (0, 0)) == (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)
    assert philox4x32_10((0xffffffff,) * 4, (0xffffffff, 0xffffffff)) == (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)
    assert philox4x32_10((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0)) == \
        (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)
    pytest.importorskip("ctypes")
    try:
        from pytorch_distributed_nn_b200.parallel.delays import DeviceDelay
        dd = DeviceDelay("uniform:lo=300:hi=1200:scale=0.005:seed=11")
    except Exception as ex:          # the ctypes struct needs the built library
        pytest.skip(f"library not built: {ex}")
    a = [dd.sample_ns(w, s) for w in range(4) for s in range(1, 40)]
    assert a == [DeviceDelay("uniform:lo=300:hi=1200:scale=0.005:seed=11").sample_ns(w, s) for w in range(4) for s in range(1, 40)]
    assert all(1_500_000 <= v <= 6_000_000 for v in a) and len(set(a)) > 100
    assert a != [DeviceDelay("uniform:lo=300:hi=1200:scale=0.005:seed=12").sample_ns(w, s) for w in range(4) for s in range(1, 40)]
    sub = DeviceDelay("const:ranks=1,2,3:ms=500:scale=0.01")
    assert [sub.sample_ns(w, 7) for w in range(5)] == [5_000_000] * 3 + [0, 0]
    per_layer = DeviceDelay("exp:mean=1:per=layer:seed=2")
    assert per_layer.per_layer and len({per_layer.sample_ns(0, 3, layer) for layer in range(8)}) == 8
    sparse = DeviceDelay("exp:mean=1:p=0.25:seed=5")
    hits = sum(1 for s in range(1, 801) if sparse.sample_ns(1, s) > 0)
    assert 150 < hits < 250
    # the tails the straggler study sweeps: moments of the draws match the named distributions
    n = 4000
    ex = [DeviceDelay("exp:shift=0.5:mean=2:seed=3").sample_ns(0, s) / 1e6 for s in range(1, n + 1)]
    assert min(ex) >= 0.5 and abs(sum(ex) / n - 2.5) < 0.15
    ln = sorted(DeviceDelay("lognormal:mu=0.7:sigma=0.5:seed=4").sample_ns(2, s) / 1e6 for s in range(1, n + 1))
    assert abs(ln[n // 2] - 2.0138) < 0.12                                  # median = e^mu
    pa = sorted(DeviceDelay("pareto:xm=1.5:alpha=2:seed=6").sample_ns(1, s) / 1e6 for s in range(1, n + 1))
    assert pa[0] >= 1.5 and abs(pa[n // 2] - 1.5 * 2 ** 0.5) < 0.1           # median = xm * 2^(1/alpha)
