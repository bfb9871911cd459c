"""GPU: the public entry points on one B200 (master + colocated worker): distributed_nn, checkpoint layout, resume,
single_machine."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_distributed_nn_single_gpu_checkpoint_and_resume(tmp_path):
    from pytorch_distributed_nn_b200 import distributed_nn
    from pytorch_distributed_nn_b200.models import build_model
    from pytorch_distributed_nn_b200.utils import checkpoint as ckpt
    d = str(tmp_path)
    common = ["--network", "ResNet18", "--dataset", "Cifar10", "--synthetic", "--batch-size", "64", "--backend", "b200", "--lr", "0.05",
              "--train-dir", d, "--eval-freq", "4", "--log-interval", "4"]
    summary = distributed_nn.main(common + ["--max-steps", "8"])
    assert summary["steps"] >= 2
    path, step = ckpt.latest_checkpoint(d)
    assert step == 8 and os.path.exists(os.path.join(d, "model_step_4"))
    sd = torch.load(path)
    ref = build_model("ResNet18")
    assert list(sd.keys()) == list(ref.state_dict().keys())          # plain torch state_dict, torch (OIHW) layout
    ref.load_state_dict(sd)
    assert os.path.exists(path + ".optim")
    # resume continues the global step numbering from the file name
    distributed_nn.main(common + ["--max-steps", "4", "--resume", path])
    assert ckpt.latest_checkpoint(d)[1] == 12


def test_single_machine_gpu_trains_lenet():
    from pytorch_distributed_nn_b200 import single_machine
    hist = single_machine.main(["--network", "LeNet", "--dataset", "MNIST", "--synthetic", "--batch-size", "64", "--max-steps", "40",
                                "--lr", "0.05", "--log-interval", "10"])
    assert len(hist) >= 4 and hist[-1]["loss"] < hist[0]["loss"] + 0.05 and all(h["loss"] == h["loss"] for h in hist)
