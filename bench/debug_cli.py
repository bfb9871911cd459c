import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("PDNN_PS_TIMEOUT_S", "3")
import torch
from pytorch_distributed_nn_b200 import distributed_nn
d = tempfile.mkdtemp()
net = sys.argv[1] if len(sys.argv) > 1 else "ResNet18"
ds = "MNIST" if net == "LeNet" else "Cifar10"
try:
    print(distributed_nn.main(["--network", net, "--dataset", ds, "--synthetic", "--batch-size", "64", "--backend", "b200", "--lr", "0.05",
                               "--train-dir", d, "--eval-freq", "4", "--log-interval", "4", "--max-steps", "8"]))
except Exception as e:
    print("FAILED", e)
