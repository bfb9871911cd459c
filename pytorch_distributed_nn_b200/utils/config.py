"""Flag surface. The reference's argparse flags are kept verbatim (pytorch_code/distributed_nn.py:42-65,
single_machine.py:34-53); additions are grouped separately and marked as such."""
from __future__ import annotations

import argparse


def add_fit_args(parser: argparse.ArgumentParser, distributed: bool = True) -> argparse.ArgumentParser:
    g = parser.add_argument_group("reference flags")
    g.add_argument("--batch-size", type=int, default=128, metavar="N", help="per-worker training batch size (default: 128)")
    g.add_argument("--test-batch-size", type=int, default=1000, metavar="N", help="evaluation batch size (default: 1000)")
    g.add_argument("--epochs", type=int, default=100, metavar="N", help="number of epochs (default: 100)")
    g.add_argument("--lr", type=float, default=0.01, metavar="LR", help="learning rate (default: 0.01)")
    g.add_argument("--momentum", type=float, default=0.5, metavar="M", help="SGD momentum (default: 0.5)")
    g.add_argument("--no-cuda", action="store_true", default=False, help="force the CPU (gloo) backend")
    g.add_argument("--seed", type=int, default=1, metavar="S", help="random seed (default: 1)")
    g.add_argument("--log-interval", type=int, default=10, metavar="N", help="steps between log lines")
    g.add_argument("--network", type=str, default="LeNet", metavar="N",
                   help="LeNet | ResNet18 | ResNet34 | ResNet50 | ResNet101 | ResNet152 | VGG11 | VGG13 | VGG16 | VGG19 | FC | FCDeep")
    g.add_argument("--dataset", type=str, default="MNIST", metavar="N", help="MNIST | Cifar10")
    if distributed:
        g.add_argument("--comm-type", type=str, default="Bcast", metavar="N", help="weight distribution: Bcast | Async")
        g.add_argument("--num-aggregate", type=int, default=5, metavar="N",
                       help="k: number of gradient sets the master aggregates per step (clamped to #workers)")
    a = parser.add_argument_group("additions of this framework")
    a.add_argument("--backend", default="auto", choices=["auto", "b200", "gloo", "nccl-ref"],
                   help="b200 = device-resident PS over NVLink peer memory; gloo = CPU message passing; auto picks b200 on GPU")
    a.add_argument("--max-steps", type=int, default=0, help="stop after this many master steps (0 = reference cap 20000)")
    a.add_argument("--train-dir", default="", help="master writes model_step_<n> state_dict checkpoints here")
    a.add_argument("--eval-freq", type=int, default=0, help="checkpoint / evaluate every N steps (0 = off)")
    a.add_argument("--resume", default="", help="path of a model_step_<n> checkpoint to continue from")
    a.add_argument("--synthetic", action="store_true", default=False, help="synthetic data of the dataset's shape")
    a.add_argument("--data-dir", default="", help="directory holding the dataset files")
    if distributed:
        a.add_argument("--straggler-policy", default="kofn", choices=["all", "kofn", "interval"],
                       help="all = wait for every worker (the *_normal roles); kofn = first k + kill; interval = deadline")
        a.add_argument("--interval-ms", type=float, default=0.0, help="deadline of the interval policy")
        a.add_argument("--avg-mode", default="k", choices=["k", "n", "received"], help="gradient normalisation (n = python reference)")
        a.add_argument("--inject-delay", default="", help="straggler injection spec, see parallel/delays.py")
        a.add_argument("--grad-dtype", default="bf16", choices=["bf16", "fp32", "fp64"], help="gradient wire dtype")
        a.add_argument("--colocate-worker", action="store_true", default=False, help="GPU 0 hosts master AND a worker (b200 backend)")
        a.add_argument("--evaluator", action="store_true", default=False, help="dedicate the last rank to evaluation (C++/TF reference role)")
        a.add_argument("--n-buckets", type=int, default=8, help="gradient buckets per step (b200 backend)")
        a.add_argument("--no-cuda-graph", dest="cuda_graph", action="store_false", default=True,
                       help="b200 backend: enqueue every kernel of every step from the host instead of replaying a CUDA graph")
    return parser


def kwargs_from_args(args, master: bool):
    """The reference's constructor kwargs contract (distributed_nn.py:85-89) plus our additions."""
    kw = {
        "batch_size": args.batch_size, "learning_rate": args.lr, "max_epochs": args.epochs, "momentum": args.momentum,
        "network": args.network, "comm_method": getattr(args, "comm_type", "Bcast"),
    }
    # the reference hands k to the master only (distributed_nn.py:85-89); on the device backend every rank derives its
    # data path (NVLS gather vs. pushed slots) from k and the deadline, so workers receive them too
    kw["kill_threshold"] = getattr(args, "num_aggregate", None)
    for extra in ("max_steps", "train_dir", "eval_freq", "resume", "straggler_policy", "interval_ms", "avg_mode", "inject_delay",
                  "grad_dtype", "colocate_worker", "n_buckets", "log_interval", "seed", "evaluator", "cuda_graph", "test_batch_size"):
        if hasattr(args, extra):
            kw[extra] = getattr(args, extra)
    return kw
