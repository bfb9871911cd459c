"""Single-machine entry point (reference pytorch_code/single_machine.py:184-205): same flags minus the
distributed ones; runs ``NN_Trainer`` on one GPU (B200 engine + fused momentum-SGD kernel) or on the CPU."""
from __future__ import annotations

import argparse

import torch

from .nn_ops import NN_Trainer
from .utils.config import add_fit_args


class _Loader:
    """Adapts a ``next_batch(batch_size)`` dataset to the ``next_batch()`` / ``len()`` loader contract."""

    def __init__(self, dataset, batch_size):
        self.dataset, self.batch_size = dataset, batch_size

    def __len__(self):
        return max(1, len(self.dataset) // self.batch_size)

    def next_batch(self):
        return self.dataset.next_batch(self.batch_size)


def main(argv=None):
    parser = add_fit_args(argparse.ArgumentParser(description="single-machine trainer"), distributed=False)
    args = parser.parse_args(argv)
    torch.manual_seed(args.seed)
    from .distributed_nn import load_dataset
    train_set, _ = load_dataset(args, 0, 1)
    device = "cpu" if (args.no_cuda or not torch.cuda.is_available()) else "cuda"
    kwargs = {"batch_size": args.batch_size, "learning_rate": args.lr, "max_epochs": args.epochs, "momentum": args.momentum,
              "network": args.network, "device": device, "max_steps": args.max_steps, "log_interval": args.log_interval}
    trainer = NN_Trainer(**kwargs)
    trainer.build_model()
    hist = trainer.train(_Loader(train_set, args.batch_size))
    if args.train_dir:
        from .utils.checkpoint import save_checkpoint
        save_checkpoint(args.train_dir, hist[-1]["step"] if hist else 0, trainer.state_dict())   # named after the last step taken
    return hist


if __name__ == "__main__":
    main()
