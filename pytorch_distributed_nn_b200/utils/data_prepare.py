"""Prepare datasets once so that N ranks do not race (reference pytorch_code/data_prepare.py:1-17). There is no
network here: this verifies that the files exist under the given directories and otherwise reports that the
synthetic generators will be used."""
import sys

from .data import read_cifar10, read_mnist


def main(mnist_dir="MNIST_data", cifar_dir="cifar10_data"):
    m = read_mnist(mnist_dir)
    c = read_cifar10(cifar_dir)
    print(f"MNIST train={m.train.num_examples} test={m.validation.num_examples}; "
          f"CIFAR-10 train={c.train.num_examples} test={c.validation.num_examples}")


if __name__ == "__main__":
    main(*sys.argv[1:])
