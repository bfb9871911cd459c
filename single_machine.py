#!/usr/bin/env python3
"""``python single_machine.py ...`` -- same spelling as the reference's entry script."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from pytorch_distributed_nn_b200.single_machine import main  # noqa: E402

if __name__ == "__main__":
    main()
