"""Checkpoint layout (north-star addition; the reference snapshot has no torch.save anywhere).

The master writes ``{train_dir}/model_step_{cur_step}``: a plain ``torch.save`` of the model's ``state_dict()`` --
the model's own keys, fp32, nothing else -- every ``--eval-freq`` steps. Optimizer momentum goes to the sidecar
``model_step_{cur_step}.optim`` so that the main file stays loadable with ``model.load_state_dict(torch.load(p))``.
``latest_checkpoint`` parses the step out of the file name (as the TF evaluator does, distributed_TF/src/nn_eval.py:84).
"""
from __future__ import annotations

import os
import re
import threading
from typing import Optional, Tuple

import torch

_PAT = re.compile(r"model_step_(\d+)$")


def checkpoint_path(train_dir: str, step: int) -> str:
    return os.path.join(train_dir, f"model_step_{step}")


def save_checkpoint(train_dir: str, step: int, state_dict, momentum: Optional[torch.Tensor] = None, background: bool = False):
    os.makedirs(train_dir, exist_ok=True)
    path = checkpoint_path(train_dir, step)
    cpu_sd = {k: v.detach().to("cpu", copy=True) for k, v in state_dict.items()}
    mom = None if momentum is None else momentum.detach().to("cpu", copy=True)

    def write():
        # sidecar first, each through tmp + rename: a reader that sees model_step_N also sees its complete .optim
        if mom is not None:
            tmp = path + ".optim.tmp"
            torch.save({"step": step, "momentum": mom}, tmp)
            os.replace(tmp, path + ".optim")
        tmp = path + ".tmp"
        torch.save(cpu_sd, tmp)
        os.replace(tmp, path)

    if background:
        t = threading.Thread(target=write, daemon=True)
        t.start()
        return path, t
    write()
    return path, None


def latest_checkpoint(train_dir: str) -> Optional[Tuple[str, int]]:
    if not train_dir or not os.path.isdir(train_dir):
        return None
    best = None
    for name in os.listdir(train_dir):
        m = _PAT.match(name)
        if m and (best is None or int(m.group(1)) > best[1]):
            best = (os.path.join(train_dir, name), int(m.group(1)))
    return best


def load_checkpoint(path: str):
    """Returns (state_dict, step, momentum or None)."""
    sd = torch.load(path, map_location="cpu")
    m = _PAT.search(path)
    step = int(m.group(1)) if m else 0
    mom = None
    if os.path.exists(path + ".optim"):
        mom = torch.load(path + ".optim", map_location="cpu").get("momentum")
    return sd, step, mom
