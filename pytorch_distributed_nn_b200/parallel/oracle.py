"""Pure-torch oracle of the parameter-server step semantics (what every backend must compute).

Restates the reference's master loop (pytorch_code/sync_replicas_master_nn.py:148-228) with the fixes listed in
SURVEY §2.11: a worker contributes iff its WHOLE gradient set arrived among the first k for THIS step (stale or
partial sets are dropped, as the C++ master does by step tag, MPI_code/src/distributed/sync_replicas_master_nn.h:85);
the sum is scaled by 1/k (``avg_mode='k'``, north star), 1/n (``'n'``, the python reference) or 1/received
(``'received'``, the C++ reference); the update is torch momentum SGD (no dampening, no nesterov, no weight decay):
m <- mu*m + g ; w <- w - lr*m (mu = 0 gives the reference's plain ``param -= lr*grad``).
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch


def select_first_k(arrival_order: Sequence[int], k: int, n_workers: int) -> Tuple[List[int], List[int]]:
    """First k distinct workers of ``arrival_order`` (complete sets only) -> (selected sorted, killed sorted)."""
    seen: List[int] = []
    for w in arrival_order:
        if w not in seen:
            seen.append(w)
        if len(seen) == k:
            break
    selected = sorted(seen)
    killed = sorted(set(range(n_workers)) - set(selected))
    return selected, killed


def aggregate(grads_by_worker: Dict[int, torch.Tensor], selected: Sequence[int], k: int, n_workers: int,
              avg_mode: str = "k") -> torch.Tensor:
    """Sum of the selected workers' gradients in rank order, scaled per ``avg_mode``."""
    acc = None
    for w in sorted(selected):
        g = grads_by_worker[w].to(torch.float32)
        acc = g.clone() if acc is None else acc + g
    if avg_mode == "n":
        scale = 1.0 / n_workers
    else:                      # 'k' and 'received' coincide: only complete sets are ever counted
        scale = 1.0 / max(1, len(selected))
    return acc * scale


def kofn_update(w: torch.Tensor, m: torch.Tensor, grads_by_worker: Dict[int, torch.Tensor], arrival_order: Sequence[int],
                k: int, n_workers: int, lr: float, mu: float = 0.0, avg_mode: str = "k"):
    """One master step. Returns (new_w, new_m, selected, killed)."""
    selected, killed = select_first_k(arrival_order, k, n_workers)
    g = aggregate(grads_by_worker, selected, k, n_workers, avg_mode)
    new_m = mu * m + g
    new_w = w - lr * new_m
    return new_w, new_m, selected, killed
