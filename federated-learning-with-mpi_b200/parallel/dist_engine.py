"""Process-per-GPU federation: every rank hosts the clients placed on its GPU.

``make_local_engine`` is the single-process case (all clients on this GPU).  The multi-GPU
engine (cross-GPU aggregation inside one kernel over NVLink peer memory) lives in
``DistFedEngine``.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence

import numpy as np
import torch

from ..fed.engine import FedEngine
from ..fed.problem import FedProblem


def init_distributed(backend: Optional[str] = None) -> None:
    import torch.distributed as dist
    if dist.is_initialized():
        return
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group(backend)


def _compact(data, rows: Sequence[np.ndarray]):
    """Gather the rows the clients use into one dense matrix; shards become index ranges."""
    allr = np.concatenate(rows)
    X, y = data.X_train[allr], data.y_train[allr]
    shards, o = [], 0
    for r in rows:
        shards.append(np.arange(o, o + len(r), dtype=np.int32))
        o += len(r)
    return X, y, shards


def make_local_engine(data, rows: Sequence[np.ndarray], max_rounds: int = 300, precision: Optional[str] = None,
                      patience: int = 0, device: Optional[str] = None) -> FedEngine:
    X, y, shards = _compact(data, rows)
    prob = FedProblem.custom(data.n_features, data.n_classes, shards, patience=patience)
    return FedEngine([prob], X, y, data.X_test, data.y_test, device=device, max_rounds=max_rounds, precision=precision)
