"""Client shards as index lists into the (device-resident) training set.

The reference copies rows into per-rank arrays; here a shard is an ``int32`` index vector,
so overlapping shards (the custom script's behaviour), contiguous shards and label-skewed
shards all feed the same kernels without duplicating the data.

* ``contiguous``  — reference: FL_SkLearn_MLPClassifier_Limitation.py:17-22 /
  hyperparameters_tuning.py:17-22 (chunk = max(1, n // size); last client takes the tail).
* ``per_client_shuffle`` — reference: FL_Custom…Multiple_Rounds.py:48-61: every client
  permutes the *whole* train set independently and then takes its rank's slice, so shards
  are random subsets that may overlap.  The reference is unseeded; here client ``k`` uses
  ``RandomState(seed + k)`` so runs are reproducible.
* ``dirichlet`` — label-skew non-IID shards (BASELINE config 5; no reference counterpart).
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np


def _bounds(n: int, size: int, rank: int):
    chunk = max(1, n // size)
    start = rank * chunk
    end = start + chunk if rank != size - 1 else n
    return min(start, n), min(end, n)


def contiguous(n: int, size: int) -> List[np.ndarray]:
    out = []
    for r in range(size):
        s, e = _bounds(n, size, r)
        out.append(np.arange(s, e, dtype=np.int32))
    return out


def per_client_shuffle(n: int, size: int, seed: Optional[int] = 0) -> List[np.ndarray]:
    out = []
    for r in range(size):
        rs = np.random.RandomState(None if seed is None else seed + r)
        perm = rs.permutation(n)
        s, e = _bounds(n, size, r)
        out.append(perm[s:e].astype(np.int32))
    return out


def fixed_size(n: int, size: int, rows_per_client: int, seed: int = 0) -> List[np.ndarray]:
    """``size`` clients with exactly ``rows_per_client`` rows each: consecutive blocks of one
    global permutation, wrapping around when ``size * rows_per_client > n`` (weak-scaling
    benchmarks: per-client work is fixed while the number of clients grows)."""
    perm = np.random.RandomState(seed).permutation(n)
    out = []
    for r in range(size):
        idx = (np.arange(rows_per_client, dtype=np.int64) + r * rows_per_client) % n
        out.append(perm[idx].astype(np.int32))
    return out


def dirichlet(y: np.ndarray, size: int, alpha: float = 0.5, seed: int = 0, min_rows: int = 8) -> List[np.ndarray]:
    """Label-skewed partition: for each class, split its rows among clients with
    Dirichlet(alpha) proportions.  Every client is topped up to ``min_rows`` rows."""
    rng = np.random.default_rng(seed)
    y = np.asarray(y)
    buckets: List[list] = [[] for _ in range(size)]
    for c in np.unique(y):
        idx = np.flatnonzero(y == c)
        rng.shuffle(idx)
        p = rng.dirichlet(np.full(size, alpha))
        cuts = (np.cumsum(p)[:-1] * len(idx)).astype(np.int64)
        for k, part in enumerate(np.split(idx, cuts)):
            buckets[k].extend(part.tolist())
    out = []
    for k in range(size):
        b = np.asarray(buckets[k], dtype=np.int64)
        if len(b) < min_rows:
            b = np.concatenate([b, rng.integers(0, len(y), min_rows - len(b))])
        rng.shuffle(b)
        out.append(b.astype(np.int32))
    return out


def make_shards(kind: str, n: int, size: int, y: Optional[np.ndarray] = None, seed: int = 0,
                alpha: float = 0.5, rows_per_client: Optional[int] = None) -> List[np.ndarray]:
    if kind == "contiguous":
        return contiguous(n, size)
    if kind in ("per_client_shuffle", "per_rank_shuffle"):
        return per_client_shuffle(n, size, seed)
    if kind == "fixed":
        assert rows_per_client is not None
        return fixed_size(n, size, rows_per_client, seed)
    if kind in ("dirichlet", "non_iid"):
        assert y is not None
        return dirichlet(y, size, alpha, seed)
    raise ValueError(f"unknown partition kind {kind!r}")
