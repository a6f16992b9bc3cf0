"""Tabular data pipeline: CSV -> label-encode -> scale -> train/test split.

Restates, with NumPy only, what every reference driver does inline before training
(reference: FL_SkLearn_MLPClassifier_Limitation.py:163-197,
FL_CustomMLPCLassifierImplementation_Multiple_Rounds.py:216-246,
hyperparameters_tuning.py:143-177):

* every string column (label included) is replaced by its rank in the sorted set of
  distinct values (what ``sklearn.preprocessing.LabelEncoder`` computes);
* the feature matrix is scaled with statistics of *all* rows, before the split
  (``StandardScaler()`` in the custom script, ``StandardScaler(with_mean=False)`` in the
  sklearn scripts);
* an 80/20 split driven by ``RandomState(42).permutation(n)``: the first ``ceil(0.2 n)``
  permuted rows are the test set, the rest the train set (this is what
  ``train_test_split(test_size=0.2, random_state=42)`` does for unstratified input).

Unlike the reference, which repeats this on every rank and then broadcasts the rank-0
result, this runs once per process and the result is uploaded once and stays resident on
the device for the whole run.

The income table ships re-encoded column-wise in ``income_columns.npz`` (lossless; see
``tools/pack_income_csv.py``); ``load_table`` also reads any CSV with a header row.
"""
from __future__ import annotations

import csv
import math
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
BUNDLED_INCOME = os.path.join(_HERE, "income_columns.npz")
INCOME_LABEL = "income"


@dataclass
class Table:
    """A column-oriented table: ``columns[name]`` is an int64 array or an array of str."""

    names: List[str]
    columns: Dict[str, np.ndarray]

    @property
    def n_rows(self) -> int:
        return len(self.columns[self.names[0]])


def _read_csv(path: str) -> Table:
    with open(path, newline="") as f:
        reader = csv.reader(f)
        names = next(reader)
        raw = list(zip(*reader))
    columns: Dict[str, np.ndarray] = {}
    for name, values in zip(names, raw):
        try:
            columns[name] = np.asarray(values, dtype=np.int64)
        except ValueError:
            try:
                columns[name] = np.asarray(values, dtype=np.float64)
            except ValueError:
                columns[name] = np.asarray(values, dtype=str)
    return Table(list(names), columns)


def _read_packed(path: str) -> Table:
    with np.load(path, allow_pickle=False) as z:
        names = [str(n) for n in z["__names__"]]
        columns: Dict[str, np.ndarray] = {}
        for i, name in enumerate(names):
            if f"codes_{i}" in z.files:
                columns[name] = z[f"vocab_{i}"][z[f"codes_{i}"]]
            else:
                columns[name] = z[f"num_{i}"]
    return Table(names, columns)


def pack_table(table: Table, path: str) -> None:
    """Write ``table`` column-wise (numeric columns verbatim, string columns as
    vocabulary + codes).  Inverse of ``_read_packed``."""
    out = {"__names__": np.asarray(table.names, dtype=str)}
    for i, name in enumerate(table.names):
        col = table.columns[name]
        if col.dtype.kind in "US":
            vocab, codes = np.unique(col, return_inverse=True)
            out[f"vocab_{i}"] = vocab
            out[f"codes_{i}"] = codes.astype(np.uint8 if len(vocab) < 256 else np.int32)
        else:
            out[f"num_{i}"] = col
    np.savez_compressed(path, **out)


def load_table(path: Optional[str] = None) -> Table:
    """Load a table.  ``None`` -> the bundled income data (or ``$FEDB200_DATA``)."""
    path = path or os.environ.get("FEDB200_DATA") or BUNDLED_INCOME
    if path.endswith(".npz"):
        return _read_packed(path)
    return _read_csv(path)


def label_encode(table: Table) -> Tuple[Dict[str, np.ndarray], Dict[str, np.ndarray]]:
    """Sorted-unique rank encoding of every string column.  Returns (encoded columns as
    int64/float64 arrays, {column: vocabulary})."""
    encoded, vocab = {}, {}
    for name in table.names:
        col = table.columns[name]
        if col.dtype.kind in "US":
            v, inv = np.unique(col, return_inverse=True)
            encoded[name] = inv.astype(np.int64)
            vocab[name] = v
        else:
            encoded[name] = col
    return encoded, vocab


@dataclass
class Prepared:
    """Output of :func:`prepare` — everything a federation needs, float64 like the
    reference's arrays; device upload converts to padded fp32."""

    X_train: np.ndarray
    y_train: np.ndarray
    X_test: np.ndarray
    y_test: np.ndarray
    feature_names: List[str]
    classes: np.ndarray
    scale_mean: np.ndarray
    scale_std: np.ndarray
    with_mean: bool
    vocab: Dict[str, np.ndarray] = field(default_factory=dict)

    @property
    def n_features(self) -> int:
        return self.X_train.shape[1]

    @property
    def n_classes(self) -> int:
        return int(len(self.classes))


def split_indices(n: int, test_size: float = 0.2, seed: int = 42) -> Tuple[np.ndarray, np.ndarray]:
    """(train_idx, test_idx) of an unstratified shuffle split with a legacy-MT19937 seed."""
    n_test = int(math.ceil(test_size * n))
    perm = np.random.RandomState(seed).permutation(n)
    return perm[n_test:], perm[:n_test]


def prepare(
    table: Optional[Table] = None,
    label: str = INCOME_LABEL,
    with_mean: bool = True,
    test_size: float = 0.2,
    split_seed: int = 42,
) -> Prepared:
    """Encode, scale on all rows, split.  ``with_mean=True`` reproduces the custom-MLP
    driver (…Multiple_Rounds.py:235), ``False`` the sklearn drivers (…Limitation.py:184)."""
    table = table if table is not None else load_table()
    if label not in table.names:
        raise KeyError(f"'{label}' not found in dataset columns. Available columns: {table.names}")
    enc, vocab = label_encode(table)
    feats = [n for n in table.names if n != label]
    X = np.stack([enc[n].astype(np.float64) for n in feats], axis=1)
    y_raw = enc[label]
    classes, y = np.unique(y_raw, return_inverse=True)
    y = y.astype(np.int64)
    mean = X.mean(axis=0)
    std = X.std(axis=0)  # ddof=0
    std_safe = np.where(std == 0.0, 1.0, std)
    Xs = (X - mean) / std_safe if with_mean else X / std_safe
    tr, te = split_indices(len(X), test_size, split_seed)
    return Prepared(
        X_train=np.ascontiguousarray(Xs[tr]),
        y_train=np.ascontiguousarray(y[tr]),
        X_test=np.ascontiguousarray(Xs[te]),
        y_test=np.ascontiguousarray(y[te]),
        feature_names=feats,
        classes=classes,
        scale_mean=mean,
        scale_std=std,
        with_mean=with_mean,
        vocab=vocab,
    )


def synthetic_income_like(
    n_rows: int,
    n_features: int = 14,
    seed: int = 0,
    n_classes: int = 2,
    with_mean: bool = True,
    test_size: float = 0.2,
    teacher_hidden: int = 32,
) -> Prepared:
    """Income-*shaped* synthetic data for the scaled benchmarks: a mix of heavy-tailed
    integer-like and small-cardinality categorical columns, labelled by a fixed random
    two-layer teacher so that an MLP can actually learn it; balanced by thresholding the
    teacher's score at its median (the bundled table is 5000/5000 as well)."""
    rng = np.random.default_rng(seed)
    cols = []
    for j in range(n_features):
        kind = j % 3
        if kind == 0:  # age / hours-like
            cols.append(np.clip(rng.normal(40.0, 12.0, n_rows).round(), 17, 99))
        elif kind == 1:  # categorical codes
            card = int(rng.integers(2, 41))
            p = rng.dirichlet(np.full(card, 0.6))
            cols.append(rng.choice(card, size=n_rows, p=p).astype(np.float64))
        else:  # capital-gain-like: mostly zero, heavy tail
            z = rng.random(n_rows) < 0.1
            cols.append(np.where(z, rng.lognormal(8.0, 1.2, n_rows).round(), 0.0))
    X = np.stack(cols, axis=1)
    mean, std = X.mean(0), X.std(0)
    std_safe = np.where(std == 0, 1.0, std)
    Z = (X - mean) / std_safe
    W1 = rng.normal(0, 1.0 / np.sqrt(n_features), (n_features, teacher_hidden))
    W2 = rng.normal(0, 1.0 / np.sqrt(teacher_hidden), (teacher_hidden, n_classes))
    score = np.maximum(Z @ W1, 0.0) @ W2 + 0.3 * rng.normal(size=(n_rows, n_classes))
    if n_classes == 2:
        d = score[:, 1] - score[:, 0]
        y = (d > np.median(d)).astype(np.int64)
    else:
        y = score.argmax(1).astype(np.int64)
    Xs = Z if with_mean else X / std_safe
    tr, te = split_indices(n_rows, test_size, 42)
    return Prepared(
        X_train=np.ascontiguousarray(Xs[tr]),
        y_train=np.ascontiguousarray(y[tr]),
        X_test=np.ascontiguousarray(Xs[te]),
        y_test=np.ascontiguousarray(y[te]),
        feature_names=[f"f{j}" for j in range(n_features)],
        classes=np.arange(n_classes),
        scale_mean=mean,
        scale_std=std,
        with_mean=with_mean,
    )
