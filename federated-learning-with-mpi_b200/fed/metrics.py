"""Classification metrics from confusion-matrix counts, and the round early-stop rule.

The reference calls four sklearn functions on host arrays every round
(FL_Custom…Multiple_Rounds.py:85-90, FL_SkLearn…Limitation.py:56-66) — measured at ~40 % of
its CPU round time.  Here predictions never leave the device: the eval kernel accumulates a
``C x C`` count matrix per client (rows = true class, cols = predicted) and everything below
is arithmetic on those counts.  The same formulas run on the device (csrc: round_metrics)
and here (host restatement used on CPU and in tests):

* accuracy = trace / total
* per class: precision = tp/(tp+fp), recall = tp/(tp+fn), f1 = 2tp/(2tp+fp+fn), each 0 when
  its denominator is 0 (``zero_division=0``)
* ``average='weighted'``: weights = per-class support (true counts) / total

Aggregations over clients: ``mean_of_local`` (unweighted mean of per-client metrics,
…Multiple_Rounds.py:169) and ``pooled`` (metrics of the summed confusion matrix, which is what
concatenating all clients' predictions gives, …Limitation.py:130-134).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Sequence

import numpy as np

METRIC_NAMES = ("accuracy", "precision", "recall", "f1")


def confusion_matrix(y_true: np.ndarray, y_pred: np.ndarray, n_classes: int) -> np.ndarray:
    cm = np.zeros((n_classes, n_classes), dtype=np.int64)
    np.add.at(cm, (np.asarray(y_true, dtype=np.int64), np.asarray(y_pred, dtype=np.int64)), 1)
    return cm


def metrics_from_confusion(cm: np.ndarray) -> Dict[str, float]:
    cm = np.asarray(cm, dtype=np.float64)
    total = cm.sum()
    if total == 0:
        return {k: 0.0 for k in METRIC_NAMES}
    tp = np.diag(cm)
    support = cm.sum(axis=1)
    predicted = cm.sum(axis=0)
    with np.errstate(divide="ignore", invalid="ignore"):
        prec = np.where(predicted > 0, tp / predicted, 0.0)
        rec = np.where(support > 0, tp / support, 0.0)
        den = support + predicted
        f1 = np.where(den > 0, 2.0 * tp / den, 0.0)
    w = support / total
    return {
        "accuracy": float(tp.sum() / total),
        "precision": float((prec * w).sum()),
        "recall": float((rec * w).sum()),
        "f1": float((f1 * w).sum()),
    }


def mean_of_local(cms: Sequence[np.ndarray]) -> Dict[str, float]:
    per = [metrics_from_confusion(c) for c in cms]
    return {k: float(np.mean([m[k] for m in per])) for k in METRIC_NAMES}


def pooled(cms: Sequence[np.ndarray]) -> Dict[str, float]:
    return metrics_from_confusion(np.sum(np.asarray(cms), axis=0))


@dataclass
class EarlyStop:
    """The custom script's stopping rule (…Multiple_Rounds.py:122-127, :181-192).

    An *anchor* metric vector is kept; each round whose four global metrics are all within
    ``atol + rtol*|anchor|`` of the anchor decrements a counter that starts at ``patience``;
    any other round re-anchors to the current metrics and resets the counter.  The anchor is
    NOT moved by a matching round.  ``stop`` becomes true when the counter hits zero; the
    engine acts on it at the top of the *next* round, after this round's aggregation has
    run (…Multiple_Rounds.py:195-198, :132-136)."""

    patience: int = 10
    atol: float = 1e-4
    rtol: float = 1e-5
    count: Optional[int] = None
    anchor: Optional[np.ndarray] = None
    stop: bool = False

    def __post_init__(self):
        if self.count is None:
            self.count = self.patience

    def update(self, metrics: Dict[str, float]) -> bool:
        cur = np.asarray([metrics[k] for k in METRIC_NAMES], dtype=np.float64)
        if self.anchor is not None and bool(np.all(np.abs(cur - self.anchor) <= self.atol + self.rtol * np.abs(self.anchor))):
            self.count -= 1
            if self.count == 0:
                self.stop = True
        else:
            self.anchor = cur
            self.count = self.patience
        return self.stop
