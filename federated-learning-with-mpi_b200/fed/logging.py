"""Human-readable round reports in the reference's own line formats, plus JSONL.

Formats mirror C:139-179 (custom script), S:87-150 (sklearn script) and H:82-132 (grid search)
so that a run here can be diffed against a reference run; the engine keeps every number on the
device, and these functions format the history after the fact (or once per polled chunk).
"""
from __future__ import annotations

import json
import sys
from typing import Dict, IO, List, Optional, Sequence

import numpy as np

from .metrics import METRIC_NAMES


def _fmt4(m: Sequence[float]) -> str:
    return (f"[accuracy: {m[0]:.4f}, precision: {m[1]:.4f}, recall: {m[2]:.4f}, f1: {m[3]:.4f}]")


def custom_round_lines(round_no: int, local: np.ndarray, glob: Sequence[float]) -> List[str]:
    """Lines the custom script prints for one round (C:139, C:154-161, C:172-179)."""
    out = [f"\nRound {round_no}:\n"]
    for k in range(local.shape[0]):
        out.append(f"  RANK {k} - Local Metrics (Round {round_no}): {_fmt4(local[k])}")
    out.append(f"  Global Metrics (Round {round_no}): {_fmt4(glob)}")
    return out


def custom_stop_lines(patience: int, stopped_round: int) -> List[str]:
    return [f"Early stopping triggered: No significant change in metrics for {patience} rounds.",
            f"Training stopped early at round {stopped_round}."]


def sklearn_round_lines(round_no: int, local: np.ndarray, glob: Sequence[float]) -> List[str]:
    out = []
    for k in range(local.shape[0]):
        out.append(f"\n[Rank {k}] Starting Round {round_no}")
    for k in range(local.shape[0]):
        d = {n: float(local[k][i]) for i, n in enumerate(METRIC_NAMES)}
        out.append(f"[Rank {k}] Local Metrics after training (Round {round_no}): {d}")
    out.append(f"[Rank 0] Computed global weights after Round {round_no}")
    out += [f"\n[Rank 0] Global Metrics for Round {round_no}:", f"  Accuracy: {glob[0]:.4f}", f"  Precision: {glob[1]:.4f}",
            f"  Recall: {glob[2]:.4f}", f"  F1: {glob[3]:.4f}", "-" * 50]
    return out


def weight_stat_lines(stats) -> List[str]:
    out = ["\nFinal Global Weight Statistics:"]
    for i, (shape, mean, std) in enumerate(stats):
        out += [f"Layer {i + 1} - Shape: {shape}", f"Mean: {mean:.6f}, Std: {std:.6f}"]
    return out


def tune_config_lines(hidden, lr, local: np.ndarray, glob: Sequence[float]) -> List[str]:
    out = []
    for k in range(local.shape[0]):
        d = {n: float(local[k][i]) for i, n in enumerate(METRIC_NAMES)}
        out += ["\n\tLOCAL MEASURED RESULTS\n", f"\t[Rank {k}] Local Metrics (Hidden Layers: {tuple(hidden)}, LR: {lr}): {d}\n", "-" * 50]
    d = {n: float(glob[i]) for i, n in enumerate(METRIC_NAMES)}
    out += ["\nGLOBAL MEASURED RESULTS", f"\t[Rank 0] Global Metrics (Hidden Layers: {tuple(hidden)}, LR: {lr}): {d}\n", "-" * 50]
    return out


def tune_best_lines(best_params: Dict, best_metrics: Dict, weights: Sequence[np.ndarray]) -> List[str]:
    out = ["\n\nBest MEASURED RESULTS", f"\nBest Global Hyperparameters: {best_params}", f"Best Global Metrics: {best_metrics}",
           "\nBest Global Weights:"]
    for i, w in enumerate(weights):
        out.append(f"Layer {i + 1}: {w.shape}\n{w}")
    return out


class RunLog:
    """stdout (reference formats) + optional JSONL (one object per round)."""

    def __init__(self, jsonl_path: Optional[str] = None, stream: Optional[IO] = None, quiet: bool = False):
        self.stream = stream or sys.stdout
        self.quiet = quiet
        self.jsonl = open(jsonl_path, "w") if jsonl_path else None

    def lines(self, lines: Sequence[str]) -> None:
        if not self.quiet:
            for l in lines:
                print(l, file=self.stream, flush=True)

    def record(self, **obj) -> None:
        if self.jsonl:
            self.jsonl.write(json.dumps(obj, default=lambda o: o.tolist() if hasattr(o, "tolist") else str(o)) + "\n")
            self.jsonl.flush()

    def close(self) -> None:
        if self.jsonl:
            self.jsonl.close()
