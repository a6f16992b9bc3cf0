"""The sklearn-``MLPClassifier``-equivalent federation (S:68-153): every round each client is
*fitted to convergence* (minibatch 200, Adam with sklearn's bias correction, L2 1e-4, stop when
the training loss has not improved by ``tol`` for more than ``n_iter_no_change`` epochs or at
``max_iter``), local predictions are scored, coefficients are averaged unweighted, and pooled
metrics over all clients' (pre-averaging) predictions are reported.

What "apply the global weights, then fit" means is the subtle part of the reference:

* ``reference_faithful`` — what the shipped script actually does: ``warm_start`` is never set, so
  ``fit()`` re-initialises from ``random_state=42`` and the averaged weights applied at S:95-98
  are discarded; every round is identical (the "Limitation" in the file name).
* ``warm_start`` (default) — what the script evidently intends and BASELINE.json config 2 asks
  for: round r+1 starts from the averaged weights, with a fresh optimizer and fresh
  no-improvement bookkeeping, so clients really train ``<= max_iter`` epochs per round.
* ``sklearn_literal`` — sklearn's own ``warm_start=True`` behaviour: the optimizer is new but
  ``_no_improvement_count`` / ``best_loss_`` persist (SK:433-441), so every fit after the first
  stops after about one epoch (measured in BASELINE.md: n_iter_ = 1 in rounds 2-5).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np

from ..models.checkpoint import weight_statistics
from . import logging as flog
from .engine import FedEngine

POLICIES = ("warm_start", "reference_faithful", "sklearn_literal")


def run_sklearn_rounds(engine: FedEngine, rounds: int = 5, policy: str = "warm_start", log: Optional[flog.RunLog] = None,
                       shuffle: bool = True, fresh: bool = True) -> Dict:
    """``fresh=False`` continues a run this engine already started (round 1 of this call starts from the state the previous
    call left: the averaged weights under warm_start) instead of re-initialising."""
    if policy not in POLICIES:
        raise ValueError(f"policy must be one of {POLICIES}")
    n_iter_hist: List[np.ndarray] = []
    for r in range(rounds):
        if policy == "reference_faithful" or (r == 0 and fresh):
            engine.init_params()
        engine.reset_fit_state(keep_no_improvement=(policy == "sklearn_literal" and r > 0))
        if engine._begin_is_fused():
            engine.prime()                      # multi-GPU: the aggregation kernel begins the next round itself
        else:
            engine.enqueue_round_begin()
        n_iter = engine.fit_local(shuffle=shuffle)
        engine.enqueue_eval_local()
        engine.enqueue_aggregate()
        n_iter_hist.append(n_iter)
        if log is not None:
            h = engine.history()
            g0 = 0
            for f in range(engine.n_feds):          # history rows hold ALL clients of the federation, also under torchrun
                nk = len(getattr(engine, "global_problems", engine.problems)[f].shards)
                log.lines(flog.sklearn_round_lines(r + 1, h["local"][r][g0: g0 + nk], h["global"][f, r]))
                g0 += nk
            log.record(round=r + 1, global_metrics=h["global"][:, r], local_metrics=h["local"][r], n_iter=n_iter)
    engine.check_errors()
    h = engine.history()
    h["n_iter"] = np.asarray(n_iter_hist)
    if log is not None:
        coefs, inter = engine.problems[0].spec.to_coefs_intercepts(engine.global_slab(0))
        log.lines(flog.weight_stat_lines(weight_statistics(coefs, inter)))
    return h
