"""Hyper-parameter grid as packed concurrent federations (H:68-132).

The reference runs its 10 x 9 = 90 (hidden sizes x learning rate) federations one after
another, each with every rank fitting an ``MLPClassifier`` for up to 400 epochs (303 s at np=4 on
the survey box).  Here every grid point is one federation of the same engine: all 90 x clients
local fits advance together through the same grouped launches, then one grouped aggregation and
metric pass scores them all.  A local-epochs axis (BASELINE.json config 4) is optional.
Selection rule as in H:114-119: first strictly greater pooled accuracy wins, in grid order.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np

from . import logging as flog
from .engine import FedEngine
from .problem import FedProblem

# H:73-74, in the reference's iteration order
HIDDEN_GRID = [(50,), (100,), (50, 50), (100, 50), (50, 100), (50, 200), (50, 400), (100, 400), (400, 200), (200, 400)]
LR_GRID = [0.002, 0.005, 0.004, 0.008, 0.01, 0.02, 0.05, 0.1, 0.2]


def build_grid(n_features: int, shards: Sequence[np.ndarray], hidden_grid=HIDDEN_GRID, lr_grid=LR_GRID,
               max_iter_grid: Sequence[int] = (400,), seed: int = 42, init: str = "sklearn_exact") -> List[FedProblem]:
    probs = []
    for hl in hidden_grid:
        for lr in lr_grid:
            for mi in max_iter_grid:
                probs.append(FedProblem.sklearn_like(n_features, shards, hidden=tuple(hl), lr=lr, max_iter=mi, seed=seed,
                                                     init=init, name=f"hl={tuple(hl)},lr={lr},max_iter={mi}"))
    return probs


def grid_cost(p: FedProblem) -> float:
    """Relative cost of fitting one federation: parameters x rows (epochs to convergence are not known in advance)."""
    return float(p.spec.n_params) * float(sum(len(s) for s in p.shards))


def place_federations(costs: Sequence[float], world: int) -> List[List[int]]:
    """Longest-processing-time placement of the federations over ``world`` ranks: in decreasing cost, each to the least
    loaded rank ((400,200) costs ~100 x (50,): round-robin would leave most GPUs idle at the end).  Deterministic."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    load, out = [0.0] * world, [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda q: (load[q], q))
        out[r].append(i)
        load[r] += costs[i]
    return [sorted(x) for x in out]


def run_grid(data, n_clients: int = 4, hidden_grid=HIDDEN_GRID, lr_grid=LR_GRID, max_iter_grid: Sequence[int] = (400,),
             device: Optional[str] = None, log: Optional[flog.RunLog] = None, seed: int = 0, init: str = "sklearn_exact",
             shards: Optional[Sequence[np.ndarray]] = None, rank: int = 0, world: int = 1, timing: Optional[Dict] = None,
             shuffle_stream: str = "feistel") -> Dict:
    """The whole grid.  With ``world`` > 1 (one process per GPU under torchrun; H:83-106 runs every grid point over all
    ranks) the federations are placed over the ranks by cost, every rank fits its own with all of their clients resident —
    a grid point never crosses a GPU boundary — and the results are gathered once at the end; a federation's result does
    not depend on where it ran or what it was packed with, so the table is bit-identical to the single-GPU one."""
    from ..data.partition import contiguous
    shards = shards if shards is not None else contiguous(len(data.X_train), n_clients)
    probs = [p.with_hyper(shuffle_stream=shuffle_stream) for p in build_grid(data.n_features, shards, hidden_grid, lr_grid, max_iter_grid, init=init)]
    mine = place_federations([grid_cost(p) for p in probs], world)[rank] if world > 1 else list(range(len(probs)))
    results = []
    # federations with different max_iter stop at different epochs; run one engine per max_iter value
    for mi in max_iter_grid:
        ids = [i for i in mine if probs[i].hyper.max_iter == mi]
        if not ids:
            continue
        sub = [probs[i] for i in ids]
        eng = FedEngine(sub, data.X_train, data.y_train, data.X_test, data.y_test, device=device, max_rounds=2, seed=seed)
        if timing is not None and eng.device.type == "cuda":
            import torch
            torch.cuda.synchronize()
            t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
            t0.record()
        eng.reset_fit_state()
        eng.enqueue_round_begin()
        n_iter = eng.fit_local(max_iter=mi)
        eng.enqueue_eval_local()
        eng.enqueue_aggregate()
        if timing is not None and eng.device.type == "cuda":
            t1.record()
            torch.cuda.synchronize()
            timing["device_ms"] = timing.get("device_ms", 0.0) + t0.elapsed_time(t1)
            tcm = eng._fit_tc_mask()                                   # clients fitted by the one tensor-core launch
            cc = n_iter[~tcm] if tcm is not None else n_iter
            timing["epochs"] = timing.get("epochs", 0) + (int(cc.max()) if len(cc) else 0)
            timing["tc_clients"] = int(tcm.sum()) if tcm is not None else 0
            timing["clients"] = int(len(n_iter))
            timing["launches_per_epoch"] = 1 + eng.n_batches * 2 + 1      # shuffle, (fwd_bwd + reduce_opt) per mini-batch, epoch_end
        eng.check_errors()
        h = eng.history()
        test = eng.evaluate_global("test") if eng.n_test else [None] * len(sub)
        for f, p in enumerate(sub):
            gs = eng.group_of[f]
            results.append({"grid_index": ids[f], "hidden_layer_sizes": tuple(p.spec.dims[1:-1]), "learning_rate": p.hyper.lr, "max_iter": mi,
                            "global": h["global"][f, 0], "local": h["local"][0][gs], "n_iter": n_iter[gs],
                            "test_averaged_model": test[f], "weights": p.spec.to_coefs_intercepts(eng.global_slab(f)), "spec": p.spec})
    if world > 1:
        import torch.distributed as dist
        parts = [None] * world
        dist.all_gather_object(parts, results)
        results = [r for part in parts for r in part]
    results.sort(key=lambda r: r["grid_index"])          # the reference's iteration order (H:83-84) decides ties
    best = None
    for r in results:  # H:114: strictly greater, first wins
        if log is not None:
            log.lines(flog.tune_config_lines(r["hidden_layer_sizes"], r["learning_rate"], r["local"], r["global"]))
            log.record(hidden=r["hidden_layer_sizes"], lr=r["learning_rate"], max_iter=r["max_iter"], global_metrics=r["global"],
                       n_iter=r["n_iter"])
        if best is None or r["global"][0] > best["global"][0]:
            best = r
    out = {"results": results, "best": best,
           "best_params": {"hidden_layer_sizes": best["hidden_layer_sizes"], "learning_rate": best["learning_rate"]},
           "best_metrics": {n: float(best["global"][i]) for i, n in enumerate(("accuracy", "precision", "recall", "f1"))}}
    if len(max_iter_grid) > 1:
        out["best_params"]["max_iter"] = best["max_iter"]
    if log is not None:
        coefs, inter = best["weights"]
        log.lines(flog.tune_best_lines(out["best_params"], out["best_metrics"], list(coefs) + list(inter)))
    return out


def accuracy_table(results: List[Dict], hidden_grid=HIDDEN_GRID, lr_grid=LR_GRID) -> np.ndarray:
    """[len(hidden_grid), len(sorted lr)] pooled-accuracy table in the layout of SURVEY.md §6.2."""
    lrs = sorted(lr_grid)
    t = np.full((len(hidden_grid), len(lrs)), np.nan)
    for r in results:
        t[hidden_grid.index(tuple(r["hidden_layer_sizes"])), lrs.index(r["learning_rate"])] = r["global"][0]
    return t
