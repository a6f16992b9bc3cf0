"""Declarative description of a federation: model, clients' shards, local optimisation,
aggregation rule, metric aggregation, stopping rule.

Defaults are the literals of the reference's custom script so that ``FedProblem.custom(...)``
with no extra arguments reproduces it: hidden ``[50, 200]`` (C:40), ``Adam(lr=0.004)`` (C:44),
``StepLR(30, 0.5)`` (C:46), one full-batch step per round (C:63-73), sample-weighted
averaging (C:110-115), mean-of-local metrics (C:169), patience 10 / tolerance 1e-4
(C:122).  ``FedProblem.sklearn_like`` carries sklearn's ``MLPClassifier`` defaults the other
two scripts rely on (S:77-83, H:90): sigmoid/log-loss head, Adam with sklearn's bias
correction, ``alpha=1e-4``, minibatch 200, ``tol=1e-4``, ``n_iter_no_change=10``,
unweighted averaging (S:114-117, H:37), pooled metrics (S:130-134).
"""
from __future__ import annotations

from dataclasses import dataclass, field, replace
from typing import List, Optional, Sequence

import numpy as np

from ..models.spec import MLPSpec


@dataclass
class TrainHyper:
    opt: str = "adam_torch"          # 'sgd' | 'adam_torch' | 'adam_sklearn'
    lr: float = 0.004
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    momentum: float = 0.0
    nesterov: bool = False
    alpha: float = 0.0                # L2, sklearn semantics
    lr_step_size: int = 30            # StepLR; 0 = constant
    lr_gamma: float = 0.5
    batch_size: int = 0               # 0 = full batch
    local_epochs: int = 1             # E: passes over the shard per round
    # sklearn-style "fit until converged" (used when max_iter > 0 instead of local_epochs)
    max_iter: int = 0
    tol: float = 1e-4
    n_iter_no_change: int = 10
    # per-epoch reshuffle of the shard (SK:704-718): 'feistel' = a keyed bijection evaluated on the device (default);
    # 'mt19937' = sklearn's own stream for random_state = the federation's init_seed, from a host-generated table
    # (parity runs against the reference: same mini-batches, epoch for epoch)
    shuffle_stream: str = "feistel"


@dataclass
class FedProblem:
    spec: MLPSpec
    shards: List[np.ndarray]                   # per client: int32 row indices into the train set
    hyper: TrainHyper = field(default_factory=TrainHyper)
    weighting: str = "samples"                 # 'samples' (C:110-115) | 'uniform' (S:114-117)
    metric_mode: str = "mean_of_local"         # | 'pooled'
    patience: int = 10                         # 0 disables early stopping
    atol: float = 1e-4
    rtol: float = 1e-5
    init: str = "torch_default"
    init_seed: int = 0
    init_per_client: bool = True               # reference: every rank draws its own init (C:42)
    name: str = ""
    # multi-GPU: this process hosts only some clients of the federation; the FedAvg denominator
    # (total samples N, or total client count) still refers to the whole federation
    total_weight: Optional[float] = None
    client_offset: int = 0                     # global index of this process's first client (init streams)

    @property
    def n_clients(self) -> int:
        return len(self.shards)

    @staticmethod
    def custom(n_features: int, n_classes: int, shards: Sequence[np.ndarray], hidden=(50, 200), **kw) -> "FedProblem":
        hyper = kw.pop("hyper", None) or TrainHyper()
        return FedProblem(MLPSpec.classifier(n_features, hidden, n_classes, "softmax_ce"), list(shards), hyper, **kw)

    @staticmethod
    def sklearn_like(n_features: int, shards: Sequence[np.ndarray], hidden=(50, 400), lr: float = 0.004,
                     max_iter: int = 300, seed: int = 42, **kw) -> "FedProblem":
        hyper = kw.pop("hyper", None) or TrainHyper(
            opt="adam_sklearn", lr=lr, alpha=1e-4, lr_step_size=0, lr_gamma=1.0, batch_size=200,
            max_iter=max_iter, tol=1e-4, n_iter_no_change=10)
        kw.setdefault("weighting", "uniform")
        kw.setdefault("metric_mode", "pooled")
        kw.setdefault("patience", 0)
        kw.setdefault("init", "sklearn_exact")
        kw.setdefault("init_seed", seed)
        kw.setdefault("init_per_client", False)  # random_state=42 on every rank (S:82)
        return FedProblem(MLPSpec.classifier(n_features, hidden, 2, "sigmoid_logloss"), list(shards), hyper, **kw)

    def with_hyper(self, **kw) -> "FedProblem":
        return replace(self, hyper=replace(self.hyper, **kw))
