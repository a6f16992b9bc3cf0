"""The straightforward baseline BASELINE.json asks for: the same federation written the obvious way
with stock PyTorch — ``nn.Linear`` (cuBLAS), ``CrossEntropyLoss``, ``torch.optim.Adam`` + ``StepLR``,
evaluation with torch ops, and FedAvg as a sample-weighted ``torch.distributed.all_reduce`` (NCCL)
of the flattened parameters.  Data and weights stay on the GPU (unlike the reference), so this is
the *strong* baseline: what a competent PyTorch user would write before reaching for custom kernels.
None of the framework's kernels are used here.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn


def make_mlp(n_in: int, hidden: Sequence[int], n_out: int) -> nn.Sequential:
    layers: List[nn.Module] = []
    d = n_in
    for h in hidden:
        layers += [nn.Linear(d, h), nn.ReLU()]
        d = h
    layers.append(nn.Linear(d, n_out))
    return nn.Sequential(*layers)


class TorchFedAvg:
    """``clients`` local clients on this rank; ``world`` ranks in total."""

    def __init__(self, X: np.ndarray, y: np.ndarray, shards: Sequence[np.ndarray], total_samples: int, hidden=(50, 200),
                 n_classes: int = 2, lr: float = 0.004, step_size: int = 30, gamma: float = 0.5, device: str = "cuda", seed: int = 0):
        self.dev = torch.device(device)
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        torch.manual_seed(seed + (dist.get_rank() if dist.is_initialized() else 0))
        self.Xs = [torch.tensor(X[s], dtype=torch.float32, device=self.dev) for s in shards]
        self.ys = [torch.tensor(y[s], dtype=torch.long, device=self.dev) for s in shards]
        self.models = [make_mlp(X.shape[1], hidden, n_classes).to(self.dev) for _ in shards]
        self.opts = [torch.optim.Adam(m.parameters(), lr=lr) for m in self.models]
        self.scheds = [torch.optim.lr_scheduler.StepLR(o, step_size=step_size, gamma=gamma) for o in self.opts]
        self.weights = [len(s) / float(total_samples) for s in shards]
        self.n_classes = n_classes
        self.loss_fn = nn.CrossEntropyLoss()
        self.last_metrics: Optional[torch.Tensor] = None

    def _metrics(self, yt: torch.Tensor, yp: torch.Tensor) -> torch.Tensor:
        C = self.n_classes
        cm = torch.bincount(yt * C + yp, minlength=C * C).view(C, C).double()
        tp, sup, prd, tot = cm.diag(), cm.sum(1), cm.sum(0), cm.sum()
        prec = torch.where(prd > 0, tp / prd.clamp(min=1), torch.zeros_like(tp))
        rec = torch.where(sup > 0, tp / sup.clamp(min=1), torch.zeros_like(tp))
        f1 = torch.where(sup + prd > 0, 2 * tp / (sup + prd).clamp(min=1), torch.zeros_like(tp))
        w = sup / tot
        return torch.stack([tp.sum() / tot, (prec * w).sum(), (rec * w).sum(), (f1 * w).sum()])

    def round(self) -> None:
        local = []
        for m, o, s, X, y in zip(self.models, self.opts, self.scheds, self.Xs, self.ys):
            m.train()
            o.zero_grad(set_to_none=True)
            self.loss_fn(m(X), y).backward()
            o.step()
            s.step()
            m.eval()
            with torch.no_grad():
                local.append(self._metrics(y, m(X).argmax(1)))
        with torch.no_grad():
            met = torch.stack(local).sum(0)
            flat = sum(w * torch.nn.utils.parameters_to_vector(m.parameters()) for w, m in zip(self.weights, self.models))
            n_clients = torch.tensor([float(len(self.models))], device=self.dev, dtype=torch.float64)
            if self.world > 1:
                dist.all_reduce(flat)
                pack = torch.cat([met, n_clients])
                dist.all_reduce(pack)
                met, n_clients = pack[:4], pack[4:]
            self.last_metrics = met / n_clients
            for m in self.models:
                torch.nn.utils.vector_to_parameters(flat, m.parameters())
