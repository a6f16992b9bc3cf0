"""MLP specification and the flat parameter-arena layout shared by every kernel.

The reference builds ``nn.Sequential(Linear, ReLU, …, Linear)``
(FL_Custom…Multiple_Rounds.py:12-25) or an sklearn ``MLPClassifier``
(FL_SkLearn…Limitation.py:77-83) per rank and ships weights around as dicts / lists of
arrays.  Here a model is a *slab*: one contiguous fp32 vector per client, identical layout
for every client of a federation, so aggregation is a single pass over ``[clients, P]`` and
the optimizer state arenas mirror it 1:1.

Slab layout (floats), for layers ``l = 0 … L-1`` with ``fan_in=d[l]``, ``fan_out=d[l+1]``::

    W_l : [fan_out, ld_l]   row-major, ld_l = round_up(fan_in, 4); pad columns are 0
    b_l : [round_up(fan_out, 4)]                                    pad entries are 0

i.e. torch's ``named_parameters()`` order and orientation (``model.0.weight`` is
``[out, in]``), with leading dimensions padded so every row starts 16-byte aligned.  Pad
entries stay exactly zero under every optimizer here (their gradient is identically zero
because the padded activations are zero).  The total is rounded up to 32 floats so slabs
are 128-byte aligned.  sklearn's ``coefs_`` / ``intercepts_`` orientation ``[in, out]`` is a
transpose performed only at the I/O edge (checkpoint, parity tests).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch

HEADS = ("softmax_ce", "sigmoid_logloss")
ACTS = ("relu", "logistic", "tanh", "identity")
MAX_LAYERS = 6  # Linear layers; must match FEDB_MAX_LAYERS in include/fedb200.h
FEATURE_PAD = 16  # device feature rows are padded to a multiple of 4; 14 -> 16


def _ru(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclass(frozen=True)
class MLPSpec:
    """``dims = (in, h1, …, hk, out)``.  ``out`` is the number of *output units*: the number
    of classes for ``softmax_ce``; 1 for ``sigmoid_logloss`` (binary, sklearn-style)."""

    dims: Tuple[int, ...]
    head: str = "softmax_ce"
    activation: str = "relu"

    def __post_init__(self):
        object.__setattr__(self, "dims", tuple(int(d) for d in self.dims))
        if self.head not in HEADS:
            raise ValueError(f"head must be one of {HEADS}")
        if self.activation not in ACTS:
            raise ValueError(f"activation must be one of {ACTS}")
        if not (2 <= len(self.dims) <= MAX_LAYERS + 1):
            raise ValueError(f"need 1..{MAX_LAYERS} Linear layers, got dims={self.dims}")
        if self.head == "sigmoid_logloss" and self.dims[-1] != 1:
            raise ValueError("sigmoid_logloss head has exactly one output unit")

    # ---- factory helpers -------------------------------------------------------------
    @staticmethod
    def classifier(n_features: int, hidden: Sequence[int], n_classes: int, head: str = "softmax_ce",
                   activation: str = "relu") -> "MLPSpec":
        out = n_classes if head == "softmax_ce" else 1
        if head == "sigmoid_logloss" and n_classes != 2:
            raise ValueError("sigmoid_logloss supports binary problems only")
        return MLPSpec((n_features, *hidden, out), head, activation)

    # ---- layout ----------------------------------------------------------------------
    @property
    def n_layers(self) -> int:
        return len(self.dims) - 1

    @property
    def n_classes(self) -> int:
        return 2 if self.head == "sigmoid_logloss" else self.dims[-1]

    def ld(self, l: int) -> int:
        return _ru(self.dims[l], 4)

    @property
    def w_offsets(self) -> Tuple[int, ...]:
        return self._layout()[0]

    @property
    def b_offsets(self) -> Tuple[int, ...]:
        return self._layout()[1]

    @property
    def slab_size(self) -> int:
        """Padded floats per client (P)."""
        return self._layout()[2]

    @property
    def n_params(self) -> int:
        """Real (unpadded) parameter count — what the reference would pickle."""
        return sum(self.dims[l] * self.dims[l + 1] + self.dims[l + 1] for l in range(self.n_layers))

    def _layout(self):
        w_off, b_off, off = [], [], 0
        for l in range(self.n_layers):
            w_off.append(off)
            off += self.dims[l + 1] * self.ld(l)
            b_off.append(off)
            off += _ru(self.dims[l + 1], 4)
        return tuple(w_off), tuple(b_off), _ru(off, 32)

    @property
    def max_width(self) -> int:
        return max(self.dims)

    # ---- views -----------------------------------------------------------------------
    def weight_view(self, slab: torch.Tensor, l: int, padded: bool = False) -> torch.Tensor:
        """``[..., out, in]`` view (torch orientation) of layer ``l`` inside ``slab[..., P]``."""
        o, i, ld = self.dims[l + 1], self.dims[l], self.ld(l)
        w = slab[..., self.w_offsets[l]: self.w_offsets[l] + o * ld].unflatten(-1, (o, ld))
        return w if padded else w[..., :i]

    def bias_view(self, slab: torch.Tensor, l: int) -> torch.Tensor:
        o = self.dims[l + 1]
        return slab[..., self.b_offsets[l]: self.b_offsets[l] + o]

    def param_mask(self, device=None) -> torch.Tensor:
        """bool[P]: True on real parameters, False on padding."""
        m = torch.zeros(self.slab_size, dtype=torch.bool, device=device)
        for l in range(self.n_layers):
            self.weight_view(m, l)[...] = True
            self.bias_view(m, l)[...] = True
        return m

    # ---- interchange with the reference's two in-memory formats ---------------------
    def to_state_dict(self, slab: torch.Tensor) -> Dict[str, np.ndarray]:
        """``{'model.0.weight': [out,in], 'model.0.bias': [out], 'model.2.weight': …}`` —
        the dict the custom script gathers (…Multiple_Rounds.py:93-94)."""
        sd = {}
        for l in range(self.n_layers):
            sd[f"model.{2 * l}.weight"] = self.weight_view(slab, l).detach().cpu().numpy().copy()
            sd[f"model.{2 * l}.bias"] = self.bias_view(slab, l).detach().cpu().numpy().copy()
        return sd

    def from_state_dict(self, sd: Dict[str, np.ndarray], slab: torch.Tensor) -> None:
        with torch.no_grad():
            slab.zero_()
            for l in range(self.n_layers):
                self.weight_view(slab, l).copy_(torch.as_tensor(np.asarray(sd[f"model.{2 * l}.weight"]), dtype=slab.dtype))
                self.bias_view(slab, l).copy_(torch.as_tensor(np.asarray(sd[f"model.{2 * l}.bias"]), dtype=slab.dtype))

    def to_coefs_intercepts(self, slab: torch.Tensor, dtype=np.float64) -> Tuple[List[np.ndarray], List[np.ndarray]]:
        """sklearn layout: ``coefs_[l]`` is ``[fan_in, fan_out]``, ``intercepts_[l]`` is
        ``[fan_out]`` (FL_SkLearn…Limitation.py:52-54, :109)."""
        coefs = [self.weight_view(slab, l).detach().cpu().numpy().T.astype(dtype).copy() for l in range(self.n_layers)]
        inter = [self.bias_view(slab, l).detach().cpu().numpy().astype(dtype).copy() for l in range(self.n_layers)]
        return coefs, inter

    def from_coefs_intercepts(self, coefs: Sequence[np.ndarray], intercepts: Sequence[np.ndarray], slab: torch.Tensor) -> None:
        if len(coefs) != self.n_layers or len(intercepts) != self.n_layers:
            raise ValueError("layer count mismatch")
        with torch.no_grad():
            slab.zero_()
            for l in range(self.n_layers):
                c = np.asarray(coefs[l])
                if c.shape != (self.dims[l], self.dims[l + 1]):
                    raise ValueError(f"coefs_[{l}] has shape {c.shape}, expected {(self.dims[l], self.dims[l + 1])}")
                self.weight_view(slab, l).copy_(torch.as_tensor(c.T.copy(), dtype=slab.dtype))
                self.bias_view(slab, l).copy_(torch.as_tensor(np.asarray(intercepts[l]), dtype=slab.dtype))

    def describe(self) -> str:
        return f"mlp{'-'.join(map(str, self.dims))}/{self.head}/{self.activation}"
