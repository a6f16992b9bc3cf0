"""Weight checkpoints in the reference's ``coefs_`` / ``intercepts_`` per-layer layout.

The reference never writes a file — it prints statistics (FL_SkLearn…Limitation.py:146-150)
or whole arrays (hyperparameters_tuning.py:130-132) of the list
``coefs_ + intercepts_ = [W1(in,h1), …, Wk(h,out), b1(h1,), …, bk(out,)]``
(FL_SkLearn…Limitation.py:109).  The checkpoint is that list as an ``.npz``:

    coefs_0 … coefs_{L-1}            float64 [fan_in, fan_out]
    intercepts_0 … intercepts_{L-1}  float64 [fan_out]
    meta                              JSON string: dims, head, activation, round, extras

plus, optionally (``resume=True``), the optimizer state of every client so a run can be
continued exactly: ``opt_m`` / ``opt_v`` ``[clients, P]`` fp32, ``opt_step`` ``[clients]``.
``to_state_dict`` gives the torch-orientation view (``model.{2l}.weight`` ``[out,in]``) the
custom script uses (FL_Custom…Multiple_Rounds.py:93-99).
"""
from __future__ import annotations

import json
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .spec import MLPSpec


def save_checkpoint(path: str, spec: MLPSpec, slab: torch.Tensor, round_idx: int = 0,
                    extras: Optional[Dict[str, Any]] = None,
                    opt_state: Optional[Dict[str, torch.Tensor]] = None) -> None:
    coefs, inter = spec.to_coefs_intercepts(slab)
    arrays: Dict[str, np.ndarray] = {}
    for l, (c, b) in enumerate(zip(coefs, inter)):
        arrays[f"coefs_{l}"] = c
        arrays[f"intercepts_{l}"] = b
    meta = {"dims": list(spec.dims), "head": spec.head, "activation": spec.activation,
            "round": int(round_idx), "format": "fedb200-coefs-intercepts-v1"}
    if extras:
        meta["extras"] = extras
    arrays["meta"] = np.asarray(json.dumps(meta))
    if opt_state:
        for k, v in opt_state.items():
            arrays[f"opt_{k}"] = v.detach().cpu().numpy()
    np.savez(path, **arrays)


def load_checkpoint(path: str) -> Tuple[MLPSpec, List[np.ndarray], List[np.ndarray], Dict[str, Any], Dict[str, np.ndarray]]:
    """Returns ``(spec, coefs_, intercepts_, meta, opt_state)``."""
    with np.load(path, allow_pickle=False) as z:
        meta = json.loads(str(z["meta"]))
        spec = MLPSpec(tuple(meta["dims"]), meta["head"], meta["activation"])
        coefs = [z[f"coefs_{l}"] for l in range(spec.n_layers)]
        inter = [z[f"intercepts_{l}"] for l in range(spec.n_layers)]
        opt = {k[4:]: z[k] for k in z.files if k.startswith("opt_")}
    return spec, coefs, inter, meta, opt


def load_into_slab(path: str, slab: torch.Tensor, expect: Optional[MLPSpec] = None) -> Tuple[MLPSpec, Dict[str, Any]]:
    spec, coefs, inter, meta, _ = load_checkpoint(path)
    if expect is not None and (expect.dims != spec.dims or expect.head != spec.head):
        raise ValueError(f"checkpoint is {spec.describe()}, expected {expect.describe()}")
    spec.from_coefs_intercepts(coefs, inter, slab)
    return spec, meta


def weight_statistics(coefs: Sequence[np.ndarray], intercepts: Sequence[np.ndarray]) -> List[Tuple[Tuple[int, ...], float, float]]:
    """(shape, mean, std) per array of ``coefs_ + intercepts_`` — the numbers the sklearn
    script prints after its last round (FL_SkLearn…Limitation.py:146-150)."""
    return [(tuple(w.shape), float(np.mean(w)), float(np.std(w))) for w in list(coefs) + list(intercepts)]
