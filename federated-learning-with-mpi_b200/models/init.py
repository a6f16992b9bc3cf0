"""Parameter initialisation.

Two distributions, selected per federation:

* ``torch_default`` — what ``nn.Linear`` does in the custom script
  (FL_Custom…Multiple_Rounds.py:18-21): ``W, b ~ U(-1/sqrt(fan_in), 1/sqrt(fan_in))``.
* ``sklearn_glorot`` — what ``MLPClassifier._init_coef`` does for the sklearn scripts
  (FL_SkLearn…Limitation.py:77-83): ``W, b ~ U(-s, s)``, ``s = sqrt(6/(fan_in+fan_out))``
  (``sqrt(2/…)`` for a logistic hidden activation).

The random stream is counter-based Philox4x32-10 keyed by ``(seed, client)`` and indexed by
the element's position in the slab, so the CUDA kernel (csrc/fedb_simt.cu: k_init) and this
NumPy restatement produce *bit-identical* slabs and every client/federation can be
initialised independently in one launch.  ``sklearn_exact`` additionally reproduces
sklearn's own MT19937 draw order for ``random_state=seed`` on the host (used for parity
tests against the reference's ``random_state=42`` models).
"""
from __future__ import annotations

import math

import numpy as np

from .spec import MLPSpec

INIT_KINDS = ("torch_default", "sklearn_glorot", "sklearn_exact", "zeros")

_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32(c0, c1, c2, c3, k0: int, k1: int):
    """Vectorised Philox4x32-10.  Counters are uint32 arrays (or scalars), key two ints."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & _MASK for c in (c0, c1, c2, c3))
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)) & _MASK, lo1, (hi0 ^ c3 ^ np.uint64(k1)) & _MASK, lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def layer_bound(spec: MLPSpec, l: int, kind: str) -> float:
    fan_in, fan_out = spec.dims[l], spec.dims[l + 1]
    if kind == "torch_default":
        return 1.0 / math.sqrt(fan_in)
    if kind in ("sklearn_glorot", "sklearn_exact"):
        factor = 2.0 if spec.activation == "logistic" else 6.0
        return math.sqrt(factor / (fan_in + fan_out))
    if kind == "zeros":
        return 0.0
    raise ValueError(f"unknown init {kind!r}")


def bound_vector(spec: MLPSpec, kind: str) -> np.ndarray:
    """float32[P]: per-element uniform bound, 0 on padding."""
    b = np.zeros(spec.slab_size, dtype=np.float32)
    for l in range(spec.n_layers):
        s = np.float32(layer_bound(spec, l, kind))
        o, i, ld = spec.dims[l + 1], spec.dims[l], spec.ld(l)
        w = b[spec.w_offsets[l]: spec.w_offsets[l] + o * ld].reshape(o, ld)
        w[:, :i] = s
        b[spec.b_offsets[l]: spec.b_offsets[l] + o] = s
    return b


def init_slab_philox(spec: MLPSpec, kind: str, seed: int, stream: int) -> np.ndarray:
    """float32[P] for one client.  Element ``i`` uses word ``i % 4`` of
    ``philox(counter=(i // 4, stream, 0, 0), key=(seed_lo, seed_hi))``; the uniform is
    ``u = (x >> 8) * 2^-24`` and the value ``(2u - 1) * bound`` computed in fp32."""
    P = spec.slab_size
    q = np.arange(P // 4, dtype=np.uint32)
    words = philox4x32(q, np.uint32(stream & 0xFFFFFFFF), np.uint32(0), np.uint32(0), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    x = np.stack(words, axis=1).reshape(-1)
    u = (x >> np.uint32(8)).astype(np.float32) * np.float32(2.0 ** -24)
    val = (np.float32(2.0) * u - np.float32(1.0)) * bound_vector(spec, kind)
    return val.astype(np.float32)


def init_slab_sklearn_exact(spec: MLPSpec, seed: int) -> np.ndarray:
    """Reproduce ``MLPClassifier(random_state=seed)._initialize`` draw-for-draw:
    for each layer, ``uniform(-s, s, (fan_in, fan_out))`` then ``uniform(-s, s, fan_out)``."""
    rs = np.random.RandomState(seed)
    out = np.zeros(spec.slab_size, dtype=np.float32)
    for l in range(spec.n_layers):
        s = layer_bound(spec, l, "sklearn_exact")
        fan_in, fan_out, ld = spec.dims[l], spec.dims[l + 1], spec.ld(l)
        coef = rs.uniform(-s, s, (fan_in, fan_out))
        inter = rs.uniform(-s, s, fan_out)
        w = out[spec.w_offsets[l]: spec.w_offsets[l] + fan_out * ld].reshape(fan_out, ld)
        w[:, :fan_in] = coef.T.astype(np.float32)
        out[spec.b_offsets[l]: spec.b_offsets[l] + fan_out] = inter.astype(np.float32)
    return out


def init_slab(spec: MLPSpec, kind: str, seed: int, stream: int = 0) -> np.ndarray:
    if kind == "sklearn_exact":
        return init_slab_sklearn_exact(spec, seed)
    return init_slab_philox(spec, kind, seed, stream)


def sklearn_shuffle_table(spec: MLPSpec, seed: int, n_rows: int, n_epochs: int) -> np.ndarray:
    """int32[n_epochs, n_rows]: the order in which ``MLPClassifier(random_state=seed, shuffle=True).fit`` visits its
    ``n_rows`` samples in each epoch, draw for draw: the model's MT19937 stream first serves the Glorot initialisation
    (as in :func:`init_slab_sklearn_exact`), then every epoch applies ``sklearn.utils.shuffle`` to the index vector left
    by the previous one (SK:704-718: ``sample_idx = shuffle(sample_idx, random_state=self._random_state)``, which is
    ``sample_idx[perm]`` with ``perm`` = an in-place ``RandomState.shuffle`` of ``arange(n)``)."""
    rs = np.random.RandomState(seed)
    for l in range(spec.n_layers):
        s_ = layer_bound(spec, l, "sklearn_exact")
        rs.uniform(-s_, s_, (spec.dims[l], spec.dims[l + 1]))
        rs.uniform(-s_, s_, spec.dims[l + 1])
    cur = np.arange(n_rows)
    out = np.empty((n_epochs, n_rows), dtype=np.int32)
    for ep in range(n_epochs):
        perm = np.arange(n_rows)
        rs.shuffle(perm)
        cur = cur[perm]
        out[ep] = cur
    return out
