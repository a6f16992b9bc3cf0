"""Grouped tensor-core GEMM (csrc/fedb_gemm.cu) — the building block of the wide-MLP path."""
from __future__ import annotations

import os
from typing import Optional

import numpy as np
import torch

from . import _native as nat
from . import kernels as K

GEMM_DTYPE = np.dtype([
    ("A", "u8"), ("B", "u8"), ("a_group_stride", "i8"), ("b_group_stride", "i8"), ("lda", "i4"), ("ldb", "i4"),
    ("M", "i4"), ("N", "i4"), ("K", "i4"), ("a_mn", "i4"), ("b_mn", "i4"), ("epilogue", "i4"),
    ("bias", "u8"), ("bias_group_stride", "i8"), ("C", "u8"), ("c_group_stride", "i8"), ("ldc", "i4"), ("ldact", "i4"),
    ("act", "u8"), ("act_group_stride", "i8"), ("W", "u8"), ("Mom", "u8"), ("Vel", "u8"), ("Wsh", "u8"),
    ("w_group_stride", "i8"), ("ldw", "i4"), ("l2_scale", "f4"), ("groups", "u8"), ("step", "u8"),
], align=True)
FWD_BIAS_RELU, DGRAD_MASK, WGRAD_OPT, STORE_F32 = 0, 1, 2, 3


def _p(t) -> int:
    return 0 if t is None else (t if isinstance(t, int) else t.data_ptr())


def gemm(n_groups: int, M: int, N: int, K_: int, A, a_stride: int, lda: int, a_mn: bool, B, b_stride: int, ldb: int, b_mn: bool,
         epilogue: int, C=None, c_stride: int = 0, ldc: int = 0, bias=None, bias_stride: int = 0, act=None, act_stride: int = 0,
         ldact: int = 0, W=None, Mom=None, Vel=None, Wsh=None, w_stride: int = 0, ldw: int = 0, l2_scale: float = 0.0,
         groups=None, step=None) -> None:
    """Pointers may be tensors or raw addresses (for views into arenas use ``tensor[offset:].data_ptr()``)."""
    lib = nat.lib(required=True)
    assert lib.fedb_gemm_args_size() == GEMM_DTYPE.itemsize, (lib.fedb_gemm_args_size(), GEMM_DTYPE.itemsize)
    a = np.zeros(1, dtype=GEMM_DTYPE)
    a["A"], a["B"], a["a_group_stride"], a["b_group_stride"], a["lda"], a["ldb"] = _p(A), _p(B), a_stride, b_stride, lda, ldb
    a["M"], a["N"], a["K"], a["a_mn"], a["b_mn"], a["epilogue"] = M, N, K_, int(a_mn), int(b_mn), epilogue
    a["bias"], a["bias_group_stride"], a["C"], a["c_group_stride"], a["ldc"] = _p(bias), bias_stride, _p(C), c_stride, ldc
    a["act"], a["act_group_stride"], a["ldact"] = _p(act), act_stride, ldact
    a["W"], a["Mom"], a["Vel"], a["Wsh"], a["w_group_stride"], a["ldw"], a["l2_scale"] = _p(W), _p(Mom), _p(Vel), _p(Wsh), w_stride, ldw, l2_scale
    a["groups"], a["step"] = _p(groups), _p(step)
    # FEDB200_GEMM: "auto" (default) = CTA-pair kernels when M and N are multiples of 256 (the persistent one for the
    # HBM-bound optimizer epilogue, the one-tile-per-cluster one otherwise: measured faster for fwd/dgrad), else the
    # single-CTA TMA kernel; "tma2" / "tma2p" / "tma" force one kernel; "cpasync" = the cp.async fallback.
    mode = os.environ.get("FEDB200_GEMM", "auto")
    if mode == "auto":
        mode = "tma2p" if epilogue == WGRAD_OPT else "tma2"
    if mode == "tma2p" and M % 256 == 0 and N % 256 == 0:       # persistent CTA-pair kernel (csrc/fedb_gemm_tma2p.cu)
        nat.check(lib.fedb_gemm_tma2p(a.ctypes.data, n_groups, _p(_pair_err()), nat.cuda_stream()))
    elif mode in ("tma2", "tma2p") and M % 256 == 0 and N % 256 == 0 and PROF is not None:
        nat.check(lib.fedb_gemm_tma2_prof(a.ctypes.data, n_groups, _p(_pair_err()), _p(PROF), nat.cuda_stream()))
    elif mode in ("tma2", "tma2p") and M % 256 == 0 and N % 256 == 0:      # CTA-pair kernel (csrc/fedb_gemm_tma2.cu)
        nat.check(lib.fedb_gemm_tma2(a.ctypes.data, n_groups, _p(_pair_err()), nat.cuda_stream()))
    else:
        fn = lib.fedb_gemm_tc if mode == "cpasync" else lib.fedb_gemm_tma
        nat.check(fn(a.ctypes.data, n_groups, nat.cuda_stream()))
    K._count()


_PAIR_ERR: Optional[torch.Tensor] = None
PROF: Optional[torch.Tensor] = None      # set to an int64 device tensor of 8 * n_ctas to collect phase stamps (tools/gemm_timeline.py)


def _pair_err() -> torch.Tensor:
    """Device flag the CTA-pair kernel sets if its barrier protocol stalls (see ``pair_status``)."""
    global _PAIR_ERR
    if _PAIR_ERR is None:
        _PAIR_ERR = torch.zeros(1, dtype=torch.int32, device="cuda")
    return _PAIR_ERR


def pair_status() -> int:
    """0 if every CTA-pair GEMM so far completed its pipeline; a FEDB_E_* code otherwise (synchronises)."""
    return 0 if _PAIR_ERR is None else int(_PAIR_ERR.item())
