"""Tensor-core (tcgen05) support code on the Python side: operand-panel layout helpers and the
hardware probe used by tests/test_gpu_tc.py to validate the descriptor conventions."""
from __future__ import annotations

import ctypes
from typing import Sequence

import numpy as np
import torch

from . import _native as nat

_P, _I32 = ctypes.c_void_p, ctypes.c_int32

PANEL_COLS = 64


def panel_offsets(rows: int) -> np.ndarray:
    """int64[rows, 64]: byte offset of element (r, c) in a 128B-swizzled bf16 panel
    (csrc/tc_common.cuh: panel_off)."""
    r = np.arange(rows)[:, None]
    c = np.arange(PANEL_COLS)[None, :]
    return r * 128 + (((c >> 3) ^ (r & 7)) << 4) + ((c & 7) << 1)


def to_panels(mat: torch.Tensor, dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """[rows, cols] float -> uint8 image of ceil(cols/64) consecutive panels (bf16 or fp16, swizzled).
    ``rows`` must be a multiple of 8."""
    rows, cols = mat.shape
    assert rows % 8 == 0
    npan = (cols + PANEL_COLS - 1) // PANEL_COLS
    padded = torch.zeros(rows, npan * PANEL_COLS, dtype=torch.float32)
    padded[:, :cols] = mat
    bits = padded.to(dtype).view(torch.int16).numpy().view(np.uint16)
    img = np.zeros(npan * rows * 128, dtype=np.uint8)
    off = panel_offsets(rows)
    img16 = img.view(np.uint16)
    for p in range(npan):
        img16[(p * rows * 128 + off) // 2] = bits[:, p * 64: (p + 1) * 64]
    return torch.from_numpy(img)


def idesc_bf16(M: int, N: int, a_mn: bool = False, b_mn: bool = False) -> int:
    return (1 << 4) | (1 << 7) | (1 << 10) | (int(a_mn) << 15) | (int(b_mn) << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def idesc_f16(M: int, N: int, a_mn: bool = False, b_mn: bool = False) -> int:
    """kind::f16 instruction descriptor with fp16 operands (format field 0), fp32 accumulation."""
    return (1 << 4) | (int(a_mn) << 15) | (int(b_mn) << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def probe(a_img: torch.Tensor, b_img: torch.Tensor, N: int, a_steps: Sequence[int], b_steps: Sequence[int],
          a_lbo: int, a_sbo: int, b_lbo: int, b_sbo: int, a_mn: bool = False, b_mn: bool = False,
          repeat: int = 1, fp16: bool = False) -> torch.Tensor:
    """Run a chain of M=128 bf16 (or fp16) MMAs on the given smem images; returns D as fp32 [128, N]."""
    lib = nat.lib(required=True)
    assert lib.fedb_tc_probe_args_size() == 74 * 4
    args = np.zeros(74, dtype=np.int32)
    args[0], args[1] = a_img.numel(), b_img.numel()
    args[2:6] = [a_lbo, a_sbo, b_lbo, b_sbo]
    args[6] = (idesc_f16 if fp16 else idesc_bf16)(128, N, a_mn, b_mn)
    args[7] = len(a_steps)
    args[8: 8 + len(a_steps)] = a_steps
    args[40: 40 + len(b_steps)] = b_steps
    args[72], args[73] = N, repeat
    out = torch.empty(128, N, dtype=torch.float32, device="cuda")
    a_d, b_d = a_img.cuda(), b_img.cuda()
    nat.check(lib.fedb_tc_probe(a_d.data_ptr(), b_d.data_ptr(), args.ctypes.data_as(ctypes.POINTER(_I32)), out.data_ptr(),
                                nat.cuda_stream()))
    torch.cuda.synchronize()
    return out


# ---- fused tensor-core local round -------------------------------------------------------------
# Two kernels with the same arguments: variant "bf16" (csrc/fedb_tc.cu: bf16 operands, 128-row tiles, clusters <= 8) and
# variant "x2" (csrc/fedb_tcx.cu: every operand split into two fp16 numbers, three MMAs per GEMM — fp32-grade results —
# 64-row tiles, clusters <= 16).
TC_ROWS = 128
TC_MAX_SLAB = 11920   # floats; fedb_tc_max_slab(): the gradient slices of all 8 CTAs must fit one CTA's exchange area
TCX_ROWS = 64


def tc_eligible(groups: np.ndarray, variant: str = "bf16") -> bool:
    """True when every group fits the fused tcgen05 kernel: 2 hidden ReLU layers, in <= 15, h1 <= 63, h2 <= 256, out <= 8,
    and the cross-CTA exchange areas hold the model (bf16: slab size; x2: fedb_tcx_fits at the planned cluster size)."""
    for g in groups:
        d = g["dims"]
        if int(g["n_layers"]) != 3 or int(g["act"]) != nat.ACT_IDS["relu"]:
            return False
        if not (d[0] <= 15 and d[1] <= 63 and d[2] <= 256 and d[3] <= 8):
            return False
        if variant != "x2" and int(g["slab_size"]) > TC_MAX_SLAB:
            return False
    return variant != "x2" or tc_cluster_size(groups, "x2") > 0


def _rows_per_step(groups: np.ndarray) -> int:
    rows = 1
    for g in groups:
        n = int(g["n_rows"])
        rows = max(rows, min(int(g["batch_size"]), n) if int(g["batch_size"]) > 0 else n)
    return rows


def tc_cluster_size(groups: np.ndarray, variant: str = "bf16") -> int:
    """CTAs per client: one per row tile, up to the largest cluster the kernel runs with.  x2: 16 where the device can
    co-schedule it and the grid then still fits the GPU in one wave, else <= 8; the largest size <= that at which every
    model's exchange blocks fit (0: none does)."""
    rows = _rows_per_step(groups)
    if variant != "x2":
        return int(min(8, max(1, (rows + TC_ROWS - 1) // TC_ROWS)))
    lib = nat.lib(required=True)
    want = max(1, (rows + TCX_ROWS - 1) // TCX_ROWS)
    if want > 8:
        want = 16 if (int(lib.fedb_tcx_max_cluster()) >= 16 and len(groups) * 16 <= 128) else 8
    shapes = {tuple(int(x) for x in g["dims"][:4]) for g in groups}
    for cs in range(want, 0, -1):
        if all(lib.fedb_tcx_fits(d0, h1, h2, c, cs) for d0, h1, h2, c in shapes):
            return cs
    return 0


def local_round_tc(tb, cluster_size: int, X, y, idx, params, m, v, step, active, fed_state, loss_out, conf, scratch,
                   batch: int = 0, n_steps: int = 1, do_eval: bool = True, prof=None, variant: str = "bf16") -> None:
    """``prof``: optional int64[32] device tensor receiving clock64 stamps of CTA 0 (debug)."""
    lib = nat.lib(required=True)
    fn = lib.fedb_local_round_tcx if variant == "x2" else lib.fedb_local_round_tc_prof
    nat.check(fn(nat.ptr(tb.d_groups), tb.n_groups, cluster_size, nat.ptr(X), X.shape[1], nat.ptr(y),
                                           nat.ptr(idx), nat.ptr(params), nat.ptr(m), nat.ptr(v), nat.ptr(step), nat.ptr(active),
                                           nat.ptr(fed_state), nat.ptr(loss_out), nat.ptr(conf), nat.ptr(scratch), batch, n_steps,
                                           int(do_eval), nat.ptr(prof), nat.cuda_stream()))
    from . import kernels as K
    K._count()


def round_tc(tb, cluster_size: int, X, y, idx, params, m, v, step, active, fed_state, loss_out, conf, scratch, n_steps: int,
             global_, hist_local, hist_global, tail_count, prof=None, state_mirror=None, host_inputs=None,
             variant: str = "bf16") -> None:
    """The whole round in one launch (see fedb_round_tc / fedb_round_tc_hostin in include/fedb200.h).
    ``host_inputs`` = (x_host, y_host) pinned tensors: the kernel pulls the step's rows from them itself."""
    lib = nat.lib(required=True)
    xh, yh = host_inputs if host_inputs is not None else (None, None)
    fn = lib.fedb_round_tcx if variant == "x2" else lib.fedb_round_tc_hostin
    nat.check(fn(nat.ptr(tb.d_groups), tb.n_groups, cluster_size, nat.ptr(X), X.shape[1], nat.ptr(y),
                                nat.ptr(xh), nat.ptr(yh), nat.ptr(idx),
                                nat.ptr(params), nat.ptr(m), nat.ptr(v), nat.ptr(step), nat.ptr(active), nat.ptr(fed_state),
                                nat.ptr(loss_out), nat.ptr(conf), nat.ptr(scratch), n_steps, nat.ptr(tb.d_feds), tb.n_feds,
                                nat.ptr(global_), nat.ptr(hist_local), nat.ptr(hist_global), conf.numel(), nat.ptr(tail_count),
                                nat.ptr(state_mirror), nat.ptr(prof), nat.cuda_stream()))
    from . import kernels as K
    K._count()


def round_tc_p2p(tb, cluster_size: int, X, y, idx, params, m, v, step, active, fed_state, loss_out, conf, scratch, n_steps: int,
                 rank_dev, state_mirror=None, host_inputs=None, variant: str = "bf16") -> None:
    """Multi-GPU round in one launch per rank (one federation, one resident client): fedb_round_tc_p2p[_hostin]."""
    lib = nat.lib(required=True)
    xh, yh = host_inputs if host_inputs is not None else (None, None)
    fn = lib.fedb_round_tcx_p2p if variant == "x2" else lib.fedb_round_tc_p2p_hostin
    nat.check(fn(nat.ptr(tb.d_groups), cluster_size, nat.ptr(X), X.shape[1], nat.ptr(y), nat.ptr(xh), nat.ptr(yh), nat.ptr(idx),
                                    nat.ptr(params), nat.ptr(m), nat.ptr(v), nat.ptr(step), nat.ptr(active), nat.ptr(fed_state),
                                    nat.ptr(loss_out), nat.ptr(conf), nat.ptr(scratch), n_steps, nat.ptr(rank_dev), nat.ptr(state_mirror),
                                    nat.cuda_stream()))
    from . import kernels as K
    K._count()


def round_tcx_p2p_many(tb, cluster_size: int, X, y, idx, params, m, v, step, active, fed_state, loss_out, conf, scratch, n_steps: int,
                       rank_dev, tail_count, state_mirror=None, host_inputs=None) -> None:
    """Multi-GPU round in one launch per rank with any number of resident clients (fedb_round_tcx_p2p_many)."""
    lib = nat.lib(required=True)
    xh, yh = host_inputs if host_inputs is not None else (None, None)
    nat.check(lib.fedb_round_tcx_p2p_many(nat.ptr(tb.d_groups), tb.n_groups, cluster_size, nat.ptr(X), X.shape[1], nat.ptr(y), nat.ptr(xh),
                                          nat.ptr(yh), nat.ptr(idx), nat.ptr(params), nat.ptr(m), nat.ptr(v), nat.ptr(step), nat.ptr(active),
                                          nat.ptr(fed_state), nat.ptr(loss_out), nat.ptr(conf), nat.ptr(scratch), n_steps, nat.ptr(rank_dev),
                                          nat.ptr(tail_count), nat.ptr(state_mirror), nat.cuda_stream()))
    from . import kernels as K
    K._count()
