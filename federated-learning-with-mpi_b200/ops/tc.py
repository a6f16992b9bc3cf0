"""Tensor-core (tcgen05) support code on the Python side: operand-panel layout helpers and the
hardware probe used by tests/test_gpu_tc.py to validate the descriptor conventions."""
from __future__ import annotations

import ctypes
from typing import List, Sequence

import numpy as np
import torch

from . import _native as nat

_P, _I32 = ctypes.c_void_p, ctypes.c_int32
nat.OPTIONAL_SIGS["fedb_tc_probe"] = ([_P, _P, ctypes.POINTER(_I32), _P, _P], _I32)
nat.OPTIONAL_SIGS["fedb_tc_probe_args_size"] = ([], _I32)

PANEL_COLS = 64


def panel_offsets(rows: int) -> np.ndarray:
    """int64[rows, 64]: byte offset of element (r, c) in a 128B-swizzled bf16 panel
    (csrc/tc_common.cuh: panel_off)."""
    r = np.arange(rows)[:, None]
    c = np.arange(PANEL_COLS)[None, :]
    return r * 128 + (((c >> 3) ^ (r & 7)) << 4) + ((c & 7) << 1)


def to_panels(mat: torch.Tensor) -> torch.Tensor:
    """[rows, cols] float -> uint8 image of ceil(cols/64) consecutive panels (bf16, swizzled).
    ``rows`` must be a multiple of 8."""
    rows, cols = mat.shape
    assert rows % 8 == 0
    npan = (cols + PANEL_COLS - 1) // PANEL_COLS
    padded = torch.zeros(rows, npan * PANEL_COLS, dtype=torch.float32)
    padded[:, :cols] = mat
    bits = padded.to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    img = np.zeros(npan * rows * 128, dtype=np.uint8)
    off = panel_offsets(rows)
    img16 = img.view(np.uint16)
    for p in range(npan):
        img16[(p * rows * 128 + off) // 2] = bits[:, p * 64: (p + 1) * 64]
    return torch.from_numpy(img)


def idesc_bf16(M: int, N: int, a_mn: bool = False, b_mn: bool = False) -> int:
    return (1 << 4) | (1 << 7) | (1 << 10) | (int(a_mn) << 15) | (int(b_mn) << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def probe(a_img: torch.Tensor, b_img: torch.Tensor, N: int, a_steps: Sequence[int], b_steps: Sequence[int],
          a_lbo: int, a_sbo: int, b_lbo: int, b_sbo: int, a_mn: bool = False, b_mn: bool = False,
          repeat: int = 1) -> torch.Tensor:
    """Run a chain of M=128 bf16 MMAs on the given smem images; returns D as fp32 [128, N]."""
    lib = nat.lib(required=True)
    assert lib.fedb_tc_probe_args_size() == 74 * 4
    args = np.zeros(74, dtype=np.int32)
    args[0], args[1] = a_img.numel(), b_img.numel()
    args[2:6] = [a_lbo, a_sbo, b_lbo, b_sbo]
    args[6] = idesc_bf16(128, N, a_mn, b_mn)
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
