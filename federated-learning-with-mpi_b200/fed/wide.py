"""Layer-wise local training for MLPs too wide for the whole-network kernels (BASELINE config 5:
4 x 4096 hidden units, 128 clients per GPU).  Hidden layers run on the grouped tcgen05 GEMM
(csrc/fedb_gemm.cu) with fused bias+ReLU / ReLU-mask / weight-gradient+optimizer epilogues; the
narrow head and the bias gradients run on CUDA cores (csrc/fedb_wide.cu).  Master weights, Adam
moments and the optimizer are fp32; GEMM operands are bf16 (activations, and a bf16 shadow of the
weights that every optimizer epilogue refreshes in place).
"""
from __future__ import annotations

from typing import List

import torch

from ..ops import _native as nat
from ..ops import gemm as gm
from ..ops import kernels as K


class WideLocal:
    """Buffers + launch sequences of the wide path for one :class:`FedEngine` (all groups must share
    one model shape and one shard size that is a multiple of 128)."""

    def __init__(self, eng):
        self.eng = eng
        g0 = eng.tb.groups[0]
        self.L = int(g0["n_layers"])
        self.dims = [int(d) for d in g0["dims"][: self.L + 1]]
        self.P = int(g0["slab_size"])
        rows = {int(g["n_rows"]) for g in eng.tb.groups}
        shapes = {tuple(int(d) for d in g["dims"]) for g in eng.tb.groups}
        if len(shapes) != 1 or len(rows) != 1:
            raise ValueError("the wide path needs one model shape and equal shard sizes for all resident clients")
        self.B = rows.pop()
        if self.B % 128 or any(d % 128 for d in self.dims[1:-1]) or int(g0["act"]) != nat.ACT_IDS["relu"] or self.L < 2:
            raise ValueError("the wide path needs ReLU, >= 1 hidden layer, shard size % 128 == 0 and hidden widths % 128 == 0")
        if any(int(g["batch_size"]) > 0 for g in eng.tb.groups):
            raise ValueError("the wide path trains full-batch")
        self.G = eng.n_groups
        dev = eng.device
        spec = eng.problems[0].spec
        self.w_off, self.b_off = spec.w_offsets, spec.b_offsets
        self.ld = [spec.ld(l) for l in range(self.L)]
        bf = torch.bfloat16
        # activations: A[0] = inputs (padded to ld[0]), A[l] = output of hidden layer l-1
        self.A: List[torch.Tensor] = [torch.zeros(self.G * self.B * self.ld[0], dtype=bf, device=dev)]
        for l in range(1, self.L):
            self.A.append(torch.empty(self.G * self.B * self.dims[l], dtype=bf, device=dev))
        maxd = max(self.dims[1:-1])
        self.dZ = [torch.empty(self.G * self.B * maxd, dtype=bf, device=dev) for _ in range(2)]
        self.dz_head = torch.zeros(self.G * self.B * nat.MAX_CLASSES, dtype=torch.float32, device=dev)
        self.loss_rows = torch.zeros(self.G * self.B, dtype=torch.float32, device=dev)
        self.shadow = torch.empty(eng.params_total, dtype=bf, device=dev)
        self._lib = nat.lib(required=True)
        self.gather_inputs()
        self.refresh_shadow()

    # -- helpers -------------------------------------------------------------------------------------
    def _wptr(self, t: torch.Tensor, l: int) -> int:
        """Address of layer l's weight of group 0 inside an arena tensor (groups are P apart)."""
        return t.data_ptr() + (int(self.eng.tb.groups[0]["slab_off"]) + self.w_off[l]) * t.element_size()

    def _bptr(self, t: torch.Tensor, l: int) -> int:
        return t.data_ptr() + (int(self.eng.tb.groups[0]["slab_off"]) + self.b_off[l]) * t.element_size()

    def gather_inputs(self) -> None:
        e = self.eng
        nat.check(self._lib.fedb_wide_gather_x(nat.ptr(e.tb.d_groups), self.G, nat.ptr(e.X), e.X.shape[1], nat.ptr(e.idx),
                                               nat.ptr(self.A[0]), self.B, self.ld[0], nat.cuda_stream()))
        K._count()

    def refresh_shadow(self) -> None:
        e = self.eng
        nat.check(self._lib.fedb_wide_shadow(nat.ptr(e.params), nat.ptr(self.shadow), e.params_total, nat.cuda_stream()))
        K._count()

    # -- forward through the hidden layers -------------------------------------------------------------
    def enqueue_forward(self) -> None:
        e, G, B = self.eng, self.G, self.B
        for l in range(self.L - 1):
            din, dout = (self.ld[0] if l == 0 else self.dims[l]), self.dims[l + 1]
            gm.gemm(G, B, dout, din, self.A[l], B * din, din, False, self._wptr(self.shadow, l), self.P, self.ld[l], False,
                    gm.FWD_BIAS_RELU, C=self.A[l + 1], c_stride=B * dout, ldc=dout, bias=self._bptr(e.params, l), bias_stride=self.P)

    def enqueue_eval(self) -> None:
        e = self.eng
        self.enqueue_forward()
        nat.check(self._lib.fedb_wide_head_fwd(nat.ptr(e.tb.d_groups), self.G, nat.ptr(e.params), nat.ptr(self.A[self.L - 1]), self.B,
                                               self.dims[self.L - 1], nat.ptr(e.y), nat.ptr(e.idx), nat.ptr(self.dz_head),
                                               nat.ptr(self.loss_rows), nat.ptr(e.conf), nat.ptr(e.step), nat.ptr(e.fed_state), 1,
                                               nat.cuda_stream()))
        K._count()

    # -- one full-batch optimizer step of every client ---------------------------------------------------
    def enqueue_step(self) -> None:
        e, G, B, L = self.eng, self.G, self.B, self.L
        lib, st = self._lib, nat.cuda_stream
        self.enqueue_forward()
        dlast = self.dims[L - 1]
        nat.check(lib.fedb_wide_head_fwd(nat.ptr(e.tb.d_groups), G, nat.ptr(e.params), nat.ptr(self.A[L - 1]), B, dlast, nat.ptr(e.y),
                                         nat.ptr(e.idx), nat.ptr(self.dz_head), nat.ptr(self.loss_rows), nat.ptr(e.conf),
                                         nat.ptr(e.step), nat.ptr(e.fed_state), 0, st()))
        cur = 0
        nat.check(lib.fedb_wide_head_bwd(nat.ptr(e.tb.d_groups), G, nat.ptr(e.params), nat.ptr(e.m), nat.ptr(e.v), nat.ptr(self.shadow),
                                         nat.ptr(self.A[L - 1]), B, dlast, nat.ptr(self.dz_head), nat.ptr(self.loss_rows),
                                         nat.ptr(self.dZ[cur]), nat.ptr(e.step), nat.ptr(e.fed_state), nat.ptr(e.loss), st()))
        K._count(2)
        for l in range(L - 2, -1, -1):          # hidden layer l: dZ[cur] is d(loss)/d(pre-activation of its output)
            din, dout = (self.ld[0] if l == 0 else self.dims[l]), self.dims[l + 1]
            nat.check(lib.fedb_wide_bias_opt(nat.ptr(e.tb.d_groups), G, nat.ptr(e.params), nat.ptr(e.m), nat.ptr(e.v),
                                             nat.ptr(self.dZ[cur]), B, dout, l, nat.ptr(e.step), nat.ptr(e.fed_state), st()))
            K._count()
            if l > 0:                            # gradient wrt the previous layer's output, with W_l as it is BEFORE its update
                gm.gemm(G, B, din, dout, self.dZ[cur], B * dout, dout, False, self._wptr(self.shadow, l), self.P, self.ld[l], True,
                        gm.DGRAD_MASK, C=self.dZ[cur ^ 1], c_stride=B * din, ldc=din, act=self.A[l], act_stride=B * din, ldact=din)
            # W_l <- opt(W_l, dZ^T A_l) inside the GEMM epilogue (fp32 master + moments + bf16 shadow)
            gm.gemm(G, dout, din, B, self.dZ[cur], B * dout, dout, True, self.A[l], B * din, din, True, gm.WGRAD_OPT,
                    W=self._wptr(e.params, l), Mom=self._wptr(e.m, l), Vel=self._wptr(e.v, l), Wsh=self._wptr(self.shadow, l),
                    w_stride=self.P, ldw=self.ld[l], l2_scale=1.0 / B, groups=e.tb.d_groups, step=e.step)
            cur ^= 1

    @property
    def launches_per_step(self) -> int:
        L = self.L
        return (L - 1) + 2 + (L - 1) * 2 + (L - 2)
