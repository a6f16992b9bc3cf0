"""Launchers for the native kernels (one Python function per C entry point).

Each function takes the device-resident tables of :class:`DeviceTables` plus torch tensors
and enqueues exactly one kernel on the current CUDA stream (graph-capturable: no allocation,
no synchronisation).  With CPU tensors it runs the fp32 restatement in ops/reference.py on
the same buffers instead — that is the CPU test path, not a second backend: on a machine
with a GPU the native library is mandatory (ops/_native.lib raises if it is missing).
"""
from __future__ import annotations

from dataclasses import dataclass

import os

import numpy as np
import torch

from . import _native as nat
from . import reference as ref

# launches enqueued through this module since import (bench.py reports the delta)
launch_count = 0


def _count(n: int = 1) -> None:
    global launch_count
    launch_count += n


@dataclass
class DeviceTables:
    """Host (NumPy structured) and device (uint8 tensor) copies of the group / federation
    descriptor tables."""

    groups: np.ndarray          # GROUP_DTYPE[G]
    feds: np.ndarray            # FED_DTYPE[F]
    d_groups: torch.Tensor
    d_feds: torch.Tensor
    max_slab: int
    smem_bytes: int

    @property
    def n_groups(self) -> int:
        return len(self.groups)

    @property
    def n_feds(self) -> int:
        return len(self.feds)

    @staticmethod
    def build(groups: np.ndarray, feds: np.ndarray, device) -> "DeviceTables":
        max_slab = int(groups["slab_size"].max())
        smem = 0
        for g in groups:
            try:
                smem = max(smem, simt_smem_bytes(g["dims"], int(g["n_layers"])))
            except ValueError:
                smem = -1   # too wide for the whole-network SIMT kernels: only the layer-wise (wide) path can run it
                break
        d_groups = torch.from_numpy(groups.view(np.uint8).copy()).to(device)
        d_feds = torch.from_numpy(feds.view(np.uint8).copy()).to(device)
        return DeviceTables(groups, feds, d_groups, d_feds, max_slab, smem)

    def refresh(self) -> None:
        """Re-upload the host tables after editing hyper-parameters in place."""
        self.d_groups.copy_(torch.from_numpy(self.groups.view(np.uint8).copy()))
        self.d_feds.copy_(torch.from_numpy(self.feds.view(np.uint8).copy()))


def simt_smem_bytes(dims, n_layers: int) -> int:
    """Dynamic shared memory of the whole-network SIMT kernels for this MLP (mirrors smem_plan / fedb_simt_smem_bytes in
    csrc/fedb_simt.cu; tests/test_models.py checks the two against each other when the library is present)."""
    d = [int(x) for x in dims[: n_layers + 1]]
    ru4 = lambda x: (x + 3) // 4 * 4
    # the backward pass ping-pongs between two dz buffers: buffer 0 holds layers L, L-2, ..., buffer 1 layers L-1, L-3, ...
    w = [4, 4]
    for par, l in enumerate(range(n_layers, 0, -1)):
        w[par & 1] = max(w[par & 1], ru4(d[l]))
    total = (sum(ru4(x) for x in d) + w[0] + w[1]) * 33 + 64
    b = total * 4
    n_w = sum(d[l + 1] * ru4(d[l]) + ru4(d[l + 1]) for l in range(n_layers))
    cap = 227 * 1024 - 64          # the kernels' few static __shared__ variables count against the same limit
    if b + n_w * 4 <= cap:
        return b + n_w * 4  # room to stage the client's weights in shared memory too
    if b > cap:
        raise ValueError(f"MLP {d} needs {b} bytes of shared memory per CTA in the SIMT kernels (max {cap}); "
                         "use the wide (layer-wise tcgen05 GEMM) path")
    # weights that do not fit stream through whatever the activations leave (csrc/fedb_simt.cu, layer_*_chunked):
    # ask for everything.  FEDB200_SIMT_STREAM=0 keeps the old behaviour (weights read from global memory in the layer loops).
    return b if os.environ.get("FEDB200_SIMT_STREAM", "1") == "0" else cap


def upload_table(arr: np.ndarray, device) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(arr).view(np.uint8).copy()).to(device)


def _cuda(t: torch.Tensor) -> bool:
    return t.is_cuda


def _L():
    l = nat.lib(required=True)
    return l


def init_params(tb: DeviceTables, params, streams: torch.Tensor, kind: str, seed: int) -> None:
    k = nat.INIT_IDS[kind]
    if _cuda(params):
        nat.check(_L().fedb_init_params(nat.ptr(tb.d_groups), tb.n_groups, tb.max_slab, nat.ptr(params), nat.ptr(streams),
                                        k, seed & (2 ** 64 - 1), nat.cuda_stream()))
        _count()
    else:
        ref.init_params(tb.groups, tb.max_slab, params, streams.numpy(), k, seed)


def round_begin(tb: DeviceTables, fed_state, conf) -> None:
    if _cuda(conf):
        nat.check(_L().fedb_round_begin(nat.ptr(tb.d_feds), nat.ptr(fed_state), tb.n_feds, nat.ptr(conf),
                                        conf.numel() // (nat.MAX_CLASSES ** 2), nat.cuda_stream()))
        _count()
    else:
        ref.round_begin(tb.feds, fed_state, conf)


def shuffle(tb: DeviceTables, idx_in, idx_out, epoch, active, seed: int) -> None:
    if _cuda(idx_in):
        nat.check(_L().fedb_shuffle(nat.ptr(tb.d_groups), tb.n_groups, nat.ptr(idx_in), nat.ptr(idx_out), nat.ptr(epoch),
                                    nat.ptr(active), seed & (2 ** 64 - 1), nat.cuda_stream()))
        _count()
    else:
        ref.shuffle(tb.groups, idx_in, idx_out, epoch, active, seed)


def fwd_bwd(tb: DeviceTables, tiles, n_tiles: int, params, X, y, idx, batch: int, scratch, step, active, fed_state) -> None:
    if _cuda(params):
        nat.check(_L().fedb_fwd_bwd(nat.ptr(tb.d_groups), nat.ptr(tiles), n_tiles, nat.ptr(params), nat.ptr(X), X.shape[1],
                                    nat.ptr(y), nat.ptr(idx), batch, nat.ptr(scratch), nat.ptr(step), nat.ptr(active),
                                    nat.ptr(fed_state), tb.smem_bytes, nat.cuda_stream()))
        _count()
    else:
        ref.fwd_bwd(tb.groups, tiles, params, X, X.shape[1], y, idx, batch, scratch, step, active, fed_state)


def reduce_opt(tb: DeviceTables, params, m, v, scratch, batch: int, step, active, fed_state, loss_out,
               batch_stats=None, max_batches: int = 0) -> None:
    if _cuda(params):
        nat.check(_L().fedb_reduce_opt(nat.ptr(tb.d_groups), tb.n_groups, tb.max_slab, nat.ptr(params), nat.ptr(m), nat.ptr(v),
                                       nat.ptr(scratch), batch, nat.ptr(step), nat.ptr(active), nat.ptr(fed_state),
                                       nat.ptr(loss_out), nat.ptr(batch_stats), max_batches, nat.cuda_stream()))
        _count()
    else:
        ref.reduce_opt(tb.groups, tb.max_slab, params, m, v, scratch, batch, step, active, fed_state, loss_out,
                       batch_stats, max_batches)


def epoch_end(tb: DeviceTables, batch_stats, max_batches: int, tol: float, n_iter_no_change: int, max_iter: int,
              active, epoch, no_improve, best_loss, last_loss) -> None:
    if _cuda(batch_stats):
        nat.check(_L().fedb_epoch_end(nat.ptr(tb.d_groups), tb.n_groups, nat.ptr(batch_stats), max_batches, tb.max_slab,
                                      float(tol), n_iter_no_change, max_iter, nat.ptr(active), nat.ptr(epoch),
                                      nat.ptr(no_improve), nat.ptr(best_loss), nat.ptr(last_loss), nat.cuda_stream()))
        _count()
    else:
        ref.epoch_end(tb.groups, batch_stats, max_batches, tb.max_slab, tol, n_iter_no_change, max_iter, active, epoch,
                      no_improve, best_loss, last_loss)


def fit_cluster(tb: DeviceTables, cluster_size: int, params, m, v, X, y, idx_in, idx_epoch, scratch, step, active, fed_state, epoch,
                no_improve, best_loss, last_loss, loss_out, batch_stats, n_batches: int, tol: float, n_iter_no_change: int,
                max_iter: int, max_epochs: int, shuffle_rows: bool, seed: int) -> None:
    """CUDA only: every client's whole local fit in one launch (fedb_fit_cluster) — the device-side equivalent of looping
    ``shuffle`` + ``n_batches`` x (``fwd_bwd`` + ``reduce_opt``) + ``epoch_end`` for ``max_epochs`` epochs."""
    nat.check(_L().fedb_fit_cluster(nat.ptr(tb.d_groups), tb.n_groups, tb.max_slab, cluster_size, nat.ptr(params), nat.ptr(m), nat.ptr(v),
                                    nat.ptr(X), X.shape[1], nat.ptr(y), nat.ptr(idx_in), nat.ptr(idx_epoch), nat.ptr(scratch),
                                    nat.ptr(step), nat.ptr(active), nat.ptr(fed_state), nat.ptr(epoch), nat.ptr(no_improve),
                                    nat.ptr(best_loss), nat.ptr(last_loss), nat.ptr(loss_out), nat.ptr(batch_stats), n_batches,
                                    n_batches, float(tol), n_iter_no_change, max_iter, max_epochs, int(shuffle_rows),
                                    seed & (2 ** 64 - 1), tb.smem_bytes, nat.cuda_stream()))
    _count()


def batch_stats_stride(max_slab: int) -> int:
    return ref.batch_stats_stride(max_slab)


def evaluate(tb: DeviceTables, jobs_np: np.ndarray, d_jobs, tiles, n_tiles: int, params, global_, X, y, idx, conf,
             fed_state) -> None:
    if _cuda(params):
        nat.check(_L().fedb_eval(nat.ptr(tb.d_groups), nat.ptr(d_jobs), nat.ptr(tiles), n_tiles, nat.ptr(params),
                                 nat.ptr(global_), nat.ptr(X), X.shape[1], nat.ptr(y), nat.ptr(idx), nat.ptr(conf),
                                 nat.ptr(fed_state), tb.smem_bytes, nat.cuda_stream()))
        _count()
    else:
        ref.evaluate(tb.groups, jobs_np, tiles, params, global_, X, X.shape[1], y, idx, conf, fed_state)


def fedavg(tb: DeviceTables, fed_state, params, global_, conf, hist_local, hist_global, write_back: bool = True,
           do_metrics: bool = True, begin_next: bool = False, shadow=None) -> None:
    if _cuda(params):
        nat.check(_L().fedb_fedavg(nat.ptr(tb.d_groups), nat.ptr(tb.d_feds), nat.ptr(fed_state), tb.n_feds, tb.n_groups,
                                   tb.max_slab, nat.ptr(params), nat.ptr(global_), nat.ptr(conf), nat.ptr(hist_local),
                                   nat.ptr(hist_global), int(write_back), int(do_metrics), int(begin_next), nat.ptr(shadow),
                                   nat.cuda_stream()))
        _count()
    else:
        ref.fedavg(tb.groups, tb.feds, fed_state, tb.max_slab, params, global_, conf, hist_local, hist_global,
                   write_back, do_metrics, begin_next)


def fitx_eligible(groups: np.ndarray) -> np.ndarray:
    """bool[n_groups]: which clients the tensor-core fit kernel (csrc/fedb_fitx.cu) can train: 2 hidden ReLU layers,
    in <= 15, h1 <= 63, h2 <= 512, out <= 8."""
    lib = nat.lib(required=True)
    out = np.zeros(len(groups), dtype=bool)
    for i, g in enumerate(groups):
        d = g["dims"]
        out[i] = (int(g["n_layers"]) == 3 and int(g["act"]) == nat.ACT_IDS["relu"]
                  and bool(lib.fedb_fitx_fits(int(d[0]), int(d[1]), int(d[2]), int(d[3]))))
    return out


def fit_tcx(tb: DeviceTables, params, m, v, X, y, idx_in, idx_epoch, step, active, fed_state, epoch, no_improve, best_loss, last_loss,
            loss_out, tol: float, n_iter_no_change: int, max_iter: int, max_epochs: int, shuffle: bool, seed: int, prof=None) -> None:
    """The whole local fit of every active client in one launch on the tensor cores (fedb_fit_tcx)."""
    nat.check(_L().fedb_fit_tcx(nat.ptr(tb.d_groups), tb.n_groups, nat.ptr(params), nat.ptr(m), nat.ptr(v), nat.ptr(X), X.shape[1],
                                nat.ptr(y), nat.ptr(idx_in), nat.ptr(idx_epoch), nat.ptr(step), nat.ptr(active), nat.ptr(fed_state),
                                nat.ptr(epoch), nat.ptr(no_improve), nat.ptr(best_loss), nat.ptr(last_loss), nat.ptr(loss_out),
                                float(tol), int(n_iter_no_change), int(max_iter), int(max_epochs), int(bool(shuffle)),
                                seed & (2 ** 64 - 1), nat.ptr(prof), nat.cuda_stream()))
    _count()
