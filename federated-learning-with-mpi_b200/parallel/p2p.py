"""Cross-GPU FedAvg over NVLink peer memory: symmetric buffers, rank descriptors, launches
(kernel: csrc/fedb_p2p.cu).  ``torch.distributed`` is used only to exchange the 64-byte CUDA
IPC handles at set-up (and as the barrier around benchmarks); the per-round aggregation is a
single kernel per rank that reads and writes peers directly.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

from ..ops import _native as nat
from ..ops import kernels as K

C2 = nat.MAX_CLASSES ** 2


@dataclass
class SymLayout:
    global_total: int   # floats per parity
    conf_slots: int     # confusion slots per parity (max local clients over ranks)

    @property
    def partial_off(self) -> int:
        return 0

    @property
    def conf_off(self) -> int:
        return (2 * self.global_total * 4 + 255) // 256 * 256

    @property
    def flag_off(self) -> int:
        return (self.conf_off + 2 * 8 * self.conf_slots * C2 * 4 + 255) // 256 * 256   # areas (parity, source rank)

    @property
    def nbytes(self) -> int:
        return self.flag_off + 256


class SymmetricBuffer:
    """A zero-initialised device buffer of this rank plus its peers' mappings."""

    def __init__(self, nbytes: int):
        self.lib = nat.lib(required=True)
        p = ctypes.c_void_p()
        nat.check(self.lib.fedb_symm_alloc(nbytes, ctypes.byref(p)))
        self.ptr = p.value
        self.nbytes = nbytes
        self.peers: List[int] = []
        self._opened: List[int] = []

    def handle(self) -> torch.Tensor:
        h = (ctypes.c_uint8 * 64)()
        nat.check(self.lib.fedb_ipc_get_handle(self.ptr, ctypes.cast(h, ctypes.c_void_p)))
        return torch.tensor(list(h), dtype=torch.uint8)

    def exchange(self, rank: int, world: int) -> List[int]:
        """all_gather the IPC handles and map every peer's buffer; returns base pointers by rank."""
        import torch.distributed as dist
        mine = self.handle().cuda()
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine)
        self.peers = []
        for q in range(world):
            if q == rank:
                self.peers.append(self.ptr)
                continue
            buf = (ctypes.c_uint8 * 64)(*allh[q].cpu().tolist())
            out = ctypes.c_void_p()
            nat.check(self.lib.fedb_ipc_open_handle(ctypes.cast(buf, ctypes.c_void_p), ctypes.byref(out)))
            self.peers.append(out.value)
            self._opened.append(out.value)
        dist.barrier()
        return self.peers

    def close(self) -> None:
        for p in self._opened:
            self.lib.fedb_ipc_close_handle(p)
        self._opened = []
        if self.ptr:
            self.lib.fedb_symm_free(self.ptr)
            self.ptr = 0


def make_rank_record(eng, rank: int, world: int, lay: SymLayout, sym_ptrs: Sequence[int], feds_global: torch.Tensor,
                     conf_all: torch.Tensor, hist_local: torch.Tensor, hist_global: torch.Tensor, src_rank: torch.Tensor,
                     src_slot: torch.Tensor, seq: torch.Tensor, bar: torch.Tensor, err: torch.Tensor, n_groups_total: int,
                     timeout_s: float = 5.0, write_back: bool = True, do_metrics: bool = True, begin_next: bool = True,
                     shadow: Optional[torch.Tensor] = None) -> np.ndarray:
    if eng.n_feds > 256:
        raise ValueError("the cross-GPU aggregation kernel handles at most 256 federations per launch")
    r = np.zeros(1, dtype=nat.P2PRANK_DTYPE)
    r["rank"], r["world"], r["n_feds"], r["n_groups_local"] = rank, world, eng.n_feds, eng.n_groups
    r["n_groups_total"], r["conf_slots"], r["global_total"] = n_groups_total, lay.conf_slots, lay.global_total
    r["groups"], r["feds"], r["feds_global"] = eng.tb.d_groups.data_ptr(), eng.tb.d_feds.data_ptr(), feds_global.data_ptr()
    r["fed_state"], r["params"], r["global"] = eng.fed_state.data_ptr(), eng.params.data_ptr(), eng.global_.data_ptr()
    r["conf"], r["conf_all"] = eng.conf.data_ptr(), conf_all.data_ptr()
    r["hist_local"], r["hist_global"] = hist_local.data_ptr(), hist_global.data_ptr()
    r["conf_src_rank"], r["conf_src_slot"] = src_rank.data_ptr(), src_slot.data_ptr()
    sym = np.zeros(8, dtype=np.uint64)
    sym[: len(sym_ptrs)] = sym_ptrs
    r["sym"][0] = sym
    r["sym_partial_off"], r["sym_conf_off"], r["sym_flag_off"] = lay.partial_off, lay.conf_off, lay.flag_off
    r["seq"], r["bar"], r["err"] = seq.data_ptr(), bar.data_ptr(), err.data_ptr()
    r["timeout_cycles"] = int(timeout_s * 1.9e9)
    r["write_back"], r["do_metrics"], r["begin_next"] = int(write_back), int(do_metrics), int(begin_next)
    r["shadow"] = 0 if shadow is None else shadow.data_ptr()
    return r


def launch(ranks_dev: torch.Tensor, n_ranks_here: int, blocks_x: int) -> None:
    lib = nat.lib(required=True)
    assert lib.fedb_p2p_rank_size() == nat.P2PRANK_DTYPE.itemsize, (lib.fedb_p2p_rank_size(), nat.P2PRANK_DTYPE.itemsize)
    nat.check(lib.fedb_fedavg_p2p(ranks_dev.data_ptr(), n_ranks_here, blocks_x, nat.cuda_stream()))
    K._count()


def blocks_for(max_slab: int, n_ranks_here: int = 1, sm_count: int = 148) -> int:
    """Enough 256-thread blocks to cover the largest slab once, bounded so the cooperative grid
    (blocks_x * emulated ranks) stays co-resident."""
    want = max(1, (max_slab // 4 + 255) // 256)
    return int(max(1, min(want, (4 * sm_count) // max(1, n_ranks_here))))   # 4 co-resident blocks per SM (launch bounds)
