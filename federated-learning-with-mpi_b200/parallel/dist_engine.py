"""Process-per-GPU federation: every rank hosts the clients placed on its GPU.

``make_local_engine`` is the single-process case (all clients on this GPU).  The multi-GPU
engine (cross-GPU aggregation inside one kernel over NVLink peer memory) lives in
``DistFedEngine``.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence

import numpy as np
import torch

from ..fed.engine import FedEngine
from ..fed.problem import FedProblem


def init_distributed(backend: Optional[str] = None) -> None:
    import torch.distributed as dist
    if dist.is_initialized():
        return
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group(backend)


def _compact(data, rows: Sequence[np.ndarray]):
    """Gather the rows the clients use into one dense matrix; shards become index ranges."""
    allr = np.concatenate(rows)
    X, y = data.X_train[allr], data.y_train[allr]
    shards, o = [], 0
    for r in rows:
        shards.append(np.arange(o, o + len(r), dtype=np.int32))
        o += len(r)
    return X, y, shards


def make_local_engine(data, rows: Sequence[np.ndarray], max_rounds: int = 300, precision: Optional[str] = None,
                      patience: int = 0, device: Optional[str] = None, hidden=(50, 200), hyper=None,
                      shared_init: bool = False) -> FedEngine:
    X, y, shards = _compact(data, rows)
    kw = {} if hyper is None else {"hyper": hyper}
    prob = FedProblem.custom(data.n_features, data.n_classes, shards, hidden=tuple(hidden), patience=patience,
                             init_per_client=not shared_init, **kw)
    return FedEngine([prob], X, y, data.X_test, data.y_test, device=device, max_rounds=max_rounds, precision=precision)


# =================================================================================================
# multi-GPU
# =================================================================================================
from dataclasses import replace as _replace
from typing import Dict

from ..ops import _native as nat
from ..ops import kernels as K
from . import p2p


def place_clients(n_clients: int, world: int) -> List[range]:
    """Contiguous blocks of clients per rank (client k -> rank k*world // n_clients)."""
    owner = [(k * world) // n_clients for k in range(n_clients)]
    return [range(owner.index(q), owner.index(q) + owner.count(q)) if q in owner else range(0, 0) for q in range(world)]


class DistFedEngine(FedEngine):
    """A :class:`FedEngine` over the clients placed on this rank, whose aggregation step is the
    cross-GPU kernel of csrc/fedb_p2p.cu instead of the single-GPU ``k_fedavg``.  Every rank
    ends each round with the same global model, the same metric history and the same
    early-stop state (computed redundantly, in the same order, from the same peer data)."""

    def __init__(self, problems: Sequence[FedProblem], X_train: np.ndarray, y_train: np.ndarray, X_test=None, y_test=None,
                 rank: int = 0, world: int = 1, max_rounds: int = 300, seed: int = 0, precision: Optional[str] = None,
                 device: Optional[str] = None, sym_ptrs: Optional[Sequence[int]] = None, timeout_s: float = 5.0):
        self.rank, self.world = rank, world
        self.global_problems = list(problems)
        local, used_rows = [], []
        self._src_rank, self._src_slot, feds_global_rows = [], [], []
        g_total = 0
        local_first = 0
        for f, p in enumerate(problems):
            blocks = place_clients(p.n_clients, world)
            mine = blocks[rank]
            if len(mine) == 0:
                raise ValueError(f"federation {f} has fewer clients ({p.n_clients}) than ranks ({world})")
            total = float(sum(len(s) for s in p.shards)) if p.weighting == "samples" else float(p.n_clients)
            local.append(_replace(p, shards=[p.shards[k] for k in mine], total_weight=total, client_offset=mine[0]))
            # where every global client's confusion slot lives
            slot_base = {q: sum(len(place_clients(pp.n_clients, world)[q]) for pp in problems[:f]) for q in range(world)}
            for q in range(world):
                for j, _k in enumerate(blocks[q]):
                    self._src_rank.append(q)
                    self._src_slot.append(slot_base[q] + j)
            feds_global_rows.append((g_total, p.n_clients))
            g_total += p.n_clients
            local_first += len(mine)
        self.n_groups_total = g_total
        # keep only the rows this rank's clients touch
        rows = np.unique(np.concatenate([s for p in local for s in p.shards]))
        remap = -np.ones(len(X_train), dtype=np.int64)
        remap[rows] = np.arange(len(rows))
        local = [_replace(p, shards=[remap[s].astype(np.int32) for s in p.shards]) for p in local]
        super().__init__(local, X_train[rows], y_train[rows], X_test, y_test, device=device or "cuda", max_rounds=max_rounds,
                         seed=seed, precision=precision)
        d = self.device
        fg = self.tb.feds.copy()
        for f, (first, n) in enumerate(feds_global_rows):
            fg[f]["first_group"], fg[f]["n_clients"] = first, n
        self.feds_global = K.upload_table(fg, d)
        self.conf_all = torch.zeros(g_total * p2p.C2, dtype=torch.int32, device=d)
        self.hist_local = torch.zeros(self.max_rounds * g_total * 4, dtype=torch.float32, device=d)
        self.src_rank = torch.tensor(self._src_rank, dtype=torch.int32, device=d)
        self.src_slot = torch.tensor(self._src_slot, dtype=torch.int32, device=d)
        self.seq = torch.zeros(1, dtype=torch.int64, device=d)
        self.bar = torch.zeros(1, dtype=torch.int32, device=d)
        self.err = torch.zeros(1, dtype=torch.int32, device=d)
        conf_slots = max(sum(len(place_clients(pp.n_clients, world)[q]) for pp in problems) for q in range(world))
        self.sym_layout = p2p.SymLayout(self.global_total, conf_slots)
        self._symbuf = None
        if sym_ptrs is None:
            self._symbuf = p2p.SymmetricBuffer(self.sym_layout.nbytes)
            sym_ptrs = self._symbuf.exchange(rank, world)
        self.sym_ptrs = list(sym_ptrs)
        self.rank_record = p2p.make_rank_record(self, rank, world, self.sym_layout, self.sym_ptrs, self.feds_global, self.conf_all,
                                                self.hist_local, self.hist_global, self.src_rank, self.src_slot, self.seq,
                                                self.bar, self.err, g_total, timeout_s,
                                                shadow=(self.wide.shadow if self.precision_name == "bf16_wide" else None))
        # FEDB200_AGG=nvls: the ranks' partial sums live in multicast-bound memory and the NVSwitch adds them (parallel/nvls.py)
        self._nvls = None
        if os.environ.get("FEDB200_AGG") == "nvls" and world > 1 and self._symbuf is not None:
            from . import nvls
            if not nvls.supported():
                raise RuntimeError("FEDB200_AGG=nvls: this device / driver does not support NVLink multicast")
            self._nvls = nvls.NvlsBuffer(3 * self.global_total * 4, rank, world)      # partial[2][total] | result[total]
            r = self.rank_record
            r["nvls_uc_partial"], r["nvls_mc_partial"] = self._nvls.uc_ptr, self._nvls.mc_ptr
            r["nvls_uc_result"] = self._nvls.uc_ptr + 2 * self.global_total * 4
            r["nvls_mc_result"] = self._nvls.mc_ptr + 2 * self.global_total * 4
        self.rank_dev = K.upload_table(self.rank_record, d)
        self.p2p_blocks = p2p.blocks_for(self.tb.max_slab)

    def _begin_is_fused(self) -> bool:
        return True   # k_fedavg_p2p ends with the next round's begin bookkeeping

    def _fused_p2p_ok(self) -> bool:
        """One federation on the tensor-core path: the whole multi-GPU round is ONE launch per rank — aggregation inside the
        round kernel (fedb_round_tc[x]_p2p for one resident client per rank, the reference's own topology C:36;
        fedb_round_tcx_p2p_many for several, split-operand kernel only)."""
        if not (self.tc_variant is not None and self.n_feds == 1 and self.world <= 8
                and os.environ.get("FEDB200_NO_FUSED_P2P") is None and getattr(self, "_nvls", None) is None):
            return False
        return self.n_groups == 1 or (self.tc_variant == "x2" and self.params_total <= self.FUSED_TAIL_MAX_FLOATS)

    def _result_is_mirrored(self) -> bool:
        return self._fused_p2p_ok()

    def pulls_host_inputs(self) -> bool:
        return self._fused_p2p_ok() and os.environ.get("FEDB200_E2E_MEMCPY") is None

    def enqueue_round(self, host_inputs=None) -> None:
        if self._fused_p2p_ok():
            from ..ops import tc
            E = max(p.hyper.local_epochs for p in self.problems)
            if self.n_groups > 1:
                tc.round_tcx_p2p_many(self.tb, self.tc_cluster, self.X, self.y, self.idx, self.params, self.m, self.v, self.step,
                                      self.active, self.fed_state, self.loss, self.conf, self.scratch, E, self.rank_dev, self.tail_count,
                                      state_mirror=self._state_pinned, host_inputs=host_inputs)
                return
            tc.round_tc_p2p(self.tb, self.tc_cluster, self.X, self.y, self.idx, self.params, self.m, self.v, self.step, self.active,
                            self.fed_state, self.loss, self.conf, self.scratch, E, self.rank_dev, state_mirror=self._state_pinned,
                            host_inputs=host_inputs, variant=self.tc_variant)
            return
        assert host_inputs is None, "host inputs are pulled by the single-launch round only"
        super().enqueue_round()

    @property
    def launches_per_round(self) -> int:
        if self._fused_p2p_ok():
            return 1
        return FedEngine.launches_per_round.fget(self)

    def enqueue_aggregate(self, write_back: bool = True, do_metrics: bool = True, begin_next: bool = True) -> None:
        p2p.launch(self.rank_dev, 1, self.p2p_blocks)   # begin_next (and the bf16 shadow of the wide path) are baked into the rank record

    @property
    def kernel_names(self) -> List[str]:
        if self._fused_p2p_ok():
            return [f"{self.tc_kernel}[train+eval+p2p-fedavg+metrics]"]
        if self.precision_name == "bf16_wide":
            return ["k_gemm_tma2[fwd|dgrad]", "k_gemm_tma2p[wgrad+optimizer]", "k_wide_head_fwd", "k_wide_head_bwd", "k_wide_bias_opt",
                    "k_fedavg_p2p[nvls multimem, +bf16 shadow]" if self._nvls is not None else "k_fedavg_p2p[+bf16 shadow]"]
        local = [f"{self.tc_kernel}[train+eval]"] if self.tc_variant is not None else ["k_fwd_bwd", "k_reduce_opt", "k_eval"]
        return local + ["k_fedavg_p2p[nvls multimem]" if self._nvls is not None else "k_fedavg_p2p"]

    def history(self) -> Dict[str, np.ndarray]:
        st = self.fed_state_host()
        R = int(st["done"].max()) if len(st) else 0
        hg = self.hist_global.cpu().numpy().reshape(self.n_feds, self.max_rounds, 4)[:, :R].astype(np.float64)
        hl = self.hist_local.cpu().numpy().reshape(self.max_rounds, self.n_groups_total, 4)[:R].astype(np.float64)
        from ..fed.metrics import METRIC_NAMES
        return {"global": hg, "local": hl, "rounds": st["done"].copy(),
                "stopped": st["stop"].astype(bool) | st["stop_pending"].astype(bool), "last": st["last"].copy(),
                "metric_names": np.asarray(METRIC_NAMES)}

    def check_errors(self) -> None:
        """The cross-GPU waits give up after ``timeout_s`` and raise a device flag instead of hanging (the reference
        aborts the job on any exception, C:203-205): surface it."""
        if int(self.err.item()) != 0:
            raise RuntimeError(f"rank {self.rank}: a peer did not arrive at the aggregation barrier within the timeout")
        super().check_errors()

    def _gather_opt_state(self, f: int):
        """Collective: every rank contributes its resident clients' rows; rank 0 returns them in global client order."""
        import torch.distributed as dist
        mine = (int(self.problems[f].client_offset), self._client_opt_state(f))
        parts = [None] * self.world
        dist.all_gather_object(parts, mine)
        if self.rank != 0:
            return None
        parts.sort(key=lambda t: t[0])
        return {k: np.concatenate([d[k] for _, d in parts]) for k in parts[0][1]}

    def close(self) -> None:
        if getattr(self, "_nvls", None) is not None:
            torch.cuda.synchronize()
            self._nvls.close()
            self._nvls = None
        if self._symbuf is not None:
            torch.cuda.synchronize()
            self._symbuf.close()
            self._symbuf = None

    @staticmethod
    def custom_income(data, rows: Sequence[np.ndarray], rank: int, world: int, max_rounds: int = 300,
                      precision: Optional[str] = None, patience: int = 0, hidden=(50, 200), hyper=None,
                      shared_init: bool = False) -> "DistFedEngine":
        """The benchmark federation: one custom-MLP federation whose client c owns ``rows[c]``."""
        kw = {} if hyper is None else {"hyper": hyper}
        prob = FedProblem.custom(data.n_features, data.n_classes, [np.asarray(r, dtype=np.int32) for r in rows], hidden=tuple(hidden),
                                 patience=patience, init_per_client=not shared_init, **kw)
        return DistFedEngine([prob], data.X_train, data.y_train, data.X_test, data.y_test, rank=rank, world=world,
                             max_rounds=max_rounds, precision=precision)


class EmulatedWorld:
    """``world`` ranks of a :class:`DistFedEngine` federation on ONE GPU: each rank has its own
    engine and its own "symmetric" buffer (a plain device tensor), and the aggregation kernel is
    launched once with gridDim.y = world (cooperative), which is the only safe way to run kernels
    that wait on each other on a single device.  Used by the single-GPU tests of the protocol."""

    def __init__(self, problems: Sequence[FedProblem], X_train, y_train, X_test=None, y_test=None, world: int = 2,
                 max_rounds: int = 64, precision: Optional[str] = None, seed: int = 0):
        dev = torch.device("cuda")
        # size the symmetric buffers first (layout depends only on the global problem list)
        global_total = sum(p.spec.slab_size for p in problems)
        conf_slots = max(sum(len(place_clients(p.n_clients, world)[q]) for p in problems) for q in range(world))
        lay = p2p.SymLayout(global_total, conf_slots)
        self.sym = [torch.zeros(lay.nbytes, dtype=torch.uint8, device=dev) for _ in range(world)]
        ptrs = [s.data_ptr() for s in self.sym]
        self.ranks = [DistFedEngine(problems, X_train, y_train, X_test, y_test, rank=q, world=world, max_rounds=max_rounds,
                                    seed=seed, precision=precision, sym_ptrs=ptrs) for q in range(world)]
        table = np.concatenate([r.rank_record for r in self.ranks])
        self.table_dev = K.upload_table(table, dev)
        self.world = world
        self.blocks = p2p.blocks_for(max(r.tb.max_slab for r in self.ranks), world)

    def enqueue_round(self) -> None:
        for r in self.ranks:
            r.prime()
            r.enqueue_local_part()
        p2p.launch(self.table_dev, self.world, self.blocks)

    def run(self, rounds: int):
        for _ in range(rounds):
            self.enqueue_round()
        torch.cuda.synchronize()
        for r in self.ranks:
            r.check_errors()
        return [r.history() for r in self.ranks]


class GlooFedEngine(FedEngine):
    """CPU multi-process federation (BASELINE config 1: "mpirun -np 3 ... on CPU, plumbing, no GPU"): every
    process hosts its share of the clients on the fp32 host path and the aggregation is a gloo
    ``all_reduce`` of the locally weighted slabs plus an ``all_gather`` of the confusion counts.  Same
    placement, arithmetic order per rank, metrics and early-stop rule as :class:`DistFedEngine`; this is the
    no-GPU stand-in for its NVLink kernel, not a hot path."""

    def __init__(self, problems: Sequence[FedProblem], X_train, y_train, X_test=None, y_test=None, rank: int = 0, world: int = 1,
                 max_rounds: int = 300, seed: int = 0, **_):
        import torch.distributed as dist
        self.rank, self.world = rank, world
        self.global_problems = list(problems)
        local = []
        for p in problems:
            mine = place_clients(p.n_clients, world)[rank]
            if len(mine) == 0:
                raise ValueError("fewer clients than ranks")
            total = float(sum(len(s) for s in p.shards)) if p.weighting == "samples" else float(p.n_clients)
            local.append(_replace(p, shards=[p.shards[k] for k in mine], total_weight=total, client_offset=mine[0]))
        super().__init__(local, X_train, y_train, X_test, y_test, device="cpu", max_rounds=max_rounds, seed=seed)
        self.n_groups_total = sum(p.n_clients for p in problems)
        self._owner = [[q for q in range(world) for _ in place_clients(p.n_clients, world)[q]] for p in problems]
        self.hist_local = torch.zeros(self.max_rounds * self.n_groups_total * 4)
        self._dist = dist
        self._group = dist.new_group(backend="gloo") if dist.get_backend() != "gloo" else None

    def enqueue_aggregate(self, write_back: bool = True, do_metrics: bool = True, begin_next: bool = False) -> None:
        from ..fed.metrics import mean_of_local, pooled
        from ..ops import reference as ref
        dist, st = self._dist, self.fed_state.numpy().view(nat.FEDSTATE_DTYPE)
        # local weighted partial sums with the single-process kernel (no write-back, no metrics) ...
        ref.fedavg(self.tb.groups, self.tb.feds, self.fed_state, self.tb.max_slab, self.params, self.global_, self.conf,
                   None, None, False, False)
        # ... summed over ranks (gloo sums in rank order for small tensors; results are identical on all ranks)
        dist.all_reduce(self.global_, group=self._group)
        confs = [torch.zeros_like(self.conf[: self.n_groups * p2p.C2]) for _ in range(self.world)]
        dist.all_gather(confs, self.conf[: self.n_groups * p2p.C2].contiguous(), group=self._group)
        g_base = 0
        for f, p in enumerate(self.global_problems):
            if st[f]["stop"]:
                g_base += p.n_clients
                continue
            P, C = p.spec.slab_size, p.spec.n_classes
            if write_back:
                for gi in self.group_of[f]:
                    o = int(self.tb.groups[gi]["slab_off"])
                    self.params[o: o + P] = self.global_slab(f)
            if do_metrics:
                # confusion matrices of ALL clients of the federation, in global client order
                cms, seen = [], {q: 0 for q in range(self.world)}
                base = {q: sum(len(place_clients(pp.n_clients, self.world)[q]) for pp in self.global_problems[:f]) for q in range(self.world)}
                for q in self._owner[f]:
                    cm = confs[q].view(-1, nat.MAX_CLASSES, nat.MAX_CLASSES)[base[q] + seen[q]].numpy()[:C, :C]
                    seen[q] += 1
                    cms.append(cm)
                r = int(st[f]["round"]) - 1
                from ..fed.metrics import metrics_from_confusion
                loc = np.array([[metrics_from_confusion(c)[k] for k in ("accuracy", "precision", "recall", "f1")] for c in cms])
                glob = pooled(cms) if p.metric_mode == "pooled" else mean_of_local(cms)
                gv = np.array([glob[k] for k in ("accuracy", "precision", "recall", "f1")])
                if 0 <= r < self.max_rounds:
                    self.hist_local.view(self.max_rounds, self.n_groups_total, 4)[r, g_base: g_base + p.n_clients] = torch.from_numpy(loc).float()
                    self.hist_global.view(self.n_feds, self.max_rounds, 4)[f, r] = torch.from_numpy(gv).float()
                st[f]["last"] = gv
                if p.patience > 0:
                    # the reference's rule (C:181-192) as restated once in fed/metrics.py, driven from the device-layout state
                    from ..fed.metrics import METRIC_NAMES, EarlyStop
                    es = EarlyStop(patience=p.patience, atol=p.atol, rtol=p.rtol, count=int(st[f]["count"]),
                                   anchor=st[f]["anchor"].copy() if st[f]["has_anchor"] else None)
                    if es.update(dict(zip(METRIC_NAMES, gv))):
                        st[f]["stop_pending"] = 1
                    st[f]["anchor"], st[f]["has_anchor"], st[f]["count"] = es.anchor, 1, es.count
                st[f]["done"] = st[f]["round"]
            g_base += p.n_clients

    _gather_opt_state = DistFedEngine._gather_opt_state

    def history(self) -> Dict[str, np.ndarray]:
        st = self.fed_state_host()
        R = int(st["done"].max()) if len(st) else 0
        from ..fed.metrics import METRIC_NAMES
        return {"global": self.hist_global.numpy().reshape(self.n_feds, self.max_rounds, 4)[:, :R].astype(np.float64),
                "local": self.hist_local.numpy().reshape(self.max_rounds, self.n_groups_total, 4)[:R].astype(np.float64),
                "rounds": st["done"].copy(), "stopped": st["stop"].astype(bool) | st["stop_pending"].astype(bool),
                "last": st["last"].copy(), "metric_names": np.asarray(METRIC_NAMES)}
