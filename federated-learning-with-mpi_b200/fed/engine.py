"""Round engine: many simulated clients (and many federations) per process, all state resident
on the device, one short fixed kernel sequence per round.

What the reference does per round with ~10 collectives, 4 host<->device dataset copies, 12
weight copies and 4 sklearn metric calls per rank (C:130-201) is here:

    round_begin -> E x [fwd_bwd -> reduce_opt] -> eval -> fedavg(+metrics +early stop)

enqueued on one stream, captured once into a CUDA graph and replayed.  Nothing returns to the
host inside a round; the stop flag (C:132-136, C:181-195) lives on the device and kernels of a
stopped federation exit immediately, so the host only polls it every few rounds.  The
sklearn-style "fit until the loss stalls" local training (S:101, H:91) is `fit_local`.
"""
from __future__ import annotations

import os

import math
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from ..models import init as minit
from ..ops import _native as nat
from ..ops import kernels as K
from ..utils import trace
from .metrics import METRIC_NAMES
from .problem import FedProblem

TILE = nat.TILE_ROWS


def _ru(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class FedEngine:
    def __init__(self, problems: Sequence[FedProblem], X_train: np.ndarray, y_train: np.ndarray,
                 X_test: Optional[np.ndarray] = None, y_test: Optional[np.ndarray] = None,
                 device: Optional[str] = None, max_rounds: int = 300, seed: int = 0, precision: Optional[str] = None):
        self.problems = list(problems)
        self._precision_req = precision
        if device is None:
            device = "cuda" if torch.cuda.is_available() else "cpu"
        self.device = torch.device(device)
        if self.device.type == "cuda":
            nat.lib(required=True)  # fail loudly rather than run a fallback on a GPU box
        self.max_rounds = int(max_rounds)
        self.seed = int(seed)
        # FEDB200_FIT=launches | cluster pins the sklearn-style fit of THIS engine to the CUDA-core kernels (read once, here)
        self._fit_mode = os.environ.get("FEDB200_FIT", "")
        self.n_train = len(X_train)
        self.n_test = 0 if X_test is None else len(X_test)
        nf = X_train.shape[1]
        self.ldx = max(16, _ru(nf, 4))
        Xall = np.zeros((self.n_train + self.n_test, self.ldx), dtype=np.float32)
        Xall[: self.n_train, :nf] = X_train
        yall = np.zeros(self.n_train + self.n_test, dtype=np.int32)
        yall[: self.n_train] = y_train
        if self.n_test:
            Xall[self.n_train:, :nf] = X_test
            yall[self.n_train:] = y_test
        self.X = torch.from_numpy(Xall).to(self.device)
        self.y = torch.from_numpy(yall).to(self.device)
        self._build_tables()
        self._alloc_state()
        self._graphs: Dict[str, object] = {}
        self._primed = False
        self._select_precision()
        self.init_params()

    def _select_precision(self) -> None:
        """'fp32' = SIMT kernels; 'tc32' = fused tcgen05 round kernel with split fp16 operands (three MMAs per
        GEMM, fp32-grade results: csrc/fedb_tcx.cu); 'bf16' = the same round with plain bf16 operands
        (csrc/fedb_tc.cu; faster, 8 significant bits per operand); fp32 accumulation, master weights and
        optimizer in all of them.  Default: tc32 when every model is eligible and training is full-batch on a
        GPU, else fp32."""
        from ..ops import tc
        full_batch = all(p.hyper.batch_size <= 0 and p.hyper.max_iter <= 0 for p in self.problems)
        ok = self.device.type == "cuda" and full_batch and tc.tc_eligible(self.tb.groups)
        ok32 = self.device.type == "cuda" and full_batch and tc.tc_eligible(self.tb.groups, "x2")
        req = self._precision_req
        self.wide = None
        self.tc_variant = None
        if req in (None, "auto"):
            req = "tc32" if ok32 else ("bf16_wide" if (self.tb.smem_bytes < 0 and self.device.type == "cuda") else "fp32")
        if req == "bf16_wide":
            from .wide import WideLocal
            if self.device.type != "cuda":
                raise ValueError("the wide (layer-wise tensor-core) path needs a CUDA device")
            self.precision_name = "bf16_wide"
            self._wide_pending = True
            return
        if req in ("tc32", "fp32x2", "tf32", "tc"):
            if not ok32:
                raise ValueError("the tensor-core path needs a CUDA device, full-batch steps and 3-layer ReLU MLPs with "
                                 "in<=15, h1<=63, h2<=256, out<=8")
            self.precision_name = "tc32"
            self.tc_variant = "x2"
            self.tc_cluster = tc.tc_cluster_size(self.tb.groups, "x2")
        elif req == "bf16":
            if not ok:
                raise ValueError("the tensor-core path needs a CUDA device, full-batch steps and 3-layer ReLU MLPs with "
                                 "in<=15, h1<=63, h2<=256, out<=8")
            self.precision_name = "bf16"
            self.tc_variant = "bf16"
            self.tc_cluster = tc.tc_cluster_size(self.tb.groups)
        elif req == "fp32":
            self.precision_name = "fp32"
        else:
            raise ValueError(f"unknown precision {req!r}")

    # ------------------------------------------------------------------------------ tables
    def _build_tables(self) -> None:
        G = sum(p.n_clients for p in self.problems)
        F = len(self.problems)
        groups = np.zeros(G, dtype=nat.GROUP_DTYPE)
        feds = np.zeros(F, dtype=nat.FED_DTYPE)
        idx_parts: List[np.ndarray] = []
        slab_off = idx_off = scratch_off = global_off = 0
        g = 0
        self.group_of: List[List[int]] = []
        for f, p in enumerate(self.problems):
            s, h = p.spec, p.hyper
            P = s.slab_size
            sizes = [len(sh) for sh in p.shards]
            total = float(sum(sizes)) if p.total_weight is None else float(p.total_weight)
            feds[f] = (g, p.n_clients, P, s.n_classes, global_off, nat.METRIC_IDS[p.metric_mode], p.patience,
                       p.atol, p.rtol, self.max_rounds, 0)
            global_off += P
            members = []
            for k, sh in enumerate(p.shards):
                sh = np.asarray(sh, dtype=np.int32)
                if len(sh) == 0:
                    raise ValueError(f"client {k} of federation {f} has an empty shard (the reference would deadlock: S:91-93)")
                rows_per_step = min(h.batch_size, len(sh)) if h.batch_size > 0 else len(sh)
                # >= 8 records: the tensor-core kernel keeps one per CTA of its (<= 8-wide) cluster
                max_tiles = max(8, (rows_per_step + TILE - 1) // TILE)
                r = groups[g]
                r["n_layers"] = s.n_layers
                r["dims"][: len(s.dims)] = s.dims
                r["head"], r["act"] = nat.HEAD_IDS[s.head], nat.ACT_IDS[s.activation]
                r["slab_size"], r["fed"], r["slab_off"], r["idx_off"] = P, f, slab_off, idx_off
                r["n_rows"], r["batch_size"] = len(sh), h.batch_size
                r["opt"], r["nesterov"] = nat.OPT_IDS[h.opt], int(h.nesterov)
                r["lr"], r["beta1"], r["beta2"], r["eps"] = h.lr, h.beta1, h.beta2, h.eps
                r["momentum"], r["alpha"] = h.momentum, h.alpha
                r["lr_step_size"], r["lr_gamma"] = h.lr_step_size, h.lr_gamma
                if p.weighting == "samples":
                    r["agg_num"], r["agg_den"] = float(len(sh)), total
                else:
                    r["agg_num"], r["agg_den"] = 1.0, float(p.n_clients if p.total_weight is None else p.total_weight)
                r["scratch_off"], r["max_tiles"], r["client"] = scratch_off, max_tiles, k + p.client_offset   # GLOBAL client index
                slab_off += P
                idx_off += len(sh)
                scratch_off += max_tiles * (P + nat.LOSS_SLOTS)
                idx_parts.append(sh)
                members.append(g)
                g += 1
            self.group_of.append(members)
        self.n_groups, self.n_feds = G, F
        self.params_total, self.global_total, self.scratch_total = slab_off, global_off, scratch_off
        # evaluation index sets appended after the shards: the whole test set, the whole train set
        self.test_idx_off = idx_off
        if self.n_test:
            idx_parts.append(np.arange(self.n_train, self.n_train + self.n_test, dtype=np.int32))
            idx_off += self.n_test
        self.train_idx_off = idx_off
        idx_parts.append(np.arange(self.n_train, dtype=np.int32))
        self.idx_host = np.concatenate(idx_parts)
        self.tb = K.DeviceTables.build(groups, feds, self.device)
        self.idx = torch.from_numpy(self.idx_host).to(self.device)
        # training tiles
        # (largest models first: CTAs are dispatched in list order, and a 32-row tile of an 86 K-parameter model runs ~50x longer
        #  than one of a 1 K-parameter model; with the long tiles at the END of the list a packed grid's launch ended with a
        #  wave of them on a mostly idle GPU)
        order = sorted(range(G), key=lambda gi: -int(groups[gi]["slab_size"]))
        tiles = [(gi, t) for gi in order for t in range(int(groups[gi]["max_tiles"]))]
        self.train_tiles_np = np.array(tiles, dtype=np.int32).view(nat.TILE_DTYPE).reshape(-1)
        self.train_tiles = K.upload_table(self.train_tiles_np, self.device)
        self.n_batches = max(
            1 if p.hyper.batch_size <= 0 else max(math.ceil(len(sh) / min(p.hyper.batch_size, len(sh))) for sh in p.shards)
            for p in self.problems)
        # local evaluation: every client's own slab on its own shard -> confusion slot g
        self.local_jobs_np, self.local_jobs, self.local_tiles, self.n_local_tiles = self._make_jobs(
            [(gi, int(groups[gi]["n_rows"]), int(groups[gi]["slab_off"]), int(groups[gi]["idx_off"]), gi, 0) for gi in range(G)])
        # held-out evaluation of each federation's global model -> slots G+f (test), G+F+f (train)
        self.n_conf_slots = G + 2 * F
        self._global_jobs = {}
        for which, off, n in (("test", self.test_idx_off, self.n_test), ("train", self.train_idx_off, self.n_train)):
            if n == 0:
                continue
            base = G if which == "test" else G + F
            self._global_jobs[which] = self._make_jobs(
                [(int(feds[f]["first_group"]), n, int(feds[f]["global_off"]), off, base + f, 1) for f in range(F)])

    def _make_jobs(self, rows):
        jobs = np.zeros(len(rows), dtype=nat.EVALJOB_DTYPE)
        tiles = []
        for j, (grp, n, slab_off, idx_off, slot, from_global) in enumerate(rows):
            jobs[j] = (grp, n, slab_off, idx_off, slot, from_global)
            tiles += [(j, t) for t in range((n + TILE - 1) // TILE)]
        tiles_np = np.array(tiles, dtype=np.int32).view(nat.TILE_DTYPE).reshape(-1)
        return jobs, K.upload_table(jobs, self.device), K.upload_table(tiles_np, self.device), len(tiles_np)

    # ------------------------------------------------------------------------------- state
    def _alloc_state(self) -> None:
        d, G, F = self.device, self.n_groups, self.n_feds
        z = lambda n, dt=torch.float32: torch.zeros(n, dtype=dt, device=d)
        self.params, self.m, self.v = z(self.params_total), z(self.params_total), z(self.params_total)
        self.global_ = z(self.global_total)
        self.scratch = z(self.scratch_total if self.tb.smem_bytes >= 0 else 32)   # (the wide path keeps no per-tile gradient records)
        self.step = z(G, torch.int32)
        self.active = torch.ones(G, dtype=torch.int32, device=d)
        self.loss = z(G)
        self.conf = z(self.n_conf_slots * nat.MAX_CLASSES ** 2, torch.int32)
        self.hist_local = z(self.max_rounds * G * 4)
        self.hist_global = z(F * self.max_rounds * 4)
        st = np.zeros(F, dtype=nat.FEDSTATE_DTYPE)
        st["count"] = self.tb.feds["patience"]
        self.fed_state = K.upload_table(st, d)
        # sklearn-style fit state
        self.epoch, self.no_improve = z(G, torch.int32), z(G, torch.int32)
        self.best_loss = torch.full((G,), float("inf"), dtype=torch.float64, device=d)
        self.last_loss = z(G, torch.float64)
        self.bs_stride = K.batch_stats_stride(self.tb.max_slab)
        self.batch_stats = z(G * self.n_batches * self.bs_stride, torch.float64)
        self.idx_epoch = self.idx.clone()
        self.tail_count = z(1, torch.int32)
        # mapped pinned host copy of the federation state: the fused round kernels store their result there too
        # ... followed (next multiple of 16 bytes) by 4 self-validating 16-byte result chunks per federation, see include/fedb200.h
        nst = self.fed_state.numel()
        self._result_off = (nst + 15) // 16 * 16
        self._mirror_full = torch.zeros(self._result_off + 64 * self.n_feds, dtype=torch.uint8).pin_memory() if d.type == "cuda" else None
        self._state_pinned = self._mirror_full[:nst] if d.type == "cuda" else None
        self._e2e_cache: Dict = {}   # (x_host ptr, y_host ptr) -> (graph replay, stream sync, result view, buffers kept alive, done view)
        self._e2e_expect = -1        # value of `done` the next polled step will publish (-1: synchronise instead)

    def _ensure_wide(self):
        if getattr(self, "_wide_pending", False):
            from .wide import WideLocal
            self._wide_pending = False
            self.wide = WideLocal(self)
        return self.wide

    def init_params(self) -> None:
        """(Re-)initialise every client's parameters and clear optimizer state (K4)."""
        by_kind: Dict[str, np.ndarray] = {}
        for f, p in enumerate(self.problems):
            if p.init == "sklearn_exact":
                slab = torch.from_numpy(minit.init_slab_sklearn_exact(p.spec, p.init_seed))
                for gi in self.group_of[f]:
                    o = int(self.tb.groups[gi]["slab_off"])
                    self.params[o: o + p.spec.slab_size].copy_(slab)
                continue
            streams = by_kind.setdefault(p.init, np.full(self.n_groups, -1, dtype=np.int32))
            for k, gi in enumerate(self.group_of[f]):
                # one Philox stream per (federation seed, client); identical streams when the
                # federation asks for a shared init
                streams[gi] = (p.init_seed * 1000003 + ((k + p.client_offset) if p.init_per_client else 0)) & 0x7FFFFFFF
        for kind, streams in by_kind.items():
            K.init_params(self.tb, self.params, torch.from_numpy(streams).to(self.device), kind, self.seed)
        self.m.zero_()
        self.v.zero_()
        self.step.zero_()
        if self.precision_name == "bf16_wide":
            w = self._ensure_wide()
            w.refresh_shadow()

    # ------------------------------------------------------------------------------ rounds
    def enqueue_local_step(self, batch: int = 0, idx: Optional[torch.Tensor] = None, stats: bool = False) -> None:
        """One optimizer step of every active client on minibatch ``batch`` (2 launches)."""
        idx = self.idx if idx is None else idx
        tiles, n_tiles = getattr(self, "_step_tiles", None) or (self.train_tiles, len(self.train_tiles_np))
        K.fwd_bwd(self.tb, tiles, n_tiles, self.params, self.X, self.y, idx, batch,
                  self.scratch, self.step, self.active, self.fed_state)
        K.reduce_opt(self.tb, self.params, self.m, self.v, self.scratch, batch, self.step, self.active, self.fed_state,
                     self.loss, self.batch_stats if stats else None, self.n_batches)

    def enqueue_eval_local(self) -> None:
        K.evaluate(self.tb, self.local_jobs_np, self.local_jobs, self.local_tiles, self.n_local_tiles, self.params,
                   self.global_, self.X, self.y, self.idx, self.conf, self.fed_state)

    def enqueue_aggregate(self, write_back: bool = True, do_metrics: bool = True, begin_next: bool = False) -> None:
        # wide path: the averaged fp32 weights replace every client's bf16 GEMM operands in the same pass
        shadow = self.wide.shadow if (self.precision_name == "bf16_wide" and write_back and self.wide is not None) else None
        K.fedavg(self.tb, self.fed_state, self.params, self.global_, self.conf, self.hist_local, self.hist_global,
                 write_back, do_metrics, begin_next, shadow)

    def _begin_is_fused(self) -> bool:
        """True when the aggregation kernel also performs the next round's begin bookkeeping (the
        single-launch tensor-core round; the cross-GPU kernel in DistFedEngine)."""
        return self._fused_round_ok()

    def prime(self) -> None:
        """With a fused begin, the very first round_begin has to be run once, eagerly."""
        if self._begin_is_fused() and not self._primed:
            self.enqueue_round_begin()
            self._primed = True

    # one CTA aggregates every federation in the fused tail: worth it while that is a few microseconds
    FUSED_TAIL_MAX_FLOATS = 1 << 20

    def _fused_round_ok(self) -> bool:
        return (self.tc_variant is not None and type(self).enqueue_aggregate is FedEngine.enqueue_aggregate
                and self.params_total <= self.FUSED_TAIL_MAX_FLOATS and self.n_feds <= 256
                and max(p.n_clients for p in self.problems) <= 32)

    def enqueue_round(self, host_inputs=None) -> None:
        """The custom script's round (C:130-201): local step(s), local eval, metrics, FedAvg (+ the next
        round's begin bookkeeping).  Requires :meth:`prime` once before the first round.  On the
        tensor-core path with a small arena this is ONE launch (fedb_round_tc).  ``host_inputs`` =
        (x_host, y_host) pinned tensors (single-launch rounds only, see :meth:`pulls_host_inputs`): the
        kernel copies the step's rows from host memory itself."""
        if self._fused_round_ok():
            from ..ops import tc
            E = max(p.hyper.local_epochs for p in self.problems)
            tc.round_tc(self.tb, self.tc_cluster, self.X, self.y, self.idx, self.params, self.m, self.v, self.step, self.active,
                        self.fed_state, self.loss, self.conf, self.scratch, E, self.global_, self.hist_local, self.hist_global,
                        self.tail_count, state_mirror=self._state_pinned, host_inputs=host_inputs, variant=self.tc_variant)
            return
        assert host_inputs is None, "host inputs are pulled by the single-launch round only"
        if not self._begin_is_fused():
            self.enqueue_round_begin()
        with trace.span("fedb200.local_train_eval"):
            self.enqueue_local_part()
        with trace.span("fedb200.aggregate"):
            self.enqueue_aggregate()

    def enqueue_round_begin(self) -> None:
        K.round_begin(self.tb, self.fed_state, self.conf)

    def enqueue_local_part(self) -> None:
        """Everything a round does that involves only resident clients: E local steps + local eval."""
        E = max(p.hyper.local_epochs for p in self.problems)
        if self.precision_name == "bf16_wide":
            for _ in range(E):
                self.wide.enqueue_step()
            self.wide.enqueue_eval()
        elif self.tc_variant is not None:
            from ..ops import tc
            tc.local_round_tc(self.tb, self.tc_cluster, self.X, self.y, self.idx, self.params, self.m, self.v, self.step,
                              self.active, self.fed_state, self.loss, self.conf, self.scratch, 0, E, True, variant=self.tc_variant)
        else:
            for _ in range(E):
                for b in range(self.n_batches):
                    self.enqueue_local_step(b)
            self.enqueue_eval_local()

    @property
    def launches_per_round(self) -> int:
        E = max(p.hyper.local_epochs for p in self.problems)
        if self._fused_round_ok():
            return 1
        if self.precision_name == "bf16_wide":
            return 1 + E * self.wide.launches_per_step + (self.wide.L - 1) + 1 + 1
        if self.tc_variant is not None:
            return 3 if not self._begin_is_fused() else 2
        return 1 + 2 * E * self.n_batches + 2

    def _graph(self, key: str, fn, unroll: int = 1):
        """Capture ``unroll`` repetitions of ``fn`` into a CUDA graph (cached)."""
        gk = f"{key}x{unroll}"
        if gk not in self._graphs:
            fn()  # warm-up outside capture: sets function attributes, touches every buffer
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(unroll):
                    fn()
            self._graphs[gk] = g
        return self._graphs[gk]

    def fed_state_host(self) -> np.ndarray:
        return self.fed_state.cpu().numpy().view(nat.FEDSTATE_DTYPE).copy()

    def run(self, rounds: int, use_graph: Optional[bool] = None, unroll: int = 8, poll: bool = True) -> Dict[str, np.ndarray]:
        """Run up to ``rounds`` rounds (early stop per federation, C:181-195).  Returns
        :meth:`history`.  On CUDA the round is replayed from a graph, ``unroll`` rounds per
        launch; the stop flags are polled once per replay."""
        st = self.fed_state_host()
        done = int(st["done"].max()) if len(st) else 0
        rounds = min(rounds, self.max_rounds - done)
        self._e2e_expect = -1
        self.prime()
        use_graph = (self.device.type == "cuda") if use_graph is None else use_graph
        if not use_graph:
            for _ in range(rounds):
                self.enqueue_round()
                if poll and bool(self.fed_state_host()["stop"].all()):
                    break
            self.check_errors()
            return self.history()
        # Graph capture runs one warm-up round eagerly; snapshot and restore state around it so
        # capture itself is invisible to the training trajectory.
        left = rounds
        while left > 0:
            u = unroll if left >= unroll else 1
            if f"roundx{u}" not in self._graphs:
                snap = self._snapshot()
                self._graph("round", self.enqueue_round, u)
                self._restore(snap)
            self._graphs[f"roundx{u}"].replay()
            left -= u
            if poll and bool(self.fed_state_host()["stop"].all()):
                break
        self.check_errors()
        return self.history()

    def _snapshot(self):
        names = ("params", "m", "v", "global_", "step", "active", "loss", "conf", "hist_local", "hist_global",
                 "fed_state", "epoch", "no_improve", "best_loss", "last_loss", "batch_stats", "idx_epoch")
        return {n: getattr(self, n).clone() for n in names}

    def _restore(self, snap) -> None:
        for n, t in snap.items():
            getattr(self, n).copy_(t)
        self._e2e_expect = -1
        self._sync_derived_state()

    def _sync_derived_state(self) -> None:
        """Bring every copy that is derived from ``params`` back in line after they were overwritten from outside a
        kernel (snapshot restore, checkpoint load): the wide path's bf16 GEMM operands (``WideLocal.shadow``)."""
        if self.precision_name == "bf16_wide" and self.wide is not None:
            self.wide.refresh_shadow()

    def check_errors(self) -> None:
        """Raise if a kernel reported a failure through a device flag (CTA-pair GEMM pipeline stall on the wide path;
        the multi-GPU engine adds the cross-GPU barrier timeout).  Synchronises the device."""
        if self.precision_name == "bf16_wide":
            from ..ops import gemm
            rc = gemm.pair_status()
            if rc != 0:
                raise RuntimeError(f"a CTA-pair GEMM reported failure code {rc} (pipeline stall)")

    # ------------------------------------------------------------- sklearn-style local fit
    def reset_fit_state(self, keep_no_improvement: bool = False) -> None:
        """What a non-incremental sklearn ``fit`` resets: a new optimizer (SK:622-640) and,
        unless ``keep_no_improvement`` (sklearn's warm_start behaviour, SK:433-441), the loss
        curve bookkeeping."""
        self.m.zero_()
        self.v.zero_()
        self.step.zero_()
        self.epoch.zero_()
        self.active.fill_(1)
        if not keep_no_improvement:
            self.no_improve.zero_()
            self.best_loss.fill_(float("inf"))

    def _uses_shuffle_table(self) -> bool:
        kinds = {p.hyper.shuffle_stream for p in self.problems}
        if kinds - {"feistel", "mt19937"} or len(kinds) > 1:
            raise ValueError(f"shuffle_stream must be 'feistel' or 'mt19937', the same for every federation of an engine: {kinds}")
        return kinds == {"mt19937"}

    def _shuffle_table(self) -> torch.Tensor:
        """int32 [T, rows of all shards]: row e = the index array of epoch e (sklearn's MT19937 order, models/init.py)."""
        if getattr(self, "_shuffle_tab", None) is None:
            T = max(max(p.hyper.max_iter, 1) for p in self.problems)
            tab = np.empty((T, self.test_idx_off), dtype=np.int32)
            cache = {}
            for f, p in enumerate(self.problems):
                for k, gi in enumerate(self.group_of[f]):
                    g = self.tb.groups[gi]
                    n, off = int(g["n_rows"]), int(g["idx_off"])
                    key = (p.spec.dims, p.init_seed, n, T)
                    if key not in cache:
                        cache[key] = minit.sklearn_shuffle_table(p.spec, p.init_seed, n, T)
                    tab[:, off: off + n] = self.idx_host[off: off + n][cache[key]]
            self._shuffle_tab = torch.from_numpy(tab).to(self.device)
        return self._shuffle_tab

    def enqueue_epoch(self, shuffle: bool = True) -> None:
        if shuffle and self._uses_shuffle_table():
            # every client still active is in the same epoch (they all started at 0); a few small torch ops, graph-capturable
            tab = self._shuffle_table()
            ep = self.epoch.max().clamp(max=tab.shape[0] - 1).to(torch.int64).view(1)
            self.idx_epoch[: tab.shape[1]].copy_(tab.index_select(0, ep).view(-1))
        elif shuffle:
            K.shuffle(self.tb, self.idx, self.idx_epoch, self.epoch, self.active, self.seed + 0x5EED)
        src = self.idx_epoch if shuffle else self.idx
        for b in range(self.n_batches):
            self.enqueue_local_step(b, src, stats=True)
        h = self.problems[0].hyper
        K.epoch_end(self.tb, self.batch_stats, self.n_batches, h.tol, h.n_iter_no_change, h.max_iter, self.active,
                    self.epoch, self.no_improve, self.best_loss, self.last_loss)

    def _fit_tc_mask(self) -> Optional[np.ndarray]:
        """bool per client: trained by the tensor-core fit kernel.  None: not on this engine (CPU, MT19937 parity stream,
        FEDB200_FIT=launches | cluster forces the CUDA-core paths)."""
        if self.device.type != "cuda" or self._uses_shuffle_table() or self._fit_mode in ("launches", "cluster"):
            return None
        if getattr(self, "_tc_fit_mask", None) is None:
            self._tc_fit_mask = K.fitx_eligible(self.tb.groups)
        return self._tc_fit_mask

    def _fit_in_one_launch(self) -> bool:
        """Whether :meth:`fit_local` runs as one persistent launch (fedb_fit_cluster: a cluster per client loops over epochs and
        mini-batches on the device) instead of one graph replay per epoch.  Opt-in with FEDB200_FIT=cluster: the results are
        bit-identical, but measured on B200 the graph of per-step launches is faster (4-client sklearn script 0.53 s against
        0.59 s, 90-federation grid 4.0 s against 4.9 s): its reduction + optimizer step spreads over 88 CTAs instead of the
        client's 7, and short graph nodes cost less than the three cluster barriers per mini-batch."""
        if self.device.type != "cuda" or self.tb.smem_bytes <= 0 or self._uses_shuffle_table():
            return False
        return self._fit_mode == "cluster"

    def fit_local(self, max_iter: Optional[int] = None, shuffle: bool = True, check_every: int = 8) -> np.ndarray:
        """Train every client until its loss stalls (SK:750-777) or ``max_iter`` epochs.
        Returns the number of epochs each client ran (sklearn's ``n_iter_``)."""
        max_iter = max_iter or max(p.hyper.max_iter for p in self.problems)
        tc_mask = self._fit_tc_mask()
        if tc_mask is not None and bool(tc_mask.any()):
            # K2b on the tensor cores (csrc/fedb_fitx.cu): the whole fit of every eligible client in ONE launch, epoch loop and
            # stopping rule on the device; clients of other shapes follow on the CUDA-core kernels below.  (Measured on the
            # 90-federation grid: putting the one launch on a second stream next to the CUDA-core epochs changes the total by
            # under 1 % - the 576 CTAs hold every SM while they run - so the two parts simply follow each other.)
            h = self.problems[0].hyper
            act_tc = torch.from_numpy(tc_mask.astype(np.int32)).to(self.device)
            # (the kernel clears a client's flag in act_tc when its fit stops, so the complement for the CUDA-core clients is
            #  built from the host mask: taken from act_tc afterwards it re-activated every converged tensor-core client for
            #  one more epoch on the CUDA cores)
            act_cc = torch.from_numpy((~tc_mask).astype(np.int32)).to(self.device)
            K.fit_tcx(self.tb, self.params, self.m, self.v, self.X, self.y, self.idx, self.idx_epoch, self.step, act_tc, self.fed_state,
                      self.epoch, self.no_improve, self.best_loss, self.last_loss, self.loss, h.tol, h.n_iter_no_change, h.max_iter,
                      max_iter, shuffle, self.seed + 0x5EED)
            if bool(tc_mask.all()):
                self.active.fill_(1)
                return self.epoch.cpu().numpy().copy()
            self.active.copy_(act_cc)
        # the fit's launches carry only the tiles they need: none of the clients the tensor-core kernel took, and per client
        # the tiles of one mini-batch (the general list reserves at least 8), largest models first
        self._step_tiles = self._fit_tile_list(tc_mask)
        try:
            self._fit_cuda_cores(max_iter, shuffle, check_every)
        finally:
            self._step_tiles = None
        self.active.fill_(1)
        return self.epoch.cpu().numpy().copy()

    def _fit_tile_list(self, tc_mask: Optional[np.ndarray]):
        if getattr(self, "_fit_tiles", None) is None:
            groups = self.tb.groups
            order = sorted(range(len(groups)), key=lambda gi: -int(groups[gi]["slab_size"]))
            tiles = []
            for gi in order:
                if tc_mask is not None and tc_mask[gi]:
                    continue
                n, bs = int(groups[gi]["n_rows"]), int(groups[gi]["batch_size"])
                rows = min(bs, n) if bs > 0 else n
                tiles += [(gi, t) for t in range(min(int(groups[gi]["max_tiles"]), -(-rows // nat.TILE_ROWS)))]
            if not tiles:
                tiles = [(0, 0)]
            arr = np.array(tiles, dtype=np.int32).view(nat.TILE_DTYPE).reshape(-1)
            self._fit_tiles = (K.upload_table(arr, self.device), len(arr))
        return self._fit_tiles

    def _fit_cuda_cores(self, max_iter: int, shuffle: bool, check_every: int) -> None:
        """The fit of the clients whose ``active`` flag is set, on the fp32 CUDA-core kernels."""
        if self._fit_in_one_launch():
            # K2b, persistent form: epochs, mini-batches and the stopping rule all run on the device (fedb_fit_cluster)
            h = self.problems[0].hyper
            tiles_per_batch = max(-(-min(int(g["batch_size"]) or int(g["n_rows"]), int(g["n_rows"])) // nat.TILE_ROWS) for g in self.tb.groups)
            K.fit_cluster(self.tb, max(1, min(8, tiles_per_batch)), self.params, self.m, self.v, self.X, self.y, self.idx, self.idx_epoch,
                          self.scratch, self.step, self.active, self.fed_state, self.epoch, self.no_improve, self.best_loss,
                          self.last_loss, self.loss, self.batch_stats, self.n_batches, h.tol, h.n_iter_no_change, h.max_iter,
                          max_iter, shuffle, self.seed + 0x5EED)
            return
        use_graph = self.device.type == "cuda"
        if use_graph and "epochx1" not in self._graphs:
            snap = self._snapshot()
            self._graph("epoch", lambda: self.enqueue_epoch(shuffle), 1)
            self._restore(snap)
        it = 0
        while it < max_iter:
            n = min(check_every, max_iter - it)
            for _ in range(n):
                if use_graph:
                    self._graphs["epochx1"].replay()
                else:
                    self.enqueue_epoch(shuffle)
            it += n
            if not bool(self.active.any()):
                break

    # ----------------------------------------------------------------------------- results
    def history(self) -> Dict[str, np.ndarray]:
        st = self.fed_state_host()
        R = int(st["done"].max()) if len(st) else 0
        hg = self.hist_global.cpu().numpy().reshape(self.n_feds, self.max_rounds, 4)[:, :R].astype(np.float64)
        hl = self.hist_local.cpu().numpy().reshape(self.max_rounds, self.n_groups, 4)[:R].astype(np.float64)
        return {"global": hg, "local": hl, "rounds": st["done"].copy(), "stopped": st["stop"].astype(bool) | st["stop_pending"].astype(bool),
                "last": st["last"].copy(), "metric_names": np.asarray(METRIC_NAMES)}

    def confusion(self) -> np.ndarray:
        return self.conf.cpu().numpy().reshape(self.n_conf_slots, nat.MAX_CLASSES, nat.MAX_CLASSES).copy()

    def global_slab(self, f: int = 0) -> torch.Tensor:
        o = int(self.tb.feds[f]["global_off"])
        return self.global_[o: o + self.problems[f].spec.slab_size]

    def client_slab(self, f: int, k: int) -> torch.Tensor:
        o = int(self.tb.groups[self.group_of[f][k]]["slab_off"])
        return self.params[o: o + self.problems[f].spec.slab_size]

    def evaluate_global(self, which: str = "test") -> List[Dict[str, float]]:
        """Metrics of every federation's *global* model on the held-out test split (or the
        whole train set).  The reference broadcasts X_test/y_test and never uses them
        (C:245-246); this is the extension BASELINE.md's accuracy band was measured with."""
        from .metrics import metrics_from_confusion
        if self.tb.smem_bytes < 0:
            return self._evaluate_global_torch(which)
        jobs_np, d_jobs, tiles, n_tiles = self._global_jobs[which]
        base = self.n_groups if which == "test" else self.n_groups + self.n_feds
        C2 = nat.MAX_CLASSES ** 2
        self.conf[base * C2: (base + self.n_feds) * C2].zero_()
        K.evaluate(self.tb, jobs_np, d_jobs, tiles, n_tiles, self.params, self.global_, self.X, self.y, self.idx,
                   self.conf, None)
        cm = self.confusion()
        return [metrics_from_confusion(cm[base + f][: p.spec.n_classes, : p.spec.n_classes]) for f, p in enumerate(self.problems)]

    def _evaluate_global_torch(self, which: str) -> List[Dict[str, float]]:
        """Held-out evaluation for models too wide for the SIMT eval kernel: a plain fp32 torch forward
        of the global slab (reporting only — not part of any round)."""
        from .metrics import confusion_matrix, metrics_from_confusion
        lo, n = (self.n_train, self.n_test) if which == "test" else (0, self.n_train)
        X, y = self.X[lo: lo + n], self.y[lo: lo + n]
        out = []
        for f, p in enumerate(self.problems):
            s, slab = p.spec, self.global_slab(f)
            a = X[:, : s.dims[0]]
            for l in range(s.n_layers):
                a = a @ s.weight_view(slab, l).T + s.bias_view(slab, l)
                if l + 1 < s.n_layers:
                    a = torch.relu(a)
            pred = a.argmax(1) if s.head == "softmax_ce" else (a[:, 0] > 0).long()
            out.append(metrics_from_confusion(confusion_matrix(y.cpu().numpy(), pred.cpu().numpy(), s.n_classes)))
        return out

    def _client_opt_state(self, f: int) -> Dict[str, np.ndarray]:
        """Optimizer state of the clients of federation ``f`` hosted here, rows in local client order."""
        p = self.problems[f]
        gs = self.group_of[f]
        P = p.spec.slab_size
        o = int(self.tb.groups[gs[0]]["slab_off"])
        n = len(gs) * P
        return {"m": self.m[o: o + n].view(len(gs), P).cpu().numpy(), "v": self.v[o: o + n].view(len(gs), P).cpu().numpy(),
                "step": self.step[gs[0]: gs[0] + len(gs)].cpu().numpy(), "params": self.params[o: o + n].view(len(gs), P).cpu().numpy()}

    def _gather_opt_state(self, f: int) -> Optional[Dict[str, np.ndarray]]:
        """All clients' optimizer state in GLOBAL client order (single process: this one's; overridden by the multi-process
        engines, where it is a collective and only rank 0 gets the result)."""
        return self._client_opt_state(f)

    def save_checkpoint(self, path: str, f: int = 0, resume: bool = False) -> None:
        """Write the global model of federation ``f`` in the coefs_/intercepts_ layout.  ``resume=True`` adds what an exact
        continuation needs: every client's parameters, moments and step count (rows = global client index) and the
        federation's round / early-stop state and metric history.  Under torchrun every rank must call this (the
        optimizer state is gathered); rank 0 writes the file."""
        from ..models.checkpoint import save_checkpoint
        p = self.problems[f]
        opt, extras = None, {"name": p.name}
        if resume:
            opt = self._gather_opt_state(f)
            st = self.fed_state_host()[f]
            fs = {k: (st[k].tolist() if st[k].ndim else int(st[k])) for k in nat.FEDSTATE_DTYPE.names}
            # stored as it stands BETWEEN rounds, before the next round's begin bookkeeping (which the fused tails have
            # already done at this point: undo it; the loader's first round_begin redoes it, stop promotion included)
            fs["round"], fs["stop"] = fs["done"], 0
            extras["fed_state"] = fs
            if opt is not None:
                R = int(st["done"])
                h = self.history()
                opt = dict(opt, hist_global=h["global"][f][:R].astype(np.float32), hist_local=h["local"][:R].astype(np.float32))
        if getattr(self, "rank", 0) != 0:
            return
        save_checkpoint(path, p.spec, self.global_slab(f), int(self.fed_state_host()[f]["done"]), extras=extras,
                        opt_state={k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in opt.items()} if opt else None)

    def load_checkpoint(self, path: str, f: int = 0) -> Dict:
        """Load global weights into every client slab of federation ``f``; when the file carries resume state, also each
        resident client's own parameters / moments / step count (selected by global client index) and the federation's
        round counter, early-stop anchor / patience count and metric history, so the run continues as if uninterrupted."""
        from ..models.checkpoint import load_checkpoint
        p = self.problems[f]
        spec, coefs, inter, meta, opt = load_checkpoint(path)
        if spec.dims != p.spec.dims or spec.head != p.spec.head:
            raise ValueError(f"checkpoint is {spec.describe()}, federation is {p.spec.describe()}")
        slab = torch.zeros(spec.slab_size)
        spec.from_coefs_intercepts(coefs, inter, slab)
        self.global_slab(f).copy_(slab)
        gs = self.group_of[f]
        for k in range(len(gs)):
            self.client_slab(f, k).copy_(slab)
        if opt and "m" in opt:
            P, o = spec.slab_size, int(self.tb.groups[gs[0]]["slab_off"])
            n = len(gs) * P
            first = int(p.client_offset)
            if opt["m"].shape[0] < first + len(gs):
                raise ValueError(f"checkpoint holds optimizer state of {opt['m'].shape[0]} clients; this process hosts clients "
                                 f"{first}..{first + len(gs) - 1}")
            rows = slice(first, first + len(gs))
            self.m[o: o + n].copy_(torch.from_numpy(np.ascontiguousarray(opt["m"][rows])).reshape(-1))
            self.v[o: o + n].copy_(torch.from_numpy(np.ascontiguousarray(opt["v"][rows])).reshape(-1))
            self.params[o: o + n].copy_(torch.from_numpy(np.ascontiguousarray(opt["params"][rows])).reshape(-1))
            self.step[gs[0]: gs[0] + len(gs)].copy_(torch.from_numpy(np.ascontiguousarray(opt["step"][rows])))
        fs = (meta.get("extras") or {}).get("fed_state")
        if fs:
            st = self.fed_state_host()
            for k in nat.FEDSTATE_DTYPE.names:
                st[f][k] = fs[k]
            self.fed_state.copy_(torch.from_numpy(st.view(np.uint8).reshape(-1)).to(self.device))
            self._primed = False          # the next round's begin bookkeeping is still to be done (see save_checkpoint)
            self._e2e_expect = -1
            R = int(fs["done"])
            if opt and "hist_global" in opt and R:
                hg = self.hist_global.view(self.n_feds, self.max_rounds, 4)
                hl = self.hist_local.view(self.max_rounds, -1, 4)
                r = min(R, self.max_rounds)
                hg[f, :r].copy_(torch.from_numpy(opt["hist_global"][:r]))
                if opt["hist_local"].shape[1] == hl.shape[1]:
                    hl[:r].copy_(torch.from_numpy(opt["hist_local"][:r]))
        self._sync_derived_state()
        return meta

    # ------------------------------------------------------------------ end-to-end stepping
    def pinned_train_inputs(self):
        """Pinned host copies of the training rows / labels, in the device layout (features
        padded to ``ldx``) — the per-step inputs of :meth:`step_e2e`."""
        xh = torch.empty((self.n_train, self.ldx), dtype=torch.float32).pin_memory() if self.device.type == "cuda" \
            else torch.empty((self.n_train, self.ldx), dtype=torch.float32)
        yh = torch.empty(self.n_train, dtype=torch.int32)
        yh = yh.pin_memory() if self.device.type == "cuda" else yh
        xh.copy_(self.X[: self.n_train].cpu())
        yh.copy_(self.y[: self.n_train].cpu())
        return xh, yh

    def step_e2e(self, x_host: torch.Tensor, y_host: torch.Tensor) -> np.ndarray:
        """One round the way a caller that streams data would run it: upload this step's
        inputs from (pinned) host memory, run the round, read the round's result (federation
        state: stop flag + the four global metrics) back to the host.  Returns ``FedbFedState``
        records whose ``done``, ``stop`` and ``last`` fields are current; the record is an internal
        buffer that the next call overwrites (copy what you keep), and on the polled path the round
        kernel may still be retiring when this returns — its other fields (anchor, count, ...) are
        valid only after a stream synchronisation.  This mirrors what the reference does every round
        (H2D of the shard C:65-66, D2H of the predictions C:83) minus the weight traffic.
        On CUDA the uploads, the round and the read-back are ONE graph launch (captured per
        pair of host buffers), followed by one stream synchronisation.  When the round is a single
        launch (:meth:`pulls_host_inputs`) that launch performs the upload itself: the kernel reads the
        rows it trains on from the pinned buffers over the host link and writes them through to the
        device-resident X / y — two copy-engine nodes (13 us for 64 KB inside a graph) become ~3 us of
        loads that overlap the weight staging."""
        # steady state: everything about this pair of host buffers is cached; the host does a dict lookup, a graph launch and
        # a stream synchronisation per step (its time between the sync and the next launch is GPU idle time)
        ent = self._e2e_cache.get((x_host.data_ptr(), y_host.data_ptr())) if self._primed else None
        if ent is not None:
            ent[0]()                                     # one graph launch
            chunks = ent[5]
            if chunks is None or self._e2e_expect < 0:
                ent[1]()                                 # stream synchronise
                if chunks is not None:                   # (re-)arm the polled path on the state this round left behind
                    self._e2e_expect = -1 if ent[2]["stop"][0] else int(ent[2]["done"][0]) + 1
                return ent[2]
            # the round kernel stores four self-validating 16-byte chunks {done, stop, metric} into mapped pinned memory: poll
            # them instead of waiting for the whole kernel to retire, so the next launch overlaps this kernel's teardown
            import time as _time
            want, spins, d = self._e2e_expect, 0, chunks["done"]
            deadline = None
            while not (d[0] == want and d[1] == want and d[2] == want and d[3] == want):
                spins += 1
                if (spins & 255) == 0:                   # a time basis, not a spin count: 2 ms without a result -> synchronise
                    now = _time.perf_counter()
                    deadline = deadline or now + 2e-3
                if deadline is not None and _time.perf_counter() > deadline:   # stopped federation (no new round) or a slow step
                    ent[1]()
                    self._e2e_expect = -1 if ent[2]["stop"][0] else int(ent[2]["done"][0]) + 1
                    return ent[2]
            out = self._e2e_out
            out["done"][0], out["stop"][0] = want, chunks["stop"][0]
            out["last"][0] = chunks["val"]
            self._e2e_expect = want + 1 if not out["stop"][0] else -1
            return out
        self.prime()
        if self.device.type != "cuda":
            self.X[: self.n_train].copy_(x_host)
            self.y[: self.n_train].copy_(y_host)
            self.enqueue_round()
            return self.fed_state.numpy().view(nat.FEDSTATE_DTYPE)
        key = f"e2e:{x_host.data_ptr()}:{y_host.data_ptr()}"
        mirrored = self._result_is_mirrored()   # the round kernel itself writes the result into mapped host memory
        pull = self.pulls_host_inputs() and x_host.is_pinned() and y_host.is_pinned() \
            and tuple(x_host.shape) == (self.n_train, self.ldx) and x_host.dtype == torch.float32 and y_host.dtype == torch.int32

        def body():
            if pull:
                self.enqueue_round(host_inputs=(x_host, y_host))
            else:
                self.X[: self.n_train].copy_(x_host, non_blocking=True)
                self.y[: self.n_train].copy_(y_host, non_blocking=True)
                self.enqueue_round()
            if not mirrored:
                self._state_pinned.copy_(self.fed_state, non_blocking=True)

        stream = torch.cuda.current_stream()
        view = self._state_pinned.numpy().view(nat.FEDSTATE_DTYPE)
        if x_host.is_pinned() and y_host.is_pinned():
            if key + "x1" not in self._graphs:
                snap = self._snapshot()
                self._graph(key, body, 1)
                self._restore(snap)
            g = self._graphs[key + "x1"]
            poll = mirrored and self.n_feds == 1 and os.environ.get("FEDB200_E2E_SYNC") is None
            chunks = self._mirror_full.numpy()[self._result_off: self._result_off + 64].view(
                np.dtype([("done", "i4"), ("stop", "i4"), ("val", "f8")])) if poll else None
            self._e2e_out = np.zeros(self.n_feds, dtype=nat.FEDSTATE_DTYPE)   # the polled path returns done / stop / last in here
            self._e2e_cache[(x_host.data_ptr(), y_host.data_ptr())] = (g.replay, stream.synchronize, view, x_host, y_host, chunks)
            g.replay()
        else:
            body()
        stream.synchronize()
        self._e2e_expect = -1 if view["stop"][0] else int(view["done"][0]) + 1
        return view

    def _result_is_mirrored(self) -> bool:
        return self._fused_round_ok()

    def pulls_host_inputs(self) -> bool:
        """True if :meth:`step_e2e` lets the round kernel read its inputs from pinned host memory
        (single-launch rounds; FEDB200_E2E_MEMCPY=1 forces the copy-engine path)."""
        return self._fused_round_ok() and os.environ.get("FEDB200_E2E_MEMCPY") is None

    def e2e_bytes(self):
        return self.n_train * self.ldx * 4 + self.n_train * 4, int(self.fed_state.numel())

    # ------------------------------------------------------------------- bench conveniences
    precision_name = "fp32"
    tc_variant = None

    @property
    def tc_kernel(self) -> str:
        return "k_round_tcx" if self.tc_variant == "x2" else "k_local_round_tc"

    @property
    def dtype_label(self) -> str:
        """What the arithmetic of the training step is, for benchmark records."""
        return {"tc32": "fp32 (operands split into 2 x fp16, 3 tcgen05 MMAs per GEMM; fp32 accumulate, weights, optimizer)",
                "bf16": "bf16", "bf16_wide": "bf16", "fp32": "fp32"}[self.precision_name]

    @property
    def kernel_names(self) -> List[str]:
        """Kernels launched per steady-state round (the first round additionally runs k_round_begin)."""
        if self._fused_round_ok():
            return [f"{self.tc_kernel}[train+eval+fedavg+metrics]"]
        if self.precision_name == "bf16_wide":
            return ["k_round_begin", "k_gemm_tma2[fwd+bias+relu|dgrad+mask] (k_gemm_tma off the 256-multiples)", "k_gemm_tma2p[wgrad+optimizer]",
                    "k_wide_head_fwd", "k_wide_head_bwd", "k_wide_bias_opt", "k_fedavg[+bf16 shadow]"]
        if self.tc_variant is not None:
            return ["k_round_begin", f"{self.tc_kernel}[train+eval]", "k_fedavg"]
        return ["k_round_begin", "k_fwd_bwd", "k_reduce_opt", "k_eval", "k_fedavg"]

    def step_graph(self) -> None:
        """Enqueue exactly one round (graph replay on CUDA); no host synchronisation."""
        self._e2e_expect = -1          # the polled end-to-end path must re-arm on the state this round leaves behind
        self.prime()
        if self.device.type != "cuda":
            self.enqueue_round()
            return
        if "roundx1" not in self._graphs:
            snap = self._snapshot()
            self._graph("round", self.enqueue_round, 1)
            self._restore(snap)
        self._graphs["roundx1"].replay()
