"""Command line: ``python -m fedb200 {custom,sklearn-avg,tune} [...]``.

The reference has no flags at all — every hyper-parameter is a literal in its three scripts
(SURVEY §5).  Here each script is a sub-command whose DEFAULTS ARE those literals, so

    python -m fedb200 custom --clients 3          ~ mpirun -np 3 python FL_Custom…Multiple_Rounds.py
    python -m fedb200 sklearn-avg --clients 4     ~ mpirun -np 4 python FL_SkLearn…Limitation.py
    python -m fedb200 tune --clients 4            ~ mpirun -np 4 python hyperparameters_tuning.py

(on the bundled income table, which is the file the two mis-named scripts evidently mean —
SURVEY §0 R4).  ``--clients`` replaces the MPI process count: clients are simulated inside one
process per GPU.  Under ``torchrun`` (WORLD_SIZE > 1) ``custom`` spreads its clients over the
ranks and aggregates across GPUs in-kernel.
"""
from __future__ import annotations

import argparse
import os
import time
from typing import List, Optional

import numpy as np


def _common(p: argparse.ArgumentParser) -> None:
    p.add_argument("--data", default=None, help="CSV (header row) or packed .npz; default: the bundled income table")
    p.add_argument("--label", default="income")
    p.add_argument("--device", default=None, help="cuda | cpu (default: cuda when available)")
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--jsonl", default=None, help="write one JSON object per round / grid point")
    p.add_argument("--checkpoint", default=None, help="save the final global weights (.npz, coefs_/intercepts_ layout)")
    p.add_argument("--quiet", action="store_true")
    p.add_argument("--synthetic-rows", type=int, default=0, help="use income-shaped synthetic data with this many rows")


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="fedb200", description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    sub = ap.add_subparsers(dest="cmd", required=True)

    c = sub.add_parser("custom", help="multi-round FedAvg with the torch-style MLP (C:)")
    _common(c)
    c.add_argument("--clients", type=int, default=3)
    c.add_argument("--rounds", type=int, default=300)                       # C:249
    c.add_argument("--hidden", type=int, nargs="+", default=[50, 200])       # C:40
    c.add_argument("--lr", type=float, default=0.004)                        # C:44
    c.add_argument("--optimizer", default="adam", choices=["adam", "sgd"])
    c.add_argument("--momentum", type=float, default=0.0)
    c.add_argument("--lr-step-size", type=int, default=30)                   # C:46
    c.add_argument("--lr-gamma", type=float, default=0.5)
    c.add_argument("--local-epochs", type=int, default=1)                    # C:63-73: one step per round
    c.add_argument("--patience", type=int, default=10)                       # C:122
    c.add_argument("--tolerance", type=float, default=1e-4)
    c.add_argument("--partition", default="per_client_shuffle", choices=["per_client_shuffle", "contiguous", "dirichlet"])
    c.add_argument("--dirichlet-alpha", type=float, default=0.5)
    c.add_argument("--weighting", default="samples", choices=["samples", "uniform"])
    c.add_argument("--precision", default=None, choices=[None, "fp32", "tc32", "bf16"],
                   help="fp32 = CUDA-core kernels; tc32 = tensor cores with split fp16 operands (fp32-grade, default where the "
                        "model fits); bf16 = tensor cores with bf16 operands")
    c.add_argument("--shared-init", action="store_true", help="all clients start from the same weights")
    c.add_argument("--resume", default=None, help="checkpoint to start from")

    s = sub.add_parser("sklearn-avg", help="MLPClassifier-equivalent coefficient averaging (S:)")
    _common(s)
    s.add_argument("--clients", type=int, default=4)
    s.add_argument("--rounds", type=int, default=5)                          # S:203
    s.add_argument("--hidden", type=int, nargs="+", default=[50, 400])       # S:72
    s.add_argument("--lr", type=float, default=0.004)                        # S:73
    s.add_argument("--max-iter", type=int, default=300)                      # S:81
    s.add_argument("--policy", default="warm_start", choices=["warm_start", "reference_faithful", "sklearn_literal"])
    s.add_argument("--reference-faithful", action="store_const", const="reference_faithful", dest="policy")
    s.add_argument("--shuffle", default="feistel", choices=["feistel", "mt19937"],
                   help="per-epoch reshuffle: device-side keyed bijection, or sklearn's MT19937 stream (parity runs)")

    t = sub.add_parser("tune", help="hidden sizes x learning rates (x max_iter) grid as packed federations (H:)")
    _common(t)
    t.add_argument("--clients", type=int, default=4)
    t.add_argument("--max-iter", type=int, nargs="+", default=[400])         # H:90
    t.add_argument("--small", action="store_true", help="3 x 3 corner of the grid (smoke runs)")
    t.add_argument("--shuffle", default="feistel", choices=["feistel", "mt19937"],
                   help="per-epoch reshuffle: device-side keyed bijection, or sklearn's MT19937 stream (parity runs)")
    return ap


def _load(args, with_mean: bool):
    from .data.income import load_table, prepare, synthetic_income_like
    if args.synthetic_rows:
        return synthetic_income_like(args.synthetic_rows, seed=args.seed, with_mean=with_mean)
    table = load_table(args.data)
    return prepare(table, label=args.label, with_mean=with_mean)


def cmd_custom(args) -> int:
    from .data.partition import make_shards
    from .fed import logging as flog
    from .fed.engine import FedEngine
    from .fed.problem import FedProblem, TrainHyper
    data = _load(args, with_mean=True)                                       # C:235 StandardScaler()
    shards = make_shards(args.partition, len(data.X_train), args.clients, y=data.y_train, seed=args.seed, alpha=args.dirichlet_alpha)
    hyper = TrainHyper(opt="adam_torch" if args.optimizer == "adam" else "sgd", lr=args.lr, momentum=args.momentum,
                       lr_step_size=args.lr_step_size, lr_gamma=args.lr_gamma, local_epochs=args.local_epochs)
    prob = FedProblem.custom(data.n_features, data.n_classes, shards, hidden=tuple(args.hidden), hyper=hyper, patience=args.patience,
                             atol=args.tolerance, weighting=args.weighting, init_seed=args.seed, init_per_client=not args.shared_init)
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    max_rounds = args.rounds
    if args.resume:                      # --rounds counts the rounds still to run: make room for the history so far
        from .models.checkpoint import load_checkpoint
        max_rounds += int(load_checkpoint(args.resume)[3].get("round", 0))
    if world > 1:
        import torch
        from .parallel.dist_engine import DistFedEngine, GlooFedEngine, init_distributed
        on_gpu = torch.cuda.is_available() and args.device != "cpu"
        init_distributed("nccl" if on_gpu else "gloo")
        Engine = DistFedEngine if on_gpu else GlooFedEngine       # NVLink kernels / CPU plumbing (BASELINE config 1)
        eng = Engine([prob], data.X_train, data.y_train, data.X_test, data.y_test, rank=rank, world=world,
                     max_rounds=max_rounds, seed=args.seed, precision=args.precision)
    else:
        eng = FedEngine([prob], data.X_train, data.y_train, data.X_test, data.y_test, device=args.device, max_rounds=max_rounds,
                        seed=args.seed, precision=args.precision)
    if args.resume:
        eng.load_checkpoint(args.resume)
    log = flog.RunLog(args.jsonl if rank == 0 else None, quiet=args.quiet or rank != 0)   # one writer per file
    t0 = time.perf_counter()
    h = eng.run(args.rounds)          # ends with check_errors(): a timed-out peer raises here
    dt = time.perf_counter() - t0
    R = int(h["rounds"][0])
    for r in range(R):
        log.lines(flog.custom_round_lines(r + 1, h["local"][r], h["global"][0, r]))
        log.record(round=r + 1, global_metrics=h["global"][0, r], local_metrics=h["local"][r])
    if bool(h["stopped"][0]):
        log.lines(flog.custom_stop_lines(args.patience, R))
    if rank == 0:
        test = eng.evaluate_global("test")[0]
        print(f"\n[fedb200] {R} rounds in {dt:.3f} s ({R / max(dt, 1e-9):.1f} rounds/s, {eng.precision_name}); "
              f"held-out accuracy of the global model: {test['accuracy']:.4f}", flush=True)
        log.record(summary=True, rounds=R, seconds=dt, test=test)
    if args.checkpoint:
        eng.save_checkpoint(args.checkpoint, resume=True)    # every rank: the optimizer state is gathered; rank 0 writes
    log.close()
    return 0


def _world():
    return int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))


def cmd_sklearn(args) -> int:
    from .data.partition import contiguous
    from .fed import logging as flog
    from .fed.engine import FedEngine
    from .fed.problem import FedProblem
    from .fed.sklearn_mode import run_sklearn_rounds
    data = _load(args, with_mean=False)                                      # S:184 StandardScaler(with_mean=False)
    shards = contiguous(len(data.X_train), args.clients)                     # S:17-22
    prob = FedProblem.sklearn_like(data.n_features, shards, hidden=tuple(args.hidden), lr=args.lr, max_iter=args.max_iter)
    prob = prob.with_hyper(shuffle_stream=args.shuffle)
    world, rank = _world()
    if world > 1:
        # client-parallel over the GPUs: every rank fits its own clients, the coefficient mean (S:108-122) and the pooled
        # counts (S:126-134) cross NVLink inside the aggregation kernel
        import torch
        from .parallel.dist_engine import DistFedEngine, GlooFedEngine, init_distributed
        on_gpu = torch.cuda.is_available() and args.device != "cpu"
        init_distributed("nccl" if on_gpu else "gloo")
        Engine = DistFedEngine if on_gpu else GlooFedEngine
        # (a local fit runs for up to max_iter epochs before a rank reaches the aggregation: generous barrier timeout)
        eng = Engine([prob], data.X_train, data.y_train, data.X_test, data.y_test, rank=rank, world=world, max_rounds=args.rounds,
                     seed=args.seed, timeout_s=60.0)
    else:
        eng = FedEngine([prob], data.X_train, data.y_train, data.X_test, data.y_test, device=args.device, max_rounds=args.rounds,
                        seed=args.seed)
    log = flog.RunLog(args.jsonl if rank == 0 else None, quiet=args.quiet or rank != 0)
    t0 = time.perf_counter()
    h = run_sklearn_rounds(eng, args.rounds, args.policy, log)
    dt = time.perf_counter() - t0
    test = eng.evaluate_global("test")[0]
    if rank == 0:
        print(f"\n[fedb200] {args.rounds} rounds in {dt:.3f} s, policy={args.policy}; epochs per round and client"
              f"{' (this rank)' if world > 1 else ''}:\n{h['n_iter']}\nheld-out accuracy of the averaged model: {test['accuracy']:.4f}", flush=True)
        if args.checkpoint:
            eng.save_checkpoint(args.checkpoint)
    log.close()
    return 0


def cmd_tune(args) -> int:
    from .fed import logging as flog
    from .fed.tune import HIDDEN_GRID, LR_GRID, accuracy_table, run_grid
    from .models.checkpoint import save_checkpoint
    import torch
    data = _load(args, with_mean=False)                                      # H:164
    hg, lg = (HIDDEN_GRID[:3], LR_GRID[:3]) if args.small else (HIDDEN_GRID, LR_GRID)
    world, rank = _world()
    if world > 1:
        from .parallel.dist_engine import init_distributed
        on_gpu = torch.cuda.is_available() and args.device != "cpu"
        init_distributed("nccl" if on_gpu else "gloo")
    log = flog.RunLog(args.jsonl if rank == 0 else None, quiet=args.quiet or rank != 0)
    if not args.quiet and rank == 0:
        print(f"Training Round 1...\n{'-' * 50}")
    t0 = time.perf_counter()
    out = run_grid(data, args.clients, hg, lg, tuple(args.max_iter), device=args.device, log=log, seed=args.seed, rank=rank, world=world,
                   shuffle_stream=args.shuffle)
    dt = time.perf_counter() - t0
    if rank == 0:
        print(f"\n[fedb200] {len(out['results'])} federations in {dt:.2f} s on {world} GPU(s); best {out['best_params']} -> "
              f"{out['best_metrics']}", flush=True)
        if len(args.max_iter) == 1:
            np.set_printoptions(precision=3, suppress=True, linewidth=160)
            print(accuracy_table(out["results"], hg, lg))
        if args.checkpoint:
            spec = out["best"]["spec"]
            slab = torch.zeros(spec.slab_size)
            spec.from_coefs_intercepts(*out["best"]["weights"], slab)
            save_checkpoint(args.checkpoint, spec, slab, extras=out["best_params"] | {"metrics": out["best_metrics"]})
    log.close()
    if world > 1:
        torch.distributed.barrier()
    return 0


def main(argv: Optional[List[str]] = None) -> int:
    args = build_parser().parse_args(argv)
    return {"custom": cmd_custom, "sklearn-avg": cmd_sklearn, "tune": cmd_tune}[args.cmd](args)


if __name__ == "__main__":
    raise SystemExit(main())
