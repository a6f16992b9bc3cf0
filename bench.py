#!/usr/bin/env python
"""Headline benchmark: FedAvg client-samples/sec (and rounds/sec) of the reference's custom-MLP
federation (BASELINE.json config 3: income data, MLP 14-50-200-2, Adam(0.004)+StepLR(30,0.5),
one full-batch local step per round, local eval + weighted P/R/F1, sample-weighted FedAvg) with
ONE CLIENT PER GPU and a fixed 1000-row shard per client (the np=8 shard size), i.e. weak
scaling: at --gpus 8 this is exactly "8 clients, one per B200".

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one complete federated round: local train step for every client, local evaluation and
metrics, early-stop bookkeeping, aggregation and write-back.  Nothing is skipped or cached.
Timing: CUDA events on the launching stream, L2 flushed (256 MiB memset) before every timed
step, max over ranks.  See DESIGN.md "Benchmark" for what each JSON key means.
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS_PER_CLIENT = 1000
BASELINE_SAMPLES_PER_SEC = 0.53e6   # reference C:, CPU, upper end of 0.45-0.53 M/s
BASELINE_ROUNDS_PER_SEC = 66.0      # upper end of 55-66 rounds/s


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "nccl"],
                    help="ours | reference (unmodified scripts) | nccl (stock torch + NCCL all_reduce baseline)")
    ap.add_argument("--clients-per-gpu", type=int, default=None, help="default: 1 (income) / 128 (wide)")
    ap.add_argument("--config", default="income", choices=["income", "wide"],
                    help="income = BASELINE config 3 (the headline; default); wide = BASELINE config 5 (4 x 4096 hidden, "
                         "synthetic income-shaped non-IID data, 1024 rows per client, SGD) — ours only")
    ap.add_argument("--width", type=int, default=4096)
    ap.add_argument("--precision", default=None, help="ours: fp32 (CUDA cores) | tc32 (tcgen05, split fp16 operands, fp32-grade; the default) | bf16 (tcgen05, bf16 operands)")
    ap.add_argument("--no-flush", action="store_true", help="debug: skip the L2 flush between timed steps")
    a = ap.parse_args()
    if a.clients_per_gpu is None:
        a.clients_per_gpu = 128 if a.config == "wide" else 1
    return a


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def base_record(args, world, impl):
    if args.config == "wide":
        rows = 1024
        return {
            "metric": "fedavg_client_samples_per_sec", "unit": "client-samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "data": "synthetic income-shaped rows (label-skewed clients), random-init weights", "impl": impl,
            "config": {"model": f"mlp-14-{args.width}x4-2 softmax-CE, SGD(lr=0.05), 1 full-batch local step/round, local eval + "
                                "weighted P/R/F1, sample-weighted FedAvg (BASELINE config 5)",
                       "clients": world * args.clients_per_gpu, "clients_per_gpu": args.clients_per_gpu, "rows_per_client": rows,
                       "global_batch": world * args.clients_per_gpu * rows, "seq_len": None,
                       "parallelism": f"federated-dp{world} (one process per GPU)",
                       "l2": "working set (weights of all clients) far larger than L2; no flush"},
        }
    return {
        "metric": "fedavg_client_samples_per_sec", "unit": "client-samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "data": "bundled income CSV (real rows), random-init weights",
        "impl": impl,
        "config": {"model": "mlp-14-50-200-2 softmax-CE, Adam(lr=0.004)+StepLR(30,0.5), 1 full-batch local step/round, "
                            "local eval + weighted P/R/F1, sample-weighted FedAvg",
                   "clients": world * args.clients_per_gpu, "clients_per_gpu": args.clients_per_gpu,
                   "rows_per_client": ROWS_PER_CLIENT, "global_batch": world * args.clients_per_gpu * ROWS_PER_CLIENT,
                   "seq_len": None, "parallelism": f"federated-dp{world} (one process per GPU)",
                   "l2": "flushed (256 MiB memset) before every timed step"},
    }


# ----------------------------------------------------------------------------------------------
def run_ours(args):
    import numpy as np
    import torch
    rank, local_rank, world = dist_env()
    torch.cuda.set_device(local_rank)
    from fedb200 import FedProblem, TrainHyper, prepare, synthetic_income_like
    from fedb200.utils.clocks import ClockSampler
    rec0 = base_record(args, world, "ours")
    cpg = args.clients_per_gpu
    n_clients = world * cpg
    extra = {}
    if args.config == "wide":
        rows_per_client = 1024
        data = synthetic_income_like(40000, seed=1)
        rng = np.random.default_rng(0)
        pos, neg = np.flatnonzero(data.y_train == 1), np.flatnonzero(data.y_train == 0)
        rows = []
        for c in range(n_clients):          # label-skewed (non-IID) clients: class mix ~ Beta(0.5, 0.5)
            k = int(rows_per_client * rng.beta(0.5, 0.5))
            rows.append(np.concatenate([rng.choice(pos, k), rng.choice(neg, rows_per_client - k)]))
        extra = dict(hidden=(args.width,) * 4, hyper=TrainHyper(opt="sgd", lr=0.05, lr_step_size=0), shared_init=True)
    else:
        rows_per_client = ROWS_PER_CLIENT
        data = prepare(with_mean=True)
        # client c owns rows [c*1000, (c+1)*1000) of the (already shuffled) train split, wrapping if needed
        rows = [(np.arange(ROWS_PER_CLIENT) + c * ROWS_PER_CLIENT) % len(data.X_train) for c in range(n_clients)]
    if world > 1:
        from fedb200.parallel.dist_engine import DistFedEngine, init_distributed
        init_distributed()
        eng = DistFedEngine.custom_income(data, rows, rank, world, max_rounds=3 * args.steps + args.warmup + 32,
                                          precision=args.precision, **extra)
    else:
        from fedb200.parallel.dist_engine import make_local_engine
        eng = make_local_engine(data, rows, max_rounds=3 * args.steps + args.warmup + 32, precision=args.precision, **extra)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    flush = None if (args.no_flush or args.config == "wide") else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    K, W = args.steps, args.warmup
    # ---- device-timed rounds ------------------------------------------------------------------
    for _ in range(W):
        eng.step_graph()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    with ClockSampler(local_rank) as clk:
        barrier()
        for s, e in ev:
            if flush is not None:
                flush.zero_()
            s.record()
            eng.step_graph()
            e.record()
        barrier()
        step_ms = sorted(s.elapsed_time(e) for s, e in ev)
        ms_flushed = sum(step_ms)
        # the metric both arms report at the same round index: after exactly warmup + steps rounds (the timing loops below train on)
        acc_round = W + K
        acc_at_round = float(eng.history()["global"][0][acc_round - 1][0])
        # steady state, back to back, no flush (reported separately, never as `value`)
        s0, e0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        s0.record()
        for _ in range(K):
            eng.step_graph()
        e0.record()
        barrier()
        ms_warm = s0.elapsed_time(e0)
        # ---- end to end through the public API: H2D inputs + round + D2H result, every step ----
        xh, yh = eng.pinned_train_inputs()
        for _ in range(min(W, 10)):
            eng.step_e2e(xh, yh)
        s1, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        s1.record()
        last = None
        for _ in range(K):
            last = eng.step_e2e(xh, yh)
        e1.record()
        barrier()
        ms_e2e = s1.elapsed_time(e1)
    t = torch.tensor([ms_flushed, ms_warm, ms_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    ms_flushed, ms_warm, ms_e2e = (float(x) for x in t.cpu())
    samples = n_clients * rows_per_client
    rec = rec0
    value = samples * K / (ms_flushed / 1e3)
    h2d, d2h = eng.e2e_bytes()
    hist = eng.history()
    rec.update({
        "value": value, "ms_per_step": ms_flushed / K,
        "vs_baseline": (value / BASELINE_SAMPLES_PER_SEC) if args.config == "income" else None,
        "dtype": eng.dtype_label, "precision": eng.precision_name, "rounds_per_sec": K / (ms_flushed / 1e3),
        "vs_baseline_rounds_per_sec": (K / (ms_flushed / 1e3)) / BASELINE_ROUNDS_PER_SEC,
        "warm_no_flush": {"ms_per_step": ms_warm / K, "client_samples_per_sec": samples * K / (ms_warm / 1e3),
                          "rounds_per_sec": K / (ms_warm / 1e3)},
        "e2e": {"value": samples * K / (ms_e2e / 1e3), "unit": "client-samples/s", "ms_per_step": ms_e2e / K,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "h2d_by": "round kernel (loads from pinned host memory, written through to the device copy)"
                          if eng.pulls_host_inputs() else "copy engine (cudaMemcpyAsync nodes in the step's graph)",
                "how": "per step, through FedEngine.step_e2e: the step's rows + labels travel from pinned host memory to the device "
                       "(single-launch rounds: the round kernel reads them over the host link, one 64-byte request per row, and writes "
                       "them through to the device-resident X / y; other paths: two cudaMemcpyAsync nodes), the round runs, the "
                       "federation state (stop flag + 4 global metrics) is stored by the round kernel into mapped pinned host memory "
                       "(a memcpy node + stream synchronise on the non-fused paths) together with four self-validating 16-byte result chunks "
                       "{round counter, stop flag, metric}; the host polls the chunks and reads the result"},
        "gpu_launches": K * eng.launches_per_round, "launches_per_step": eng.launches_per_round,
        "kernels": eng.kernel_names, "clocks": clk.summary(),
        "step_ms": {"min": step_ms[0], "median": step_ms[len(step_ms) // 2], "max": step_ms[-1],
                    "p90": step_ms[min(len(step_ms) - 1, (9 * len(step_ms)) // 10)]},
        "global_accuracy_train_shards": {"round": acc_round, "value": acc_at_round},
        "final_global_accuracy_train_shards": float(hist["last"][0][0]), "rounds_run": int(hist["rounds"][0]),
    })
    eng.check_errors()    # a timed-out peer or a stalled GEMM pipeline raises here instead of passing silently
    if rank == 0:
        print(json.dumps(rec))
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


# ----------------------------------------------------------------------------------------------
def run_reference(args):
    rank, local_rank, world = dist_env()
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    script = os.path.join(ref_dir, "FL_CustomMLPCLassifierImplementation_Multiple_Rounds.py")
    if not os.path.exists(script) and os.path.exists("/root/reference"):
        import subprocess
        subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "install_ref.py")], capture_output=True)
    if not os.path.exists(script):
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is empty: the reference is not pip-installable "
                              "(no setup.py/pyproject.toml) and /root/reference is not present to copy from"}))
        return
    # the reference picks torch.device("cuda") = current device; give every rank its own GPU
    os.environ.setdefault("CUDA_VISIBLE_DEVICES", str(local_rank))
    sys.path.insert(0, os.path.join(ROOT, "baseline", "ref_mpi"))  # serves `from mpi4py import MPI`
    import importlib.util
    import numpy as np
    import torch
    from fedb200 import prepare  # data preparation only (bitwise equal to the reference's sklearn pipeline)
    spec = importlib.util.spec_from_file_location("ref_custom", script)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)  # unmodified source; main() is not run
    from mpi4py import MPI
    comm = MPI.COMM_WORLD
    data = prepare(with_mean=True)
    n = world * ROWS_PER_CLIENT
    X, y = data.X_train[:n], data.y_train[:n]
    np.random.seed(1234 + rank)
    torch.manual_seed(1234 + rank)
    trainer = ref.FederatedMLPLearning(X, y, rank, world)  # rank r gets 1000 rows (C:48-61)
    cuda = torch.cuda.is_available()
    sync = (lambda: torch.cuda.synchronize()) if cuda else (lambda: None)
    K, W = args.steps, args.warmup
    from fedb200.utils.clocks import ClockSampler
    with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):
        trainer.train_and_evaluate(comm, rounds=W, termination_patience=10 ** 9)
        comm.Barrier(); sync()
        with ClockSampler(local_rank) as clk:  # NVML indexes physical GPUs
            t0 = time.perf_counter()
            hist = trainer.train_and_evaluate(comm, rounds=K, termination_patience=10 ** 9)
            sync(); comm.Barrier()
            dt = time.perf_counter() - t0
    dts = comm.gather(dt, root=0)
    if rank == 0:
        dt = max(dts)
        rec = base_record(args, world, "reference")
        samples = world * ROWS_PER_CLIENT
        value = samples * K / dt
        rec["arm_notes"] = {"timing": "stock reference path: wall clock around train_and_evaluate, max over ranks; no L2 flush",
                            "transport": f"{world} ranks = {world} clients over an mpi4py stand-in (gloo), device={'cuda' if cuda else 'cpu'}"}
        rec.update({
            "value": value, "ms_per_step": dt * 1e3 / K, "vs_baseline": value / BASELINE_SAMPLES_PER_SEC, "dtype": "fp32",
            "rounds_per_sec": K / dt,
            "e2e": {"value": value, "unit": "client-samples/s", "note": "the reference has only an end-to-end path: it re-uploads the "
                    "shard (C:65-66, C:77-78) and downloads predictions (C:83) and weights (C:94) every round",
                    "h2d_bytes_per_step": 2 * (ROWS_PER_CLIENT * 14 * 4 + ROWS_PER_CLIENT * 8) + 11352 * 4,
                    "d2h_bytes_per_step": ROWS_PER_CLIENT * 8 + 11352 * 4},
            "gpu_launches": None, "clocks": clk.summary(),
            "global_accuracy_train_shards": {"round": W + K, "value": float(hist["accuracy"][-1]) if hist["accuracy"] else None},
            "final_global_accuracy_train_shards": float(hist["accuracy"][-1]) if hist["accuracy"] else None,
        })
        print(json.dumps(rec))


# ----------------------------------------------------------------------------------------------
def run_nccl(args):
    """Stock-PyTorch (cuBLAS + NCCL all_reduce) FedAvg of the same model and config."""
    import numpy as np
    import torch
    rank, local_rank, world = dist_env()
    torch.cuda.set_device(local_rank)
    sys.path.insert(0, os.path.join(ROOT, "baseline", "nccl_cublas"))
    from fedavg_torch import TorchFedAvg
    from fedb200 import prepare
    from fedb200.utils.clocks import ClockSampler
    if world > 1:
        from fedb200.parallel.dist_engine import init_distributed
        init_distributed()
    data = prepare(with_mean=True)
    cpg = args.clients_per_gpu
    rows = [(np.arange(ROWS_PER_CLIENT) + c * ROWS_PER_CLIENT) % len(data.X_train) for c in range(rank * cpg, (rank + 1) * cpg)]
    fed = TorchFedAvg(data.X_train, data.y_train, rows, world * cpg * ROWS_PER_CLIENT)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    K, W = args.steps, args.warmup
    for _ in range(W):
        fed.round()
    barrier()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    with ClockSampler(local_rank) as clk:
        for s, e in ev:
            flush.zero_()
            s.record()
            fed.round()
            e.record()
        barrier()
    ms = torch.tensor([sum(s.elapsed_time(e) for s, e in ev)], dtype=torch.float64, device="cuda")
    if world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    ms = float(ms.item())
    if rank == 0:
        rec = base_record(args, world, "nccl_cublas_baseline")
        samples = world * cpg * ROWS_PER_CLIENT
        value = samples * K / (ms / 1e3)
        rec["arm_notes"] = {"transport": f"{world} ranks, torch.nn + torch.optim + NCCL all_reduce FedAvg"}
        rec.update({"value": value, "ms_per_step": ms / K, "vs_baseline": value / BASELINE_SAMPLES_PER_SEC, "dtype": "fp32",
                    "rounds_per_sec": K / (ms / 1e3), "gpu_launches": None, "clocks": clk.summary(),
                    "final_global_accuracy_train_shards": float(fed.last_metrics[0])})
        print(json.dumps(rec))
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    {"reference": run_reference, "nccl": run_nccl, "ours": run_ours}[a.impl](a)
