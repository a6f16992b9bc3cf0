#!/usr/bin/env python
"""Run one of the three reference scripts UNMODIFIED (from baseline/_ref) at np = N ranks over the mpi4py stand-in
(baseline/ref_mpi, gloo underneath) and turn what it prints into JSON: per-round global metrics, the sklearn script's final
weight statistics, the grid search's 90-cell table and winner, and the wall time of the script's own main().

    python baseline/run_reference.py custom|sklearn|grid [--np 4] [--rounds R] [--out result.json] [--echo]

The only source substitutions happen at load time and only where the shipped script cannot run as it is:
`Pakistani_Diabetes_Dataset.csv` -> `balanced_income_data.csv` and the label `'Outcome'` -> `'income'` in the custom and the
grid script.  `--rounds R` additionally replaces the literal round count in the script's call of train_and_evaluate (300 / 5 /
1); without it the literal stands.  Nothing of this repository's framework is imported: the point is the reference's own
code path, timed and parsed.
"""
from __future__ import annotations

import argparse
import contextlib
import io
import json
import os
import re
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
SCRIPTS = {"custom": "FL_CustomMLPCLassifierImplementation_Multiple_Rounds.py", "sklearn": "FL_SkLearn_MLPClassifier_Limitation.py",
           "grid": "hyperparameters_tuning.py"}
_NUM = r"(?:np\.float64\()?([-+0-9.eE]+)\)?"


def ensure_ref() -> bool:
    if all(os.path.exists(os.path.join(REF, f)) for f in list(SCRIPTS.values()) + ["balanced_income_data.csv"]):
        return True
    if os.path.exists("/root/reference"):
        subprocess.run([sys.executable, os.path.join(HERE, "install_ref.py")], capture_output=True)
    return all(os.path.exists(os.path.join(REF, f)) for f in SCRIPTS.values())


def load_source(which: str, rounds=None) -> str:
    src = open(os.path.join(REF, SCRIPTS[which])).read()
    if which in ("custom", "grid"):      # R4: the file / label these two name are not shipped
        src = src.replace("Pakistani_Diabetes_Dataset.csv", "balanced_income_data.csv").replace("'Outcome'", "'income'")
    if rounds is not None:
        src, n = re.subn(r"(train_and_evaluate\(comm,\s*rounds=)\d+(\))", rf"\g<1>{int(rounds)}\g<2>", src)
        assert n == 1, "could not find the script's train_and_evaluate(comm, rounds=...) call"
    return src


def parse(which: str, text: str) -> dict:
    out: dict = {}
    if which == "custom":
        pat = re.compile(r"Global Metrics \(Round (\d+)\): \[accuracy: " + _NUM + ", precision: " + _NUM + ", recall: " + _NUM + ", f1: " + _NUM)
        out["rounds"] = [{"round": int(m.group(1)), "accuracy": float(m.group(2)), "precision": float(m.group(3)), "recall": float(m.group(4)),
                          "f1": float(m.group(5))} for m in pat.finditer(text)]
        m = re.search(r"Training stopped early at round (\d+)", text)
        out["stopped_at"] = int(m.group(1)) if m else None
    elif which == "sklearn":
        pat = re.compile(r"Global Metrics for Round (\d+):\s*\n\s*Accuracy: " + _NUM + r"\s*\n\s*Precision: " + _NUM + r"\s*\n\s*Recall: " + _NUM +
                         r"\s*\n\s*F1: " + _NUM)
        out["rounds"] = [{"round": int(m.group(1)), "accuracy": float(m.group(2)), "precision": float(m.group(3)), "recall": float(m.group(4)),
                          "f1": float(m.group(5))} for m in pat.finditer(text)]
        pat = re.compile(r"Layer (\d+) - Shape: \(([^)]*)\)\s*\nMean: " + _NUM + ", Std: " + _NUM)
        out["weight_stats"] = [{"layer": int(m.group(1)), "shape": [int(x) for x in m.group(2).replace(" ", "").split(",") if x],
                                "mean": float(m.group(3)), "std": float(m.group(4))} for m in pat.finditer(text)]
    else:
        pat = re.compile(r"Global Metrics \(Hidden Layers: \(([^)]*)\), LR: ([0-9.eE+-]+)\): \{'accuracy': " + _NUM + ", 'precision': " + _NUM +
                         ", 'recall': " + _NUM + ", 'f1': " + _NUM)
        out["table"] = [{"hidden_layer_sizes": [int(x) for x in m.group(1).replace(" ", "").split(",") if x], "learning_rate": float(m.group(2)),
                         "accuracy": float(m.group(3)), "precision": float(m.group(4)), "recall": float(m.group(5)), "f1": float(m.group(6))}
                        for m in pat.finditer(text)]
        m = re.search(r"Best Global Hyperparameters: \{'hidden_layer_sizes': \(([^)]*)\), 'learning_rate': ([0-9.eE+-]+)\}", text)
        if m:
            out["best"] = {"hidden_layer_sizes": [int(x) for x in m.group(1).replace(" ", "").split(",") if x], "learning_rate": float(m.group(2))}
        m = re.search(r"Best Global Metrics: \{'accuracy': " + _NUM, text)
        if m:
            out["best_accuracy"] = float(m.group(1))
    return out


def worker(args) -> int:
    """One rank: exec the script as __main__ from inside baseline/_ref."""
    sys.path.insert(0, os.path.join(HERE, "ref_mpi"))       # serves `from mpi4py import MPI`
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    src = load_source(args.which, args.rounds)
    rank = int(os.environ.get("RANK", "0"))
    os.chdir(REF)
    buf = io.StringIO()

    class Tee(io.TextIOBase):
        def write(self, s):
            buf.write(s)
            if args.echo:
                sys.__stdout__.write(s)
            return len(s)

        def flush(self):
            if args.echo:
                sys.__stdout__.flush()

    import warnings
    warnings.filterwarnings("ignore")
    code = compile(src, os.path.join(REF, SCRIPTS[args.which]), "exec")
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(Tee()):
        try:
            exec(code, {"__name__": "__main__", "__file__": os.path.join(REF, SCRIPTS[args.which])})
        except SystemExit:
            pass
    wall = time.perf_counter() - t0
    if rank == 0:
        rec = {"script": SCRIPTS[args.which], "np": int(os.environ.get("WORLD_SIZE", "1")), "wall_s": wall, "rounds_override": args.rounds}
        rec.update(parse(args.which, buf.getvalue()))
        js = json.dumps(rec)
        if args.out:
            with open(args.out, "w") as f:
                f.write(js)
        print(js, file=sys.__stdout__, flush=True)
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("which", choices=list(SCRIPTS))
    ap.add_argument("--np", type=int, default=4)
    ap.add_argument("--rounds", type=int, default=None)
    ap.add_argument("--out", default=None)
    ap.add_argument("--echo", action="store_true", help="also pass the script's own output through")
    ap.add_argument("--port", type=int, default=29655)
    args = ap.parse_args()
    if not ensure_ref():
        print(json.dumps({"unavailable": "baseline/_ref is empty and /root/reference is not present to copy from"}))
        return 0
    if "RANK" in os.environ or args.np == 1:
        return worker(args)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.np), "--master-addr", "127.0.0.1",
           "--master-port", str(args.port), "--", os.path.abspath(__file__), args.which, "--np", str(args.np)]
    if args.rounds is not None:
        cmd += ["--rounds", str(args.rounds)]
    if args.out:
        cmd += ["--out", os.path.abspath(args.out)]
    if args.echo:
        cmd += ["--echo"]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("REF_CUDA_VISIBLE_DEVICES", ""), OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, env=env, capture_output=True, text=True)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    if r.returncode != 0 or not lines:
        sys.stderr.write(r.stderr[-4000:])
        return r.returncode or 1
    print(lines[-1])
    return 0


if __name__ == "__main__":
    sys.exit(main())
