"""Real multi-GPU run of the P2P aggregation (needs >= 2 GPUs; skipped otherwise): launches
bench.py under torchrun on 2 ranks and checks it against the single-GPU engine."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def test_two_gpu_bench_runs_and_learns():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "60", "--warmup", "5"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    rec = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert rec["n_gpus"] == 2 and rec["value"] > 0 and any("k_fedavg_p2p" in k for k in rec["kernels"])
    assert rec["final_global_accuracy_train_shards"] > 0.8
