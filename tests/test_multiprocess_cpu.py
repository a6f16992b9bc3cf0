"""Multi-process paths on CPU (gloo, world_size 2/3): the CPU-plumbing federation of BASELINE config 1
and the reference arm of bench.py (unmodified reference script over the mpi4py stand-in)."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(nproc, port, args, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(port)] + args
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)


def test_three_rank_cpu_federation_matches_single_process(tmp_path):
    from fedb200.cli import main as cli_main
    single, multi = str(tmp_path / "s.jsonl"), str(tmp_path / "m.jsonl")
    common = ["custom", "--clients", "3", "--rounds", "4", "--device", "cpu", "--quiet", "--partition", "contiguous", "--synthetic-rows", "900"]
    assert cli_main(common + ["--jsonl", single]) == 0
    out = _torchrun(3, 29631, ["-m", "fedb200"] + common + ["--jsonl", multi])
    assert out.returncode == 0, out.stderr[-2000:]
    a = [json.loads(l) for l in open(single)][:4]
    b = [json.loads(l) for l in open(multi)][:4]
    for ra, rb in zip(a, b):
        np.testing.assert_allclose(ra["global_metrics"], rb["global_metrics"], atol=1e-6)
        np.testing.assert_allclose(ra["local_metrics"], rb["local_metrics"], atol=1e-6)


def test_reference_arm_runs_unmodified_script_on_two_ranks():
    if not os.path.exists(os.path.join(ROOT, "baseline", "_ref", "FL_CustomMLPCLassifierImplementation_Multiple_Rounds.py")) \
            and not os.path.exists("/root/reference"):
        import pytest
        pytest.skip("reference scripts not available")
    out = _torchrun(2, 29632, ["bench.py", "--impl", "reference", "--gpus", "2", "--steps", "4", "--warmup", "1"])
    assert out.returncode == 0, out.stderr[-2000:]
    rec = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert rec["impl"] == "reference" and rec["n_gpus"] == 2 and rec["value"] > 0
    assert 0.4 < rec["final_global_accuracy_train_shards"] <= 1.0


def test_client_placement_and_symmetric_layout():
    """Host-side bookkeeping of the multi-GPU engines: every client lands on exactly one rank in contiguous blocks, and the
    symmetric buffer's regions are 256-byte aligned and do not overlap."""
    from fedb200.parallel.dist_engine import place_clients
    from fedb200.parallel.p2p import SymLayout, C2
    for n, world in [(8, 8), (3, 2), (128, 8), (1024, 8), (5, 4)]:
        blocks = place_clients(n, world)
        flat = [k for b in blocks for k in b]
        assert flat == list(range(n)) and len(blocks) == world
        assert max(len(b) for b in blocks) - min(len(b) for b in blocks) <= 1
    lay = SymLayout(global_total=11360, conf_slots=3)
    assert lay.partial_off == 0 and lay.conf_off % 256 == 0 and lay.flag_off % 256 == 0
    assert lay.conf_off >= 2 * lay.global_total * 4
    assert lay.flag_off >= lay.conf_off + 2 * 8 * lay.conf_slots * C2 * 4
    assert lay.nbytes >= lay.flag_off + 24 * 8          # flag words A, B (fused tail) and C (NVLS), one per source rank


def test_grid_placement_is_balanced_and_complete():
    """Longest-processing-time placement of the 90 grid federations over the ranks (fed/tune.py)."""
    from fedb200.fed.tune import place_federations
    costs = [float(c) for c in ([700] * 9 + [1500] * 9 + [3300] * 9 + [6500] * 9 + [6600] * 9 + [11600] * 9 + [21500] * 9 +
                                 [42000] * 9 + [86400] * 9 + [84000] * 9)]
    for world in (1, 2, 4, 8):
        parts = place_federations(costs, world)
        assert sorted(i for p in parts for i in p) == list(range(90))
        loads = [sum(costs[i] for i in p) for p in parts]
        assert max(loads) <= 1.25 * (sum(costs) / world) + max(costs) * (world > 4)


def test_two_rank_tune_returns_the_single_process_table(tmp_path):
    """`tune` under torchrun (federation-parallel, H:83-106 over all ranks): same table, same winner, bit for bit."""
    from fedb200.cli import main as cli_main
    single, multi = str(tmp_path / "s.jsonl"), str(tmp_path / "m.jsonl")
    common = ["tune", "--clients", "2", "--max-iter", "3", "--small", "--device", "cpu", "--quiet", "--synthetic-rows", "400"]
    assert cli_main(common + ["--jsonl", single]) == 0
    out = _torchrun(2, 29633, ["-m", "fedb200"] + common + ["--jsonl", multi])
    assert out.returncode == 0, out.stderr[-2000:]
    a = [json.loads(l) for l in open(single)]
    b = [json.loads(l) for l in open(multi)]
    assert len(a) == len(b) == 9
    for ra, rb in zip(a, b):
        assert ra["hidden"] == rb["hidden"] and ra["lr"] == rb["lr"]
        assert ra["global_metrics"] == rb["global_metrics"] and ra["n_iter"] == rb["n_iter"]


def test_two_rank_sklearn_avg_matches_single_process(tmp_path):
    """`sklearn-avg` under torchrun (client-parallel: coefficient mean S:108-122 and pooled counts S:126-134 across ranks)."""
    from fedb200.cli import main as cli_main
    single, multi = str(tmp_path / "s.jsonl"), str(tmp_path / "m.jsonl")
    common = ["sklearn-avg", "--clients", "4", "--rounds", "2", "--max-iter", "3", "--hidden", "8", "6", "--device", "cpu", "--quiet",
              "--synthetic-rows", "800"]
    assert cli_main(common + ["--jsonl", single]) == 0
    out = _torchrun(2, 29634, ["-m", "fedb200"] + common + ["--jsonl", multi])
    assert out.returncode == 0, out.stderr[-2000:]
    a = [json.loads(l) for l in open(single)]
    b = [json.loads(l) for l in open(multi)]
    assert len(a) == len(b) == 2
    for ra, rb in zip(a, b):
        np.testing.assert_allclose(ra["global_metrics"], rb["global_metrics"], atol=1e-6)
        np.testing.assert_allclose(ra["local_metrics"], rb["local_metrics"], atol=1e-6)
