"""Real multi-GPU run of the P2P aggregation (needs >= 2 GPUs; skipped otherwise): launches
bench.py under torchrun on 2 ranks and checks it against the single-GPU engine."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _bench(port, env_extra=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "60", "--warmup", "5"]
    env = dict(os.environ, **(env_extra or {}))
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    return json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])


def _dist_run(tmp_path, name, port, n_clients, rows, rounds, precision, env_extra=None, nproc=2):
    out = str(tmp_path / f"{name}.npz")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "_dist_worker.py"), out, str(n_clients), str(rows), str(rounds), precision]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, **(env_extra or {})))
    assert r.returncode == 0, r.stderr[-3000:]
    import numpy as np
    return np.load(out)


@pytest.mark.parametrize("precision", ["tc32", "bf16"])
def test_two_gpu_global_model_is_bitwise_the_single_gpu_one(tmp_path, precision):
    """Two clients on two GPUs (one launch per rank per round, aggregation inside the round kernel over NVLink peer memory)
    against the same two clients on ONE GPU: the global model after 12 rounds, the metric history, and every rank's copy
    of the global model are bit-identical.  Also with the stand-alone aggregation kernel."""
    import numpy as np
    import torch
    sys.path.insert(0, ROOT)
    from fedb200 import FedEngine, FedProblem, prepare
    rows, rounds = 1000, 12
    data = prepare(with_mean=True)
    shards = [np.arange(c * rows, (c + 1) * rows, dtype=np.int32) for c in range(2)]
    one = FedEngine([FedProblem.custom(data.n_features, data.n_classes, shards, patience=0)], data.X_train, data.y_train,
                    data.X_test, data.y_test, device="cuda", max_rounds=rounds + 4, precision=precision)
    h = one.run(rounds)
    want = one.global_.cpu().numpy()
    fused = _dist_run(tmp_path, "fused", 29551, 2, rows, rounds, precision)
    assert int(fused["launches"]) == 1 and "p2p-fedavg" in str(fused["kernels"][0]) and str(fused["precision"]) == precision
    assert np.array_equal(fused["global_"], want)
    assert np.array_equal(fused["globals_all"][0], fused["globals_all"][1])
    assert np.array_equal(fused["hist"], h["global"]) and np.array_equal(fused["local"], h["local"])
    sep = _dist_run(tmp_path, "sep", 29552, 2, rows, rounds, precision, {"FEDB200_NO_FUSED_P2P": "1"})
    assert int(sep["launches"]) > 1
    assert np.array_equal(sep["global_"], want) and np.array_equal(sep["hist"], h["global"])


def test_two_gpu_many_clients_per_rank_in_one_launch(tmp_path):
    """Six clients on two GPUs (three per rank): one launch per rank per round (the rank's last cluster folds its clients into
    the rank partial and runs the cross-GPU exchange) — bit-identical to the local-round + stand-alone aggregation pair."""
    import numpy as np
    fused = _dist_run(tmp_path, "many_fused", 29553, 6, 700, 10, "tc32")
    assert int(fused["launches"]) == 1 and "p2p-fedavg" in str(fused["kernels"][0])
    sep = _dist_run(tmp_path, "many_sep", 29554, 6, 700, 10, "tc32", {"FEDB200_NO_FUSED_P2P": "1"})
    assert int(sep["launches"]) > 1
    assert np.array_equal(fused["global_"], sep["global_"]) and np.array_equal(fused["globals_all"][0], fused["globals_all"][1])
    assert np.array_equal(fused["hist"], sep["hist"]) and np.array_equal(fused["local"], sep["local"])
    assert fused["hist"][0, -1, 0] > 0.6


def test_two_gpu_bench_fused_and_standalone_aggregation_agree():
    fused = _bench(29541)
    assert fused["n_gpus"] == 2 and fused["value"] > 0 and any("p2p-fedavg" in k for k in fused["kernels"])
    assert fused["final_global_accuracy_train_shards"] > 0.8
    # the stand-alone aggregation kernel performs the same arithmetic in the same order: identical trajectory
    sep = _bench(29542, {"FEDB200_NO_FUSED_P2P": "1"})
    assert any("k_fedavg_p2p" in k for k in sep["kernels"])
    assert sep["final_global_accuracy_train_shards"] == fused["final_global_accuracy_train_shards"]
    assert sep["rounds_run"] == fused["rounds_run"]


def test_two_gpu_nvls_aggregation_matches_peer_loads():
    """FEDB200_AGG=nvls: the NVSwitch adds the ranks' partial sums (multimem.ld_reduce) instead of peer loads in rank order.
    Two addends commute, so with two ranks the trajectory must be identical to the P2P kernel's."""
    import torch
    sys.path.insert(0, ROOT)
    from fedb200.parallel import nvls
    if not nvls.supported(0):
        pytest.skip("NVLink multicast is not supported on this device / driver")
    ref = _bench(29543, {"FEDB200_NO_FUSED_P2P": "1"})
    try:
        got = _bench(29544, {"FEDB200_AGG": "nvls"})
    except AssertionError as e:                      # e.g. no fabric manager / multicast not permitted inside this container
        if "cuMulticast" in str(e) or "multicast" in str(e):
            pytest.skip("the driver refused to create a multicast object here: " + str(e)[-300:])
        raise
    assert any("nvls" in k for k in got["kernels"])
    assert got["rounds_run"] == ref["rounds_run"]
    assert got["final_global_accuracy_train_shards"] == ref["final_global_accuracy_train_shards"]
