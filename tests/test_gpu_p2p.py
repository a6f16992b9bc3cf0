"""Cross-GPU aggregation protocol (csrc/fedb_p2p.cu).  On one GPU the ranks are emulated inside a
single cooperative launch; with >= 2 GPUs tests/test_multigpu.py runs the real thing."""
import numpy as np
import pytest
import torch

from fedb200 import FedEngine, FedProblem, TrainHyper, make_shards
from fedb200.parallel.dist_engine import EmulatedWorld, place_clients

pytestmark = pytest.mark.gpu


def test_placement():
    assert [list(b) for b in place_clients(8, 4)] == [[0, 1], [2, 3], [4, 5], [6, 7]]
    assert [len(b) for b in place_clients(5, 2)] == [3, 2]


@pytest.mark.parametrize("world,clients,precision", [(2, 2, "fp32"), (4, 4, "fp32"), (2, 6, "fp32"), (4, 8, "bf16")])
def test_emulated_ranks_equal_single_gpu(income, world, clients, precision):
    shards = make_shards("per_client_shuffle", 4000, clients, seed=7)
    probs = [FedProblem.custom(14, 2, shards, patience=0)]
    single = FedEngine(probs, income.X_train[:4000], income.y_train[:4000], income.X_test, income.y_test, device="cuda",
                       max_rounds=16, precision=precision)
    emu = EmulatedWorld(probs, income.X_train[:4000], income.y_train[:4000], income.X_test, income.y_test, world=world,
                        max_rounds=16, precision=precision)
    hs = single.run(6, use_graph=False)
    he = emu.run(6)
    # every rank ends with the same global model ...
    for r in emu.ranks[1:]:
        assert torch.equal(r.global_, emu.ranks[0].global_)
    # ... which is the single-GPU one (bit-exact with one client per rank: same summation order)
    if clients == world:
        assert torch.equal(emu.ranks[0].global_, single.global_)
    else:
        np.testing.assert_allclose(emu.ranks[0].global_.cpu().numpy(), single.global_.cpu().numpy(), rtol=1e-4, atol=2e-5 if precision == "fp32" else 5e-3)
    for h in he:
        np.testing.assert_allclose(h["global"], hs["global"], atol=1e-6 if clients == world else 2e-3)
        assert h["local"].shape == hs["local"].shape
        np.testing.assert_allclose(h["local"], hs["local"], atol=1e-6 if clients == world else 2e-3)
    # write-back happened on every rank
    for r in emu.ranks:
        assert torch.equal(r.client_slab(0, 0), r.global_slab(0))


def test_emulated_early_stop_is_consistent(income):
    shards = make_shards("contiguous", 2000, 4)
    probs = [FedProblem.custom(14, 2, shards, patience=2, atol=1.0)]
    emu = EmulatedWorld(probs, income.X_train[:2000], income.y_train[:2000], world=2, max_rounds=16, precision="fp32")
    hist = emu.run(8)
    for h in hist:
        assert list(h["rounds"]) == [3] and bool(h["stopped"][0])
