"""End-to-end stepping (FedEngine.step_e2e): inputs from pinned host memory every step, result read back every step.
On the single-launch round the kernel pulls the rows from host memory itself; that must be indistinguishable from the
copy-engine path and must leave the device-resident copy of the inputs up to date."""
import numpy as np
import pytest
import torch

from fedb200 import FedEngine, FedProblem, make_shards
from fedb200.ops import _native as nat

pytestmark = pytest.mark.gpu


def _engine(income, clients=3, n=3000):
    return FedEngine([FedProblem.custom(14, 2, make_shards("per_client_shuffle", n, clients, seed=2), patience=0)],
                     income.X_train[:n], income.y_train[:n], income.X_test, income.y_test, device="cuda", max_rounds=64,
                     precision="bf16")


@pytest.mark.parametrize("clients", [1, 3])
def test_kernel_pull_equals_copy_engine_path(income, clients, monkeypatch):
    a = _engine(income, clients)
    assert a.pulls_host_inputs()
    xh, yh = a.pinned_train_inputs()
    for _ in range(6):
        sa = a.step_e2e(xh, yh).copy()
    monkeypatch.setenv("FEDB200_E2E_MEMCPY", "1")
    b = _engine(income, clients)
    assert not b.pulls_host_inputs()
    for _ in range(6):
        sb = b.step_e2e(xh, yh).copy()
    assert sa["done"][0] == sb["done"][0] == 6
    assert np.array_equal(sa["last"], sb["last"]) and sa["last"][0][0] > 0.5
    assert torch.equal(a.params, b.params) and torch.equal(a.global_, b.global_)


def test_kernel_pull_reads_the_host_buffers_every_step(income):
    """Change the pinned inputs between steps: the next round must train on the new values and write them through."""
    e = _engine(income, 3)
    ref = _engine(income, 3)
    xh, yh = e.pinned_train_inputs()
    e.step_e2e(xh, yh)
    ref.step_e2e(*ref.pinned_train_inputs())
    assert torch.equal(e.params, ref.params)
    xh[:, :14] *= -1.0                                    # same buffers, new contents: no re-capture, no explicit upload
    yh.copy_(1 - yh)
    s1 = e.step_e2e(xh, yh).copy()
    s2 = ref.step_e2e(*ref.pinned_train_inputs()).copy()
    assert not torch.equal(e.params, ref.params)
    assert not np.array_equal(s1["last"], s2["last"])
    torch.cuda.synchronize()
    rows = torch.from_numpy(np.concatenate(e.problems[0].shards).astype(np.int64))
    assert torch.equal(e.X[: e.n_train].cpu()[rows], xh[rows]) and torch.equal(e.y[: e.n_train].cpu()[rows], yh[rows])
    # and the same result as uploading the changed inputs explicitly into a fresh engine at the same state
    chk = _engine(income, 3)
    chk.step_e2e(*chk.pinned_train_inputs())
    chk.X[: chk.n_train].copy_(xh); chk.y[: chk.n_train].copy_(yh)
    chk.step_graph(); torch.cuda.synchronize()
    assert torch.equal(chk.params, e.params)


def test_polled_result_matches_synchronised_and_survives_interleaving(income, monkeypatch):
    """The steady-state step polls `done` in the mapped result record instead of synchronising the stream; mixing it with
    other stepping calls, and running past an early stop, must give the same states as the synchronising path."""
    def drive(e):
        xh, yh = e.pinned_train_inputs()
        out = []
        for i in range(12):
            if i in (4, 9):
                e.step_graph()                     # a round outside the polled path: expectation must re-arm
            out.append(e.step_e2e(xh, yh).copy())
        torch.cuda.synchronize()
        return out
    a = drive(_engine(income, 1))
    monkeypatch.setenv("FEDB200_E2E_SYNC", "1")
    b = drive(_engine(income, 1))
    for sa, sb in zip(a, b):
        assert sa["done"][0] == sb["done"][0] and np.array_equal(sa["last"], sb["last"]) and sa["stop"][0] == sb["stop"][0]
    assert a[-1]["done"][0] == 14
    # early stop: patience 1 with a huge tolerance stops after two rounds; further steps return the frozen state
    monkeypatch.delenv("FEDB200_E2E_SYNC")
    e = FedEngine([FedProblem.custom(14, 2, make_shards("contiguous", 1000, 1), patience=1, atol=1.0)], income.X_train[:1000],
                  income.y_train[:1000], income.X_test, income.y_test, device="cuda", max_rounds=64, precision="bf16")
    xh, yh = e.pinned_train_inputs()
    states = [e.step_e2e(xh, yh).copy() for _ in range(8)]
    assert states[-1]["stop"][0] and states[-1]["done"][0] == states[-2]["done"][0] < 8
