"""The tensor-core sklearn-style fit (csrc/fedb_fitx.cu: split fp16 operands, a cluster of 4 CTAs per client splitting the
second hidden layer, epoch loop and stopping rule on the device) against the fp32 CUDA-core kernels."""
import numpy as np
import pytest
import torch

from fedb200 import FedEngine, FedProblem
from fedb200.data.partition import contiguous

pytestmark = pytest.mark.gpu


def _pair(income_nomean, monkeypatch, hidden=(50, 400), clients=2, rows=4000, **kw):
    d = income_nomean
    shards = contiguous(rows, clients)
    mk = lambda: FedEngine([FedProblem.sklearn_like(d.n_features, shards, hidden=hidden, **kw)], d.X_train[:rows], d.y_train[:rows],
                           device="cuda", max_rounds=2)
    monkeypatch.setenv("FEDB200_FIT", "launches")
    simt = mk()
    assert simt._fit_tc_mask() is None
    monkeypatch.delenv("FEDB200_FIT")
    tcx = mk()
    assert tcx._fit_tc_mask() is not None and bool(tcx._fit_tc_mask().all())
    assert torch.equal(simt.params, tcx.params)
    return simt, tcx


@pytest.mark.parametrize("hidden", [(50, 400), (50, 100), (33, 72)])
def test_one_epoch_matches_the_cuda_core_path(income_nomean, monkeypatch, hidden):
    """One unshuffled epoch = 10 mini-batch steps of sklearn's Adam with L2: epoch loss, weights, moments, counters."""
    simt, tcx = _pair(income_nomean, monkeypatch, hidden)
    for e in (simt, tcx):
        e.reset_fit_state()
        e.enqueue_round_begin()
        n = e.fit_local(max_iter=1, shuffle=False)
        assert n.tolist() == [1, 1]
    torch.cuda.synchronize()
    np.testing.assert_allclose(tcx.last_loss.cpu().numpy(), simt.last_loss.cpu().numpy(), rtol=2e-6)
    assert torch.equal(tcx.step, simt.step) and int(tcx.step[0]) == 10
    scale = float(simt.params.abs().max())
    assert float((tcx.params - simt.params).abs().max()) <= 2e-4 * scale      # ten Adam steps amplify rounding-level differences
    assert float((tcx.m - simt.m).abs().max()) <= 1e-5 * float(simt.m.abs().max())
    pad = ~simt.problems[0].spec.param_mask().repeat(2)
    assert float(tcx.params.cpu()[pad].abs().max()) == 0.0


def test_loss_curve_with_reshuffling_follows_the_cuda_core_path(income_nomean, monkeypatch):
    """Five epochs with the keyed per-epoch reshuffle (same bijection in both paths): the loss curve agrees epoch by epoch."""
    simt, tcx = _pair(income_nomean, monkeypatch)
    curves = []
    for e in (simt, tcx):
        e.reset_fit_state()
        e.enqueue_round_begin()
        c = []
        for ep in range(5):
            e.active.fill_(1)
            e.fit_local(max_iter=1, shuffle=True) if e is simt else None
            if e is tcx:
                from fedb200.ops import kernels as K
                h = e.problems[0].hyper
                act = torch.ones_like(e.active)
                K.fit_tcx(e.tb, e.params, e.m, e.v, e.X, e.y, e.idx, e.idx_epoch, e.step, act, e.fed_state, e.epoch, e.no_improve,
                          e.best_loss, e.last_loss, e.loss, h.tol, h.n_iter_no_change, h.max_iter, 1, True, e.seed + 0x5EED)
            torch.cuda.synchronize()
            c.append(e.last_loss.cpu().numpy().copy())
        curves.append(np.array(c))
    assert curves[0][-1, 0] < curves[0][0, 0]
    np.testing.assert_allclose(curves[1], curves[0], rtol=5e-4)
    np.testing.assert_allclose(curves[1][:2], curves[0][:2], rtol=1e-5)


def test_full_fit_stops_and_scores_like_the_cuda_core_path(income_nomean, monkeypatch):
    simt, tcx = _pair(income_nomean, monkeypatch, clients=4, rows=8000)
    out = []
    for e in (simt, tcx):
        e.reset_fit_state()
        e.enqueue_round_begin()
        n = e.fit_local()
        e.enqueue_eval_local()
        e.enqueue_aggregate()
        out.append((n, e.history()["global"][0, 0, 0]))
    (n0, a0), (n1, a1) = out
    print(f"[fitx] epochs cuda-core {n0.tolist()} tensor-core {n1.tolist()}; pooled accuracy {a0:.4f} / {a1:.4f}")
    assert np.all(n1 >= 30) and np.all(n1 <= 300) and abs(a0 - a1) < 0.02 and a1 > 0.92
    assert abs(int(n1.sum()) - int(n0.sum())) <= 0.35 * int(n0.sum())     # stopping epochs are chaotic in the last digits of the loss


def test_mixed_engine_routes_each_client_to_its_kernel(income_nomean):
    """A grid corner with shapes the tensor-core kernel takes ((50, 100)) and shapes it does not ((100,) and (100, 50))."""
    from fedb200.fed.tune import run_grid
    out = run_grid(income_nomean, 4, hidden_grid=[(100,), (50, 100), (100, 50)], lr_grid=[0.004, 0.01], max_iter_grid=(40,), device="cuda")
    assert len(out["results"]) == 6
    for r in out["results"]:
        assert 0.75 < r["global"][0] < 1.0 and np.all(r["n_iter"] >= 1), r["hidden_layer_sizes"]
