"""The tensor-core sklearn-style fit (csrc/fedb_fitx.cu: split fp16 operands, a cluster of 4 CTAs per client splitting the
second hidden layer, epoch loop and stopping rule on the device) against the fp32 CUDA-core kernels."""
import numpy as np
import pytest
import torch

from fedb200 import FedEngine, FedProblem
from fedb200.data.partition import contiguous
from fedb200.fed.problem import TrainHyper

pytestmark = pytest.mark.gpu


def _pair(income_nomean, monkeypatch, hidden=(50, 400), clients=2, rows=4000, **kw):
    """(CUDA-core engine, tensor-core engine) on the same problem.  FEDB200_FIT is read ONCE, when an engine is built."""
    d = income_nomean
    shards = contiguous(rows, clients)
    mk = lambda: FedEngine([FedProblem.sklearn_like(d.n_features, shards, hidden=hidden, **kw)], d.X_train[:rows], d.y_train[:rows],
                           device="cuda", max_rounds=2)
    monkeypatch.setenv("FEDB200_FIT", "launches")
    simt = mk()
    monkeypatch.delenv("FEDB200_FIT")
    tcx = mk()
    assert simt._fit_tc_mask() is None                       # stays on the CUDA cores although the variable is gone again
    assert tcx._fit_tc_mask() is not None and bool(tcx._fit_tc_mask().all())
    assert torch.equal(simt.params, tcx.params)
    return simt, tcx


def _fit_epochs(e, epochs, shuffle):
    e.reset_fit_state()
    e.enqueue_round_begin()
    for _ in range(epochs):
        e.active.fill_(1)
        n = e.fit_local(max_iter=1, shuffle=shuffle)
    torch.cuda.synchronize()
    return n


SGD = TrainHyper(opt="sgd", lr=0.02, momentum=0.9, alpha=1e-4, lr_step_size=0, lr_gamma=1.0, batch_size=200, max_iter=300)


@pytest.mark.parametrize("hidden", [(50, 400), (50, 100), (33, 72)])
@pytest.mark.parametrize("shuffle", [False, True])
def test_sgd_epochs_match_the_cuda_core_path_to_fp32_rounding(income_nomean, monkeypatch, hidden, shuffle):
    """Momentum SGD with L2 propagates errors linearly, so three epochs (30 steps, per-epoch reshuffle, a 72-row tail tile in every
    mini-batch) must agree with the fp32 CUDA-core kernels to rounding: loss 1e-6, weights 1e-6 of the largest weight
    (measured on B200: 2e-7 and 3e-7).  A ReLU unit whose pre-activation is within rounding of zero may legitimately flip and
    move the ~50 weights of that unit by ~lr / batch (seen with plain SGD: 1e-4); hence the quantile and the loose maximum."""
    simt, tcx = _pair(income_nomean, monkeypatch, hidden, hyper=SGD)
    for e in (simt, tcx):
        n = _fit_epochs(e, 3, shuffle)
        assert n.tolist() == [3, 3]
    np.testing.assert_allclose(tcx.last_loss.cpu().numpy(), simt.last_loss.cpu().numpy(), rtol=5e-5)
    assert torch.equal(tcx.step, simt.step) and int(tcx.step[0]) == 30
    assert torch.equal(tcx.epoch, simt.epoch) and torch.equal(tcx.no_improve, simt.no_improve)
    d = (tcx.params - simt.params).abs() / simt.params.abs().max()
    assert float(d.quantile(0.97)) <= 1e-6 and float(d.max()) <= 2e-3
    dm = (tcx.m - simt.m).abs() / simt.m.abs().max()
    assert float(dm.quantile(0.97)) <= 2e-5
    pad = ~simt.problems[0].spec.param_mask().repeat(2)
    assert float(tcx.params.cpu()[pad].abs().max()) == 0.0


@pytest.mark.parametrize("hidden", [(50, 400), (33, 72)])
def test_adam_epoch_matches_the_cuda_core_path(income_nomean, monkeypatch, hidden):
    """One unshuffled epoch = 10 steps of sklearn's Adam with L2.  Adam's first steps are lr * g / |g|: a gradient at rounding level
    gets a full-size step of either sign, so single weights differ by up to a few lr while the loss and the bulk agree
    (measured on B200, (50, 400): loss 2e-5, mean weight difference 1e-5 of the largest weight; (33, 72): 1.5e-6 / 2.4e-7).
    The exact (IEEE) square root and division give the same figures as the approximate ones the kernel uses."""
    simt, tcx = _pair(income_nomean, monkeypatch, hidden)
    for e in (simt, tcx):
        n = _fit_epochs(e, 1, False)
        assert n.tolist() == [1, 1]
    np.testing.assert_allclose(tcx.last_loss.cpu().numpy(), simt.last_loss.cpu().numpy(), rtol=1e-4)
    assert torch.equal(tcx.step, simt.step) and int(tcx.step[0]) == 10
    d = (tcx.params - simt.params).abs() / simt.params.abs().max()
    assert float(d.mean()) <= 2e-4 and float(d.max()) <= 0.2
    pad = ~simt.problems[0].spec.param_mask().repeat(2)
    assert float(tcx.params.cpu()[pad].abs().max()) == 0.0


def test_loss_curve_with_reshuffling_follows_the_cuda_core_path(income_nomean, monkeypatch):
    """Five Adam epochs with the keyed per-epoch reshuffle (same bijection in both paths), one launch per epoch: the loss
    curve agrees epoch by epoch (measured: 1e-6 in epoch 1, 5e-4 in epoch 3)."""
    simt, tcx = _pair(income_nomean, monkeypatch)
    curves = []
    for e in (simt, tcx):
        e.reset_fit_state()
        e.enqueue_round_begin()
        c = []
        for ep in range(5):
            e.active.fill_(1)
            e.fit_local(max_iter=1, shuffle=True)
            torch.cuda.synchronize()
            c.append(e.last_loss.cpu().numpy().copy())
        curves.append(np.array(c))
    assert curves[0][-1, 0] < curves[0][0, 0]
    np.testing.assert_allclose(curves[1], curves[0], rtol=5e-3)
    np.testing.assert_allclose(curves[1][:1], curves[0][:1], rtol=5e-5)


def test_full_fit_stops_and_scores_like_the_cuda_core_path(income_nomean, monkeypatch):
    """The whole fit to convergence.  The stopping epoch is decided by the 5th digit of the loss curve (tol 1e-4 against
    ten epochs' best), which Adam's amplification does not preserve between ANY two fp32 implementations: the two paths stop
    within tens of epochs of each other (measured [162, 183, 200, 188] against [212, 180, 201, 129]) at the same accuracy."""
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
    assert abs(int(n1.sum()) - int(n0.sum())) <= 0.35 * int(n0.sum())


def test_mixed_engine_routes_each_client_to_its_kernel(income_nomean):
    """A grid corner with shapes the tensor-core kernel takes ((50, 100)) and shapes it does not ((100,) and (100, 50))."""
    from fedb200.fed.tune import run_grid
    out = run_grid(income_nomean, 4, hidden_grid=[(100,), (50, 100), (100, 50)], lr_grid=[0.004, 0.01], max_iter_grid=(40,), device="cuda")
    assert len(out["results"]) == 6
    for r in out["results"]:
        assert 0.75 < r["global"][0] < 1.0 and np.all(r["n_iter"] >= 1), r["hidden_layer_sizes"]


@pytest.mark.parametrize("n_classes,hidden", [(2, (50, 100)), (4, (40, 130)), (7, (33, 72))])
def test_softmax_heads_match_the_cuda_core_path(income_nomean, monkeypatch, n_classes, hidden):
    """The kernel's other head: softmax + cross-entropy with 2, 4 and 7 outputs (the partial-logit exchange, dW3 and db3 for every
    class, the head rows beyond the first two whose optimizer state stays in global memory), momentum SGD so that rounding is all
    that may differ.  Labels: the income label refined by quantiles of two features, so every class is populated."""
    d = income_nomean
    rows = 3000
    X = d.X_train[:rows]
    y = d.y_train[:rows].astype(np.int64)
    if n_classes > 2:
        y = (y + 2 * (X[:, 0] > np.median(X[:, 0])) + 4 * (X[:, 2] > np.median(X[:, 2]))) % n_classes
    from fedb200.models.spec import MLPSpec
    hyper = TrainHyper(opt="sgd", lr=0.02, momentum=0.9, alpha=1e-4, lr_step_size=0, lr_gamma=1.0, batch_size=200, max_iter=300)
    mk = lambda: FedEngine([FedProblem(MLPSpec.classifier(d.n_features, hidden, n_classes, "softmax_ce"), contiguous(rows, 2), hyper,
                                       weighting="uniform", metric_mode="pooled", patience=0, init="sklearn_glorot", init_seed=3,
                                       init_per_client=False)], X, y.astype(d.y_train.dtype), device="cuda", max_rounds=2)
    monkeypatch.setenv("FEDB200_FIT", "launches")
    simt = mk()
    monkeypatch.delenv("FEDB200_FIT")
    tcx = mk()
    assert simt._fit_tc_mask() is None and tcx._fit_tc_mask() is not None and bool(tcx._fit_tc_mask().all())
    for e in (simt, tcx):
        n = _fit_epochs(e, 2, True)
        assert n.tolist() == [2, 2]
    np.testing.assert_allclose(tcx.last_loss.cpu().numpy(), simt.last_loss.cpu().numpy(), rtol=5e-5)
    assert float(simt.last_loss.min()) > 0.05
    dd = (tcx.params - simt.params).abs() / simt.params.abs().max()
    assert float(dd.quantile(0.97)) <= 1e-6 and float(dd.max()) <= 2e-3
    pad = ~simt.problems[0].spec.param_mask().repeat(2)
    assert float(tcx.params.cpu()[pad].abs().max()) == 0.0


def test_tensor_core_clients_are_left_alone_by_the_cuda_core_epochs(income_nomean):
    """A packed engine with one federation the tensor-core kernel takes ((50, 100)) and one it does not ((100,)): the first
    federation's clients must come out of fit_local exactly as they do from an engine that holds only them — same epochs, same
    parameters and moments bit for bit.  (The complement mask for the CUDA-core clients was once derived from the flags the fit
    kernel clears when a client stops, which re-activated every converged tensor-core client for one more CUDA-core epoch.)"""
    d = income_nomean
    rows = 2000
    shards = contiguous(rows, 2)
    tc = lambda: FedProblem.sklearn_like(d.n_features, shards, hidden=(50, 100), lr=0.004)
    cc = lambda: FedProblem.sklearn_like(d.n_features, shards, hidden=(100,), lr=0.01)
    mk = lambda probs: FedEngine(probs, d.X_train[:rows], d.y_train[:rows], device="cuda", max_rounds=2)
    mixed, pure = mk([tc(), cc()]), mk([tc()])
    assert mixed._fit_tc_mask().tolist() == [True, True, False, False] and pure._fit_tc_mask().tolist() == [True, True]
    out = []
    for e in (mixed, pure):
        e.reset_fit_state()
        e.enqueue_round_begin()
        out.append(e.fit_local(max_iter=60))
        torch.cuda.synchronize()
    assert out[0][:2].tolist() == out[1].tolist() and out[0][2:].min() >= 1
    n = 2 * pure.problems[0].spec.slab_size
    o = int(mixed.tb.groups[0]["slab_off"])
    assert int(pure.tb.groups[0]["slab_off"]) == 0
    assert torch.equal(mixed.params[o:o + n], pure.params[:n]) and torch.equal(mixed.m[o:o + n], pure.m[:n])
    assert torch.equal(mixed.epoch[:2], pure.epoch) and torch.equal(mixed.step[:2], pure.step)
