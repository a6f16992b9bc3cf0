"""CUDA SIMT kernels vs the fp32 host restatement, entry point by entry point (same tables,
same buffers), then whole rounds."""
import numpy as np
import pytest
import torch

from fedb200 import FedEngine, FedProblem, TrainHyper, make_shards
from fedb200.ops import _native as nat

pytestmark = pytest.mark.gpu


def _pair(problems, data, n=None, **kw):
    X, y = (data.X_train, data.y_train) if n is None else (data.X_train[:n], data.y_train[:n])
    cpu = FedEngine(problems, X, y, data.X_test, data.y_test, device="cpu", **kw)
    gpu = FedEngine(problems, X, y, data.X_test, data.y_test, device="cuda", **kw)
    return cpu, gpu


def _problems(data, kind="custom"):
    if kind == "custom":
        return [FedProblem.custom(14, 2, make_shards("per_client_shuffle", 3000, 3, seed=2), patience=0),
                FedProblem.custom(14, 2, make_shards("contiguous", 3000, 4), hidden=(37,), patience=0,
                                  hyper=TrainHyper(opt="sgd", lr=0.05, momentum=0.9, lr_step_size=0))]
    return [FedProblem.sklearn_like(14, make_shards("contiguous", 3000, 4), hidden=h, lr=lr, init="sklearn_glorot")
            for h, lr in (((50, 400), 0.004), ((100,), 0.01), ((400, 200), 0.002))]


def test_native_library_is_loaded():
    assert nat.lib(required=True) is not None and nat.lib().fedb_version() >= 100


def test_init_bitwise(income):
    cpu, gpu = _pair(_problems(income), income, 3000, max_rounds=4)
    assert torch.equal(cpu.params, gpu.params.cpu())
    assert float(cpu.params.abs().sum()) > 0


@pytest.mark.parametrize("kind", ["custom", "sklearn"])
def test_fwd_bwd_gradients(income, income_nomean, kind):
    data = income if kind == "custom" else income_nomean
    cpu, gpu = _pair(_problems(data, kind), data, 3000, max_rounds=4)
    for e in (cpu, gpu):
        from fedb200.ops import kernels as K
        K.fwd_bwd(e.tb, e.train_tiles, len(e.train_tiles_np), e.params, e.X, e.y, e.idx, 0, e.scratch, e.step, e.active, e.fed_state)
    torch.cuda.synchronize()
    sc_c, sc_g = cpu.scratch.numpy(), gpu.scratch.cpu().numpy()
    for gi, g in enumerate(cpu.tb.groups):
        P, stride, nt = int(g["slab_size"]), int(g["slab_size"]) + nat.LOSS_SLOTS, int(g["max_tiles"])
        o = int(g["scratch_off"])
        rc = sc_c[o: o + nt * stride].reshape(nt, stride)
        rg = sc_g[o: o + nt * stride].reshape(nt, stride)
        gc, gg = rc[0, :P], rg[:, :P].sum(0)
        scale = np.abs(gc).max()
        np.testing.assert_allclose(gg, gc, rtol=2e-4, atol=2e-6 * max(scale, 1e-3) + 1e-8)
        assert rg[:, P].sum() == pytest.approx(rc[0, P], rel=2e-5)
    assert torch.equal(cpu.step, gpu.step.cpu())


@pytest.mark.parametrize("kind", ["custom", "sklearn"])
def test_reduce_opt_given_same_gradients(income, income_nomean, kind):
    data = income if kind == "custom" else income_nomean
    cpu, gpu = _pair(_problems(data, kind), data, 3000, max_rounds=4)
    from fedb200.ops import kernels as K
    for it in range(3):
        K.fwd_bwd(cpu.tb, cpu.train_tiles, len(cpu.train_tiles_np), cpu.params, cpu.X, cpu.y, cpu.idx, 0, cpu.scratch, cpu.step,
                  cpu.active, cpu.fed_state)
        gpu.scratch.copy_(cpu.scratch)
        gpu.step.copy_(cpu.step)
        for e in (cpu, gpu):
            K.reduce_opt(e.tb, e.params, e.m, e.v, e.scratch, 0, e.step, e.active, e.fed_state, e.loss, e.batch_stats, e.n_batches)
        torch.cuda.synchronize()
        for name in ("params", "m", "v", "loss"):
            np.testing.assert_allclose(getattr(gpu, name).cpu().numpy(), getattr(cpu, name).numpy(), rtol=3e-5, atol=1e-7, err_msg=name)
        gpu.params.copy_(cpu.params); gpu.m.copy_(cpu.m); gpu.v.copy_(cpu.v)
    np.testing.assert_allclose(gpu.batch_stats.cpu().numpy().reshape(cpu.n_groups, -1).sum(1),
                               cpu.batch_stats.numpy().reshape(cpu.n_groups, -1).sum(1), rtol=1e-5)


def test_eval_confusion_and_fedavg_exact(income):
    cpu, gpu = _pair(_problems(income), income, 3000, max_rounds=4)
    for e in (cpu, gpu):
        from fedb200.ops import kernels as K
        K.round_begin(e.tb, e.fed_state, e.conf)
        e.enqueue_eval_local()
        e.enqueue_aggregate()
    torch.cuda.synchronize()
    cc, cg = cpu.confusion(), gpu.confusion()
    assert cc.sum() == cg.sum() and np.abs(cc - cg).sum() <= 4
    assert torch.equal(cpu.global_, gpu.global_.cpu()) and torch.equal(cpu.params, gpu.params.cpu())
    np.testing.assert_allclose(gpu.history()["global"], cpu.history()["global"], atol=2e-3)
    sc, sg = cpu.fed_state_host(), gpu.fed_state_host()
    assert list(sc["round"]) == list(sg["round"]) == [1, 1]


def test_fedavg_exact_with_many_clients(income):
    """300 clients in one federation: the aggregation kernel stages client descriptors 128 at a time; the result must
    still be the reference's term-by-term sum in client order, bit for bit."""
    problems = [FedProblem.custom(14, 2, make_shards("contiguous", 3000, 300), hidden=(9,), patience=0),
                FedProblem.custom(14, 2, make_shards("contiguous", 3000, 128), hidden=(5, 6), patience=0, weighting="uniform")]
    cpu, gpu = _pair(problems, income, 3000, max_rounds=2)
    for e in (cpu, gpu):
        from fedb200.ops import kernels as K
        K.round_begin(e.tb, e.fed_state, e.conf)
        e.enqueue_eval_local()
        e.enqueue_aggregate()
    torch.cuda.synchronize()
    assert torch.equal(cpu.global_, gpu.global_.cpu()) and torch.equal(cpu.params, gpu.params.cpu())


def test_shuffle_bitwise(income_nomean):
    cpu, gpu = _pair(_problems(income_nomean, "sklearn"), income_nomean, 3000, max_rounds=2)
    from fedb200.ops import kernels as K
    for e in (cpu, gpu):
        e.epoch.fill_(3)
        K.shuffle(e.tb, e.idx, e.idx_epoch, e.epoch, e.active, 1234)
    assert torch.equal(cpu.idx_epoch, gpu.idx_epoch.cpu())
    n = int(cpu.tb.groups[0]["n_rows"])
    assert sorted(cpu.idx_epoch[:n].tolist()) == sorted(cpu.idx[:n].tolist()) and not torch.equal(cpu.idx_epoch[:n], cpu.idx[:n])


def test_rounds_track_reference_and_graph_equals_eager(income):
    probs = [FedProblem.custom(14, 2, make_shards("per_client_shuffle", 8000, 3, seed=4), patience=0)]
    cpu, gpu = _pair(probs, income, max_rounds=40)
    eager = FedEngine(probs, income.X_train, income.y_train, income.X_test, income.y_test, device="cuda", max_rounds=40)
    hc = cpu.run(24, use_graph=False)
    hg = gpu.run(24, use_graph=True, unroll=8)
    he = eager.run(24, use_graph=False)
    assert list(hg["rounds"]) == [24]
    assert np.array_equal(hg["global"], he["global"]) and torch.equal(gpu.params, eager.params)  # deterministic
    np.testing.assert_allclose(hg["global"][0, :, 0], hc["global"][0, :, 0], atol=0.02)
    assert hg["global"][0, -1, 0] > 0.78
    assert abs(gpu.evaluate_global("test")[0]["accuracy"] - cpu.evaluate_global("test")[0]["accuracy"]) < 0.02


def test_sklearn_epoch_matches_reference(income_nomean):
    cpu, gpu = _pair(_problems(income_nomean, "sklearn"), income_nomean, 3000, max_rounds=2)
    for e in (cpu, gpu):
        e.enqueue_epoch(shuffle=True)
        e.enqueue_epoch(shuffle=True)
    torch.cuda.synchronize()
    np.testing.assert_allclose(gpu.last_loss.cpu().numpy(), cpu.last_loss.numpy(), rtol=2e-3)
    d = (gpu.params.cpu() - cpu.params).abs()
    assert float((d > 2e-3).float().mean()) < 1e-3
    assert torch.equal(cpu.epoch, gpu.epoch.cpu()) and torch.equal(cpu.step, gpu.step.cpu())


@pytest.mark.parametrize("shuffle", [True, False])
def test_one_launch_fit_is_bit_identical_to_per_step_launches(income_nomean, shuffle, monkeypatch):
    """fedb_fit_cluster (device-side epoch / mini-batch loop + stopping rule) against the same fit driven from the host with one
    launch per step: identical epoch counts, losses and parameters."""
    def fit(mode, max_iter):
        monkeypatch.setenv("FEDB200_FIT", mode)
        e = FedEngine(_problems(income_nomean, "sklearn"), income_nomean.X_train[:3000], income_nomean.y_train[:3000],
                      income_nomean.X_test, income_nomean.y_test, device="cuda", max_rounds=2)
        assert e._fit_in_one_launch() == (mode == "cluster")
        n_iter = e.fit_local(max_iter=max_iter, shuffle=shuffle)
        torch.cuda.synchronize()
        return e, n_iter
    for max_iter in (3, 40):
        a, na = fit("cluster", max_iter)
        b, nb = fit("launches", max_iter)
        assert np.array_equal(na, nb) and na.max() == max_iter or na.max() <= max_iter
        assert torch.equal(a.step, b.step) and torch.equal(a.epoch, b.epoch) and torch.equal(a.no_improve, b.no_improve)
        assert torch.equal(a.last_loss, b.last_loss) and torch.equal(a.best_loss, b.best_loss)
        assert torch.equal(a.params, b.params) and torch.equal(a.m, b.m) and torch.equal(a.v, b.v)


def test_early_stop_freezes_state(income):
    probs = [FedProblem.custom(14, 2, make_shards("contiguous", 2000, 2), patience=2, atol=1.0)]  # everything "close"
    gpu = FedEngine(probs, income.X_train[:2000], income.y_train[:2000], device="cuda", max_rounds=20)
    h = gpu.run(12, unroll=4)
    # round 1 anchors, rounds 2 and 3 decrement 2 -> 0, the flag takes effect at the top of round 4 (C:132-136)
    assert list(h["rounds"]) == [3] and bool(h["stopped"][0])
    before = gpu.params.clone()
    gpu.run(4, unroll=4)
    assert torch.equal(before, gpu.params)


def test_streamed_weights_are_bit_identical_to_global_reads(income_nomean, monkeypatch):
    """Models whose weights do not fit next to the activations ((100, 400), (400, 200), (200, 400)): the layer loops that stream the
    matrices through the left-over shared memory (column-blocked forward, row-blocked dgrad, cp.async double buffering, wide
    weight-gradient mapping) against the plain loops that read them from global memory (FEDB200_SIMT_STREAM=0, read when the
    tables are built): same parameters, moments and losses bit for bit after three epochs, and the same evaluation counts."""
    def fit(stream):
        monkeypatch.setenv("FEDB200_SIMT_STREAM", stream)
        monkeypatch.setenv("FEDB200_FIT", "launches")
        probs = [FedProblem.sklearn_like(14, make_shards("contiguous", 3000, 2), hidden=h, lr=0.004, init="sklearn_glorot")
                 for h in ((100, 400), (400, 200), (200, 400), (50, 400))]
        e = FedEngine(probs, income_nomean.X_train[:3000], income_nomean.y_train[:3000], income_nomean.X_test, income_nomean.y_test,
                      device="cuda", max_rounds=2)
        e.reset_fit_state()
        e.enqueue_round_begin()
        n = e.fit_local(max_iter=3)
        e.enqueue_eval_local()
        e.enqueue_aggregate()
        torch.cuda.synchronize()
        return e, n
    a, na = fit("1")
    b, nb = fit("0")
    assert a.tb.smem_bytes > b.tb.smem_bytes            # the streaming launch asks for everything
    assert np.array_equal(na, nb) and na.max() == 3
    assert torch.equal(a.params, b.params) and torch.equal(a.m, b.m) and torch.equal(a.v, b.v)
    assert torch.equal(a.last_loss, b.last_loss) and torch.equal(a.conf, b.conf)
    assert torch.equal(a.global_, b.global_)


@pytest.mark.parametrize("hidden,act", [((300, 300, 60), "tanh"), ((520, 40), "relu"), ((96, 500), "logistic"), ((24, 510, 30), "relu")])
def test_streamed_weights_odd_shapes_bitwise(income, monkeypatch, hidden, act):
    """Shapes around the limits of the streamed loops (three hidden layers, other activations, 3 classes' worth of head rows,
    d_out = 520 > 512 and ld = 520 > 512 that fall back to the plain loops layer by layer, partial last blocks): two full
    rounds (train, eval, FedAvg) with Adam on the fp32 CUDA-core path, streamed against global reads, bit for bit."""
    from fedb200.models.spec import MLPSpec

    def run(stream):
        monkeypatch.setenv("FEDB200_SIMT_STREAM", stream)
        spec = MLPSpec.classifier(14, hidden, 2, "softmax_ce", act)
        prob = FedProblem(spec, make_shards("contiguous", 900, 3), TrainHyper(batch_size=200, local_epochs=1), patience=0)
        e = FedEngine([prob], income.X_train[:900], income.y_train[:900], income.X_test, income.y_test, device="cuda", max_rounds=4,
                      precision="fp32")
        h = e.run(2, unroll=1)
        torch.cuda.synchronize()
        return e, h
    a, ha = run("1")
    b, hb = run("0")
    assert a.tb.smem_bytes >= b.tb.smem_bytes
    assert torch.equal(a.params, b.params) and torch.equal(a.m, b.m) and torch.equal(a.v, b.v) and torch.equal(a.global_, b.global_)
    assert np.array_equal(ha["global"], hb["global"]) and np.isfinite(ha["global"]).all()
