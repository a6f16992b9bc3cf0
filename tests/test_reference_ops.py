"""The fp32 host restatement (ops/reference.py) — the oracle of the GPU tests — is itself
checked here against torch.nn + torch.optim (custom mode, C:12-25, C:43-46, C:63-73) and
sklearn's MLPClassifier (sklearn mode, SK:301-393 + SKO Adam)."""
import numpy as np
import pytest
import torch

from fedb200 import FedEngine, FedProblem, TrainHyper, make_shards


def _torch_model(spec, slab):
    layers = []
    for l in range(spec.n_layers):
        lin = torch.nn.Linear(spec.dims[l], spec.dims[l + 1])
        with torch.no_grad():
            lin.weight.copy_(spec.weight_view(slab, l))
            lin.bias.copy_(spec.bias_view(slab, l))
        layers.append(lin)
        if l + 1 < spec.n_layers:
            layers.append(torch.nn.ReLU())
    return torch.nn.Sequential(*layers)


def test_custom_step_matches_torch_adam_steplr(income):
    shards = make_shards("contiguous", 900, 3)
    prob = FedProblem.custom(14, 2, shards, hyper=TrainHyper(lr_step_size=2, lr_gamma=0.5), patience=0)
    eng = FedEngine([prob], income.X_train[:900], income.y_train[:900], device="cpu", max_rounds=8)
    spec = prob.spec
    models, opts, scheds = [], [], []
    for k in range(3):
        m = _torch_model(spec, eng.client_slab(0, k).clone())
        o = torch.optim.Adam(m.parameters(), lr=0.004)
        models.append(m); opts.append(o); scheds.append(torch.optim.lr_scheduler.StepLR(o, step_size=2, gamma=0.5))
    X = torch.tensor(income.X_train[:900], dtype=torch.float32)
    y = torch.tensor(income.y_train[:900], dtype=torch.long)
    for step in range(5):
        eng.enqueue_local_step(0)
        for k in range(3):
            rows = torch.from_numpy(shards[k]).long()
            opts[k].zero_grad()
            loss = torch.nn.functional.cross_entropy(models[k](X[rows]), y[rows])
            loss.backward(); opts[k].step(); scheds[k].step()
            assert float(eng.loss[k]) == pytest.approx(float(loss), rel=2e-5)
            for l in range(spec.n_layers):
                np.testing.assert_allclose(spec.weight_view(eng.client_slab(0, k), l).numpy(),
                                           models[k][2 * l].weight.detach().numpy(), rtol=2e-4, atol=2e-6)
                np.testing.assert_allclose(spec.bias_view(eng.client_slab(0, k), l).numpy(),
                                           models[k][2 * l].bias.detach().numpy(), rtol=2e-4, atol=2e-6)
    # padding never moves
    assert float(eng.params[~spec.param_mask().repeat(3)].abs().max()) == 0.0


def test_fedavg_matches_reference_arithmetic(income):
    shards = [np.arange(0, 100, dtype=np.int32), np.arange(100, 350, dtype=np.int32), np.arange(350, 400, dtype=np.int32)]
    prob = FedProblem.custom(14, 2, shards, patience=0)
    eng = FedEngine([prob], income.X_train[:400], income.y_train[:400], device="cpu", max_rounds=4)
    sds = [prob.spec.to_state_dict(eng.client_slab(0, k)) for k in range(3)]
    sizes = [100, 250, 50]
    total = sum(sizes)
    want = {key: sum(sds[i][key] * sizes[i] / total for i in range(3)) for key in sds[0]}  # C:110-115 verbatim
    eng.enqueue_aggregate(do_metrics=False)
    got = prob.spec.to_state_dict(eng.global_slab(0))
    for key in want:
        np.testing.assert_array_equal(got[key], want[key].astype(np.float32))
        for k in range(3):
            np.testing.assert_array_equal(prob.spec.to_state_dict(eng.client_slab(0, k))[key], got[key])


def test_eval_and_metrics_match_sklearn(income):
    from sklearn.metrics import accuracy_score, f1_score, precision_score
    shards = make_shards("contiguous", 600, 2)
    prob = FedProblem.custom(14, 2, shards, patience=0)
    eng = FedEngine([prob], income.X_train[:600], income.y_train[:600], device="cpu", max_rounds=4)
    eng.run(1, use_graph=False)
    h = eng.history()
    # recompute with torch + sklearn from the pre-aggregation local weights: rerun deterministically
    eng2 = FedEngine([prob], income.X_train[:600], income.y_train[:600], device="cpu", max_rounds=4)
    eng2.enqueue_local_step(0)
    accs, f1s = [], []
    for k in range(2):
        m = _torch_model(prob.spec, eng2.client_slab(0, k))
        rows = shards[k]
        pred = m(torch.tensor(income.X_train[rows], dtype=torch.float32)).argmax(1).numpy()
        accs.append(accuracy_score(income.y_train[rows], pred))
        f1s.append(f1_score(income.y_train[rows], pred, average="weighted", zero_division=0))
        assert h["local"][0, k, 0] == pytest.approx(accs[-1], abs=1e-6)
    assert h["global"][0, 0, 0] == pytest.approx(np.mean(accs), abs=1e-6)
    assert h["global"][0, 0, 3] == pytest.approx(np.mean(f1s), abs=1e-6)


def test_sklearn_mode_matches_mlpclassifier_one_epoch(income_nomean):
    """One partial_fit epoch, no shuffle: same batches -> weights agree to fp32 accuracy."""
    from sklearn.neural_network import MLPClassifier
    n = 1000
    X, y = income_nomean.X_train[:n], income_nomean.y_train[:n]
    prob = FedProblem.sklearn_like(14, [np.arange(n, dtype=np.int32)], hidden=(50, 400))
    eng = FedEngine([prob], X, y, device="cpu", max_rounds=2)
    clf = MLPClassifier(hidden_layer_sizes=(50, 400), learning_rate_init=0.004, random_state=42, shuffle=False)
    for ep in range(2):
        clf.partial_fit(X, y, classes=[0, 1])
        eng.enqueue_epoch(shuffle=False)
        assert float(eng.last_loss[0]) == pytest.approx(clf.loss_, rel=5e-5)
    coefs, inter = prob.spec.to_coefs_intercepts(eng.client_slab(0, 0))
    for a, b in zip(coefs + inter, clf.coefs_ + clf.intercepts_):
        np.testing.assert_allclose(a, b, rtol=2e-3, atol=2e-5)
    assert int(eng.step[0]) == 10 and int(eng.epoch[0]) == 2


def test_sgd_momentum_matches_sklearn_sgd(income_nomean):
    from sklearn.neural_network import MLPClassifier
    n = 400
    X, y = income_nomean.X_train[:n], income_nomean.y_train[:n]
    hyper = TrainHyper(opt="sgd", lr=0.01, momentum=0.9, nesterov=True, alpha=1e-4, lr_step_size=0, batch_size=200, max_iter=5)
    prob = FedProblem.sklearn_like(14, [np.arange(n, dtype=np.int32)], hidden=(50,), hyper=hyper)
    eng = FedEngine([prob], X, y, device="cpu", max_rounds=2)
    clf = MLPClassifier(hidden_layer_sizes=(50,), solver="sgd", learning_rate_init=0.01, momentum=0.9, nesterovs_momentum=True,
                        random_state=42, shuffle=False)
    for ep in range(3):
        clf.partial_fit(X, y, classes=[0, 1])
        eng.enqueue_epoch(shuffle=False)
    coefs, inter = prob.spec.to_coefs_intercepts(eng.client_slab(0, 0))
    for a, b in zip(coefs + inter, clf.coefs_ + clf.intercepts_):
        np.testing.assert_allclose(a, b, rtol=2e-3, atol=2e-5)


def test_tol_stopping_rule():
    from fedb200.ops import reference as ref
    from fedb200.ops import _native as nat
    g = np.zeros(1, dtype=nat.GROUP_DTYPE)
    g["n_rows"], g["batch_size"], g["alpha"], g["slab_size"] = 100, 0, 0.0, 32
    stride = ref.batch_stats_stride(32)
    bs = torch.zeros(stride, dtype=torch.float64)
    active, epoch, ni = torch.ones(1, dtype=torch.int32), torch.zeros(1, dtype=torch.int32), torch.zeros(1, dtype=torch.int32)
    best, last = torch.full((1,), float("inf"), dtype=torch.float64), torch.zeros(1, dtype=torch.float64)
    losses = [1.0, 0.5] + [0.49999] * 20
    stopped_at = None
    for i, l in enumerate(losses):
        bs[0], bs[1] = l * 100, 100
        ref.epoch_end(g, bs, 1, 32, 1e-4, 10, 1000, active, epoch, ni, best, last)
        if not int(active[0]):
            stopped_at = i + 1
            break
    # improvement < tol from epoch 3 on: count exceeds 10 at the 11th stalled epoch (SK:759-777)
    assert stopped_at == 2 + 11


def test_shuffle_is_a_permutation_and_epoch_dependent():
    from fedb200.ops import reference as ref
    p0 = ref._feistel_perm(2000, 1, 0)
    p1 = ref._feistel_perm(2000, 1, 1)
    assert sorted(p0.tolist()) == list(range(2000)) and sorted(p1.tolist()) == list(range(2000))
    assert (p0 != p1).mean() > 0.9


def test_sklearn_mode_with_mt19937_stream_follows_sklearn_epoch_for_epoch(income_nomean):
    """Parity mode (TrainHyper.shuffle_stream='mt19937'): same init draws, same mini-batches, same Adam / L2 / loss as
    ``MLPClassifier(random_state=42)`` — the per-epoch training loss of the first ten epochs equals sklearn's float64 loss
    curve to 1e-6 on the reference's rank-0 shard (S:17-22, S:77-83).  (Beyond a few dozen epochs float32 and float64
    trajectories drift apart — sklearn itself stops after 259 epochs on float64 inputs and after 156 on the same data cast
    to float32 — so whole-run numbers are compared with a tolerance, tests/test_modes.py.)"""
    import warnings
    from sklearn.neural_network import MLPClassifier
    from fedb200 import FedEngine, FedProblem
    from fedb200.data.partition import contiguous
    d = income_nomean
    sh = contiguous(len(d.X_train), 4)[0]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        sk = MLPClassifier(activation="relu", hidden_layer_sizes=(50, 400), learning_rate_init=0.004, max_iter=10, random_state=42)
        sk.fit(d.X_train[sh], d.y_train[sh])
    prob = FedProblem.sklearn_like(d.n_features, [sh]).with_hyper(shuffle_stream="mt19937")
    eng = FedEngine([prob], d.X_train, d.y_train, device="cpu", max_rounds=2)
    eng.reset_fit_state()
    eng.enqueue_round_begin()
    ours = []
    for _ in range(10):
        eng.enqueue_epoch(True)
        ours.append(float(eng.last_loss[0]))
    np.testing.assert_allclose(ours, sk.loss_curve_, rtol=0, atol=1e-6)
    # ... and the device-side default stream is a different permutation of the same shard
    prob_f = FedProblem.sklearn_like(d.n_features, [sh])
    eng_f = FedEngine([prob_f], d.X_train, d.y_train, device="cpu", max_rounds=2)
    eng_f.reset_fit_state()
    eng_f.enqueue_round_begin()
    eng_f.enqueue_epoch(True)
    assert sorted(eng_f.idx_epoch[: len(sh)].tolist()) == sorted(sh.tolist())
    assert abs(float(eng_f.last_loss[0]) - sk.loss_curve_[0]) > 1e-6
