"""Model spec / arena layout, init, checkpoint, metrics, early-stop."""
import numpy as np
import pytest
import torch

from fedb200.fed.metrics import EarlyStop, confusion_matrix, mean_of_local, metrics_from_confusion, pooled
from fedb200.models import init as minit
from fedb200.models.checkpoint import load_checkpoint, save_checkpoint, weight_statistics
from fedb200.models.spec import MLPSpec


def test_layout_custom_model():
    s = MLPSpec.classifier(14, (50, 200), 2)
    assert s.n_params == 11352  # what the reference pickles per rank
    assert s.w_offsets == (0, 852, 11452) and s.slab_size % 32 == 0
    assert all(o % 4 == 0 for o in s.w_offsets + s.b_offsets)
    assert int(s.param_mask().sum()) == s.n_params


def test_state_dict_and_sklearn_round_trip():
    s = MLPSpec.classifier(14, (50, 400), 2, head="sigmoid_logloss")
    slab = torch.randn(s.slab_size) * s.param_mask()
    sd = s.to_state_dict(slab)
    assert list(sd) == ["model.0.weight", "model.0.bias", "model.2.weight", "model.2.bias", "model.4.weight", "model.4.bias"]
    assert sd["model.2.weight"].shape == (400, 50)
    coefs, inter = s.to_coefs_intercepts(slab)
    assert [c.shape for c in coefs] == [(14, 50), (50, 400), (400, 1)] and [b.shape for b in inter] == [(50,), (400,), (1,)]
    back = torch.zeros_like(slab)
    s.from_coefs_intercepts(coefs, inter, back)
    assert torch.equal(back, slab)
    back2 = torch.zeros_like(slab)
    s.from_state_dict(sd, back2)
    assert torch.equal(back2, slab)


def test_checkpoint_round_trip(tmp_path):
    s = MLPSpec.classifier(14, (50, 200), 2)
    slab = torch.randn(s.slab_size) * s.param_mask()
    path = str(tmp_path / "ck.npz")
    save_checkpoint(path, s, slab, round_idx=7, extras={"note": "x"})
    spec, coefs, inter, meta, opt = load_checkpoint(path)
    assert spec == s and meta["round"] == 7 and not opt
    assert coefs[0].dtype == np.float64 and coefs[1].shape == (50, 200)
    slab2 = torch.zeros_like(slab)
    spec.from_coefs_intercepts(coefs, inter, slab2)
    assert torch.equal(slab, slab2)
    assert len(weight_statistics(coefs, inter)) == 6


def test_checkpoint_loads_into_sklearn_shapes():
    from sklearn.neural_network import MLPClassifier
    s = MLPSpec.classifier(14, (50, 400), 2, head="sigmoid_logloss")
    rng = np.random.default_rng(0)
    X, y = rng.normal(size=(64, 14)), rng.integers(0, 2, 64)
    clf = MLPClassifier(hidden_layer_sizes=(50, 400), max_iter=1).fit(X, y)
    slab = torch.zeros(s.slab_size)
    s.from_coefs_intercepts(clf.coefs_, clf.intercepts_, slab)
    coefs, inter = s.to_coefs_intercepts(slab)
    for a, b in zip(coefs + inter, clf.coefs_ + clf.intercepts_):
        np.testing.assert_allclose(a, b.astype(np.float32), rtol=0, atol=0)


def test_philox_known_answer():
    # Random123 known-answer test for philox4x32-10
    out = minit.philox4x32(0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344, 0xA4093822, 0x299F31D0)
    assert [int(x) for x in out] == [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]


def test_init_bounds():
    s = MLPSpec.classifier(14, (50, 200), 2)
    a = minit.init_slab(s, "torch_default", seed=3, stream=0)
    b = minit.init_slab(s, "torch_default", seed=3, stream=1)
    assert not np.array_equal(a, b)
    w1 = a[: 50 * 16].reshape(50, 16)
    assert np.abs(w1[:, :14]).max() <= 1 / np.sqrt(14) and np.all(w1[:, 14:] == 0) and np.abs(w1[:, :14]).max() > 0.9 / np.sqrt(14)
    g = minit.init_slab(s, "sklearn_glorot", seed=3)
    assert np.abs(g[: 50 * 16]).max() <= np.sqrt(6 / 64)


def test_sklearn_exact_init_matches_sklearn():
    from sklearn.neural_network import MLPClassifier
    s = MLPSpec.classifier(14, (50, 400), 2, head="sigmoid_logloss")
    rng = np.random.default_rng(0)
    X, y = rng.normal(size=(32, 14)), np.arange(32) % 2
    clf = MLPClassifier(hidden_layer_sizes=(50, 400), random_state=42, max_iter=1, learning_rate_init=1e-30, solver="sgd", momentum=0, alpha=0)
    clf.fit(X, y)  # lr ~ 0: coefficients stay at their initial values
    slab = torch.from_numpy(minit.init_slab_sklearn_exact(s, 42))
    coefs, inter = s.to_coefs_intercepts(slab)
    for a, b in zip(coefs + inter, clf.coefs_ + clf.intercepts_):
        np.testing.assert_allclose(a, b, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("C", [2, 3])
def test_metrics_match_sklearn(C):
    from sklearn.metrics import accuracy_score, f1_score, precision_score, recall_score
    rng = np.random.default_rng(C)
    for trial in range(5):
        yt, yp = rng.integers(0, C, 300), rng.integers(0, C, 300)
        if trial == 0:
            yp[:] = 0  # a class never predicted -> zero_division path
        m = metrics_from_confusion(confusion_matrix(yt, yp, C))
        assert m["accuracy"] == pytest.approx(accuracy_score(yt, yp))
        assert m["precision"] == pytest.approx(precision_score(yt, yp, average="weighted", zero_division=0))
        assert m["recall"] == pytest.approx(recall_score(yt, yp, average="weighted", zero_division=0))
        assert m["f1"] == pytest.approx(f1_score(yt, yp, average="weighted", zero_division=0))


def test_pooled_vs_mean_of_local():
    a, b = np.array([[10, 0], [5, 5]]), np.array([[1, 3], [0, 4]])
    assert pooled([a, b])["accuracy"] == pytest.approx(20 / 28)
    assert mean_of_local([a, b])["accuracy"] == pytest.approx((15 / 20 + 5 / 8) / 2)


def test_early_stop_anchor_semantics():
    es = EarlyStop(patience=3, atol=1e-4)
    m = lambda a: {"accuracy": a, "precision": a, "recall": a, "f1": a}
    assert not es.update(m(0.5))              # anchors
    assert not es.update(m(0.50005))          # within tol: count 3 -> 2, anchor stays 0.5
    assert not es.update(m(0.50009))          # still compared to 0.5, count -> 1
    assert es.anchor[0] == 0.5
    assert not es.update(m(0.6))              # re-anchor, count reset
    assert es.count == 3
    es.update(m(0.6)); es.update(m(0.6))
    assert es.update(m(0.6)) and es.stop


def test_simt_shared_memory_plan_matches_library():
    """ops.kernels.simt_smem_bytes mirrors smem_plan in csrc/fedb_simt.cu; the engine sizes launches with the former."""
    import ctypes
    from fedb200.ops import _native as nat, kernels as K
    lib = nat.lib(required=False)
    if lib is None:
        pytest.skip("native library not built")
    for dims in [(14, 50, 400, 1), (14, 50, 200, 2), (14, 400, 200, 1), (14, 100, 1), (14, 37, 2), (14, 5, 6, 2), (14, 200, 400, 1),
                 (14, 3000, 3000, 2), (14, 50, 60, 70, 2)]:
        want = lib.fedb_simt_smem_bytes((ctypes.c_int32 * 8)(*dims), len(dims) - 1)
        try:
            got = K.simt_smem_bytes(dims, len(dims) - 1)
        except ValueError:
            got = -1
        assert got == want, (dims, got, want)
