"""sklearn-mode round policies, the packed hyper-parameter grid, the CLI and the reference-format
logging — on the CPU path with small budgets (the GPU path runs the same Python)."""
import json
import os

import numpy as np
import torch
import pytest

from fedb200 import FedEngine, FedProblem, make_shards
from fedb200.cli import main as cli_main
from fedb200.fed import logging as flog
from fedb200.fed.sklearn_mode import run_sklearn_rounds
from fedb200.fed.tune import build_grid, run_grid
from fedb200.models.checkpoint import load_checkpoint


def _sk_engine(data, clients=2, n=600, max_iter=4, hidden=(16,)):
    shards = make_shards("contiguous", n, clients)
    prob = FedProblem.sklearn_like(14, shards, hidden=hidden, max_iter=max_iter)
    return FedEngine([prob], data.X_train[:n], data.y_train[:n], data.X_test, data.y_test, device="cpu", max_rounds=4)


def test_reference_faithful_rounds_are_identical(income_nomean):
    """R5 of SURVEY §0: fit() re-initialises, so every round reproduces round 1."""
    h = run_sklearn_rounds(_sk_engine(income_nomean), rounds=3, policy="reference_faithful")
    assert np.array_equal(h["global"][0, 0], h["global"][0, 1]) and np.array_equal(h["global"][0, 1], h["global"][0, 2])
    assert np.array_equal(h["n_iter"][0], h["n_iter"][2])


def test_warm_start_improves_and_literal_policy_stalls(income_nomean):
    hw = run_sklearn_rounds(_sk_engine(income_nomean, max_iter=40), rounds=3, policy="warm_start")
    assert hw["global"][0, -1, 0] >= hw["global"][0, 0, 0] - 0.02
    hl = run_sklearn_rounds(_sk_engine(income_nomean, max_iter=400, hidden=(8,)), rounds=2, policy="sklearn_literal")
    # sklearn's warm_start keeps the no-improvement counter: once it has tripped, later fits stop at once
    if hl["n_iter"][0].max() < 400:
        assert hl["n_iter"][1].max() <= 2


def test_sklearn_round_uses_pooled_unweighted(income_nomean):
    eng = _sk_engine(income_nomean, clients=3, n=700)
    h = run_sklearn_rounds(eng, rounds=1)
    cm = eng.confusion()[:3]
    from fedb200.fed.metrics import pooled
    assert h["global"][0, 0, 0] == pytest.approx(pooled([c[:2, :2] for c in cm])["accuracy"], abs=1e-6)
    assert list(eng.tb.groups["agg_num"]) == [1.0, 1.0, 1.0] and list(eng.tb.groups["agg_den"]) == [3.0, 3.0, 3.0]


def test_grid_is_packed_and_picks_first_best(income_nomean):
    class D:  # a 500-row view of the data
        X_train, y_train = income_nomean.X_train[:500], income_nomean.y_train[:500]
        X_test, y_test = income_nomean.X_test[:200], income_nomean.y_test[:200]
        n_features = 14
    assert len(build_grid(14, make_shards("contiguous", 500, 2))) == 90
    out = run_grid(D, 2, hidden_grid=[(8,), (8, 8)], lr_grid=[0.004, 0.05], max_iter_grid=(3,), device="cpu", init="sklearn_glorot")
    assert len(out["results"]) == 4
    accs = [r["global"][0] for r in out["results"]]
    assert out["best_metrics"]["accuracy"] == pytest.approx(max(accs))
    assert out["results"][int(np.argmax(accs))]["hidden_layer_sizes"] == out["best_params"]["hidden_layer_sizes"]
    coefs, inter = out["best"]["weights"]
    assert coefs[0].shape[0] == 14 and inter[-1].shape == (1,)


def test_log_formats_match_reference_strings():
    loc = np.array([[0.5, 0.25, 0.125, 0.0625]])
    lines = flog.custom_round_lines(3, loc, [0.83456, 0.8, 0.8, 0.8])
    assert lines[0] == "\nRound 3:\n"
    assert lines[1] == "  RANK 0 - Local Metrics (Round 3): [accuracy: 0.5000, precision: 0.2500, recall: 0.1250, f1: 0.0625]"
    assert lines[2] == "  Global Metrics (Round 3): [accuracy: 0.8346, precision: 0.8000, recall: 0.8000, f1: 0.8000]"
    assert flog.custom_stop_lines(10, 162)[0] == "Early stopping triggered: No significant change in metrics for 10 rounds."
    s = flog.sklearn_round_lines(1, loc, [0.957, 0.9584, 0.957, 0.957])
    assert "  Accuracy: 0.9570" in s and "-" * 50 in s


def test_cli_custom_checkpoint_and_resume(tmp_path, capsys):
    ck, jl = str(tmp_path / "g.npz"), str(tmp_path / "log.jsonl")
    assert cli_main(["custom", "--clients", "2", "--rounds", "3", "--device", "cpu", "--checkpoint", ck, "--jsonl", jl,
                     "--synthetic-rows", "600"]) == 0
    out = capsys.readouterr().out
    assert "Global Metrics (Round 3)" in out and "RANK 1 - Local Metrics (Round 1)" in out
    spec, coefs, inter, meta, opt = load_checkpoint(ck)
    assert spec.dims == (14, 50, 200, 2) and meta["round"] == 3
    assert set(opt) == {"m", "v", "step", "params", "hist_global", "hist_local"} and opt["m"].shape[0] == 2
    assert meta["extras"]["fed_state"]["done"] == 3 and meta["extras"]["fed_state"]["round"] == 3
    recs = [json.loads(l) for l in open(jl)]
    assert len(recs) == 4 and recs[-1]["summary"]
    assert cli_main(["custom", "--clients", "2", "--rounds", "1", "--device", "cpu", "--resume", ck, "--quiet",
                     "--synthetic-rows", "600"]) == 0


def test_resume_continues_exactly(tmp_path, income):
    """3 rounds + checkpoint + 3 more rounds in a fresh engine == 6 rounds in one go: weights, moments, step counts, the
    early-stop anchor / patience count and the metric history all carry over (resume dropped the federation state)."""
    from fedb200 import FedEngine, FedProblem, make_shards
    mk = lambda: FedEngine([FedProblem.custom(14, 2, make_shards("contiguous", 600, 3), hidden=(8, 12), patience=10)],
                           income.X_train[:600], income.y_train[:600], income.X_test, income.y_test, device="cpu", max_rounds=8)
    whole, first, second = mk(), mk(), mk()
    whole.run(6)
    first.run(3)
    ck = str(tmp_path / "r.npz")
    first.save_checkpoint(ck, resume=True)
    second.load_checkpoint(ck)
    h = second.run(3)
    assert int(h["rounds"][0]) == 6
    for name in ("params", "m", "v", "step", "global_"):
        assert torch.equal(getattr(second, name), getattr(whole, name)), name
    a, b = whole.fed_state_host()[0], second.fed_state_host()[0]
    for k in ("round", "done", "count", "has_anchor", "stop", "stop_pending"):
        assert a[k] == b[k], k
    assert np.array_equal(a["anchor"], b["anchor"]) and np.array_equal(whole.history()["global"], h["global"])
    assert np.array_equal(whole.history()["local"], h["local"])


def test_cli_sklearn_and_tune_smoke(capsys):
    assert cli_main(["sklearn-avg", "--clients", "2", "--rounds", "1", "--max-iter", "2", "--hidden", "8", "--device", "cpu",
                     "--synthetic-rows", "500"]) == 0
    assert "Final Global Weight Statistics:" in capsys.readouterr().out
    assert cli_main(["tune", "--clients", "2", "--max-iter", "2", "--small", "--device", "cpu", "--quiet", "--synthetic-rows", "400"]) == 0
    assert "Best" not in capsys.readouterr().err


def test_step_e2e_on_cpu_streams_new_inputs(income):
    """The streaming API on the CPU engine: each step trains on whatever the host buffers hold at that moment."""
    from fedb200 import FedEngine, FedProblem, make_shards
    mk = lambda: FedEngine([FedProblem.custom(14, 2, make_shards("contiguous", 600, 3), hidden=(8, 12), patience=0)],
                           income.X_train[:600], income.y_train[:600], income.X_test, income.y_test, device="cpu", max_rounds=8)
    a, b = mk(), mk()
    assert not a.pulls_host_inputs()
    xh, yh = a.pinned_train_inputs()
    s1 = a.step_e2e(xh, yh).copy()
    b.run(1)
    assert s1["done"][0] == 1 and np.allclose(s1["last"][0], b.history()["global"][0, 0], atol=1e-6)
    xh[:, :14] *= -1.0
    a.step_e2e(xh, yh)
    b.run(1)
    assert not np.allclose(a.params.numpy(), b.params.numpy())          # a trained on the changed inputs
    assert a.e2e_bytes()[0] == 600 * a.ldx * 4 + 600 * 4


def test_reference_faithful_round_against_the_reference_numbers(income_nomean):
    """`sklearn-avg --policy reference_faithful --clients 4 --shuffle mt19937` against what the unmodified script prints at
    np=4 (tests/golden.py; reproduced with baseline/run_reference.py): pooled accuracy 0.9570, six weight-statistic lines.
    The arithmetic here is float32 where sklearn's is float64, and the tol / n_iter_no_change stopping rule amplifies that
    (see test_reference_ops.py): measured pooled accuracy 0.9489, weight std within 4 %."""
    import golden
    from fedb200 import FedEngine, FedProblem
    from fedb200.data.partition import contiguous
    from fedb200.fed.sklearn_mode import run_sklearn_rounds
    from fedb200.models.checkpoint import weight_statistics
    d = income_nomean
    prob = FedProblem.sklearn_like(d.n_features, contiguous(len(d.X_train), 4)).with_hyper(shuffle_stream="mt19937")
    eng = FedEngine([prob], d.X_train, d.y_train, d.X_test, d.y_test, device="cpu", max_rounds=2)
    h = run_sklearn_rounds(eng, 1, "reference_faithful")
    acc = float(h["global"][0, 0, 0])
    assert abs(acc - golden.SKLEARN_POOLED["accuracy"]) <= 0.015, acc
    stats = weight_statistics(*prob.spec.to_coefs_intercepts(eng.global_slab(0)))
    for (shape, mean, std), (gshape, gmean, gstd) in zip(stats, golden.SKLEARN_WEIGHT_STATS):
        assert tuple(shape) == tuple(gshape)
        assert abs(std - gstd) <= 0.06 * gstd + 1e-12, (shape, std, gstd)
        assert abs(mean - gmean) <= 0.05 * max(gstd, abs(gmean)), (shape, mean, gmean)


def test_warm_start_rounds_can_be_continued_call_by_call(income_nomean):
    """run_sklearn_rounds(fresh=False) continues from the state the previous call left: 1 + 1 + 1 rounds, one call each (what
    bench.py's end-to-end loop does), are the same three rounds as one call of 3."""
    a = _sk_engine(income_nomean, max_iter=6)
    ha = run_sklearn_rounds(a, rounds=3, policy="warm_start")
    b = _sk_engine(income_nomean, max_iter=6)
    for r in range(3):
        hb = run_sklearn_rounds(b, rounds=1, policy="warm_start", fresh=(r == 0))
    assert torch.equal(a.params, b.params) and torch.equal(a.global_, b.global_)
    assert np.array_equal(ha["global"], hb["global"]) and np.array_equal(ha["n_iter"][-1], hb["n_iter"][-1])


def test_simt_kernels_ask_for_all_shared_memory_when_weights_must_stream(monkeypatch):
    """ops.kernels.simt_smem_bytes: models whose weights fit are sized exactly; models whose weights do not fit get the whole
    227 KB (the layer loops stream the matrices through what the activations leave) unless FEDB200_SIMT_STREAM=0."""
    from fedb200.ops.kernels import simt_smem_bytes
    cap = 227 * 1024 - 64
    small = simt_smem_bytes([14, 50, 400, 1], 3)
    assert small < cap                                      # 14-50-400-1: activations + 86 KB of weights fit
    assert simt_smem_bytes([14, 400, 200, 1], 3) == cap
    monkeypatch.setenv("FEDB200_SIMT_STREAM", "0")
    acts_only = simt_smem_bytes([14, 400, 200, 1], 3)
    assert acts_only < cap and acts_only == (16 + 400 + 200 + 4 + 400 + 200) * 33 * 4 + 256
    with pytest.raises(ValueError):
        simt_smem_bytes([14, 4096, 4096, 2], 3)


def test_fit_tile_list_is_pruned_and_ordered(income_nomean):
    """FedEngine._fit_tile_list: one mini-batch's tiles per client (the general list reserves at least 8), largest models first,
    clients the tensor-core fit kernel takes left out."""
    from fedb200.ops import _native as nat
    shards = make_shards("contiguous", 600, 2)
    probs = [FedProblem.sklearn_like(14, shards, hidden=h, max_iter=3) for h in ((8,), (40, 30), (16,))]
    eng = FedEngine(probs, income_nomean.X_train[:600], income_nomean.y_train[:600], device="cpu", max_rounds=2)
    assert len(eng.train_tiles_np) == 6 * 8
    # heavy first in the general list too
    assert [int(o) for o in eng.train_tiles_np["owner"][:16:8]] == [2, 3]
    t, n = eng._fit_tile_list(None)
    tiles = t.numpy().view(nat.TILE_DTYPE).reshape(-1)
    need = -(-200 // nat.TILE_ROWS)                      # mini-batches of 200 rows
    assert n == 6 * need and sorted(set(int(x) for x in tiles["tile"])) == list(range(need))
    assert [int(o) for o in tiles["owner"][::need]] == [2, 3, 4, 5, 0, 1]
    eng._fit_tiles = None
    t, n = eng._fit_tile_list(np.array([False, False, True, True, False, False]))
    tiles = t.numpy().view(nat.TILE_DTYPE).reshape(-1)
    assert n == 4 * need and set(int(o) for o in tiles["owner"]) == {0, 1, 4, 5}
    # and a fit through the pruned list is the fit through the general one
    a = _sk_engine(income_nomean, max_iter=5)
    na = a.fit_local()
    b = _sk_engine(income_nomean, max_iter=5)
    b._fit_tile_list = lambda mask: (b.train_tiles, len(b.train_tiles_np))
    nb = b.fit_local()
    assert np.array_equal(na, nb) and torch.equal(a.params, b.params)
