"""Layer-wise tensor-core path (fed/wide.py, csrc/fedb_gemm.cu + fedb_wide.cu) against the fp32 SIMT
engine on a model small enough for both."""
import numpy as np
import pytest
import torch

from fedb200 import FedEngine, FedProblem, TrainHyper

pytestmark = pytest.mark.gpu


def _pair(income, hidden, rows, clients, hyper):
    shards = [np.arange(c * rows, (c + 1) * rows, dtype=np.int32) for c in range(clients)]
    n = rows * clients
    mk = lambda prec: FedEngine([FedProblem.custom(14, 2, shards, hidden=hidden, hyper=hyper, patience=0)], income.X_train[:n],
                                income.y_train[:n], income.X_test, income.y_test, device="cuda", max_rounds=64, precision=prec)
    return mk("fp32"), mk("bf16_wide")


@pytest.mark.parametrize("hidden,rows,clients", [((128, 128), 256, 2), ((128, 256, 128), 128, 3)])
def test_wide_gradients_match_fp32(income, hidden, rows, clients):
    hyper = TrainHyper(opt="sgd", lr=1.0, lr_step_size=0)
    f32, wide = _pair(income, hidden, rows, clients, hyper)
    assert wide.precision_name == "bf16_wide" and torch.equal(f32.params, wide.params)
    before = f32.params.clone()
    f32.enqueue_local_step(0)
    wide.wide.enqueue_step()
    torch.cuda.synchronize()
    g32, gw = (before - f32.params).cpu(), (before - wide.params).cpu()
    spec = f32.problems[0].spec
    P = spec.slab_size
    for c in range(clients):
        for l in range(spec.n_layers):
            for view in (spec.weight_view, spec.bias_view):
                a, b = view(g32[c * P:(c + 1) * P], l), view(gw[c * P:(c + 1) * P], l)
                cos = float((a * b).sum() / (a.norm() * b.norm() + 1e-20))
                assert cos > 0.995, (c, l, view.__name__, cos)
                assert float((a - b).abs().max()) <= 0.2 * float(a.abs().max()) + 1e-7, (c, l, view.__name__)
    np.testing.assert_allclose(wide.loss.cpu().numpy(), f32.loss.cpu().numpy(), rtol=1e-2)
    assert torch.equal(wide.step, f32.step)
    assert float(wide.params.cpu()[~spec.param_mask().repeat(clients)].abs().max()) == 0.0
    # the bf16 shadow tracks the fp32 master WEIGHTS exactly (biases are read in fp32; their shadow is unused)
    sh, pw = wide.wide.shadow.cpu().float(), wide.params.cpu().to(torch.bfloat16).float()
    for c in range(clients):
        for l in range(spec.n_layers):
            assert torch.equal(spec.weight_view(sh[c * P:(c + 1) * P], l), spec.weight_view(pw[c * P:(c + 1) * P], l)), (c, l)


def test_wide_rounds_track_fp32_and_eval(income):
    f32, wide = _pair(income, (128, 128), 1024, 4, TrainHyper())
    h32, hw = f32.run(30), wide.run(30)
    a32, aw = h32["global"][0, :, 0], hw["global"][0, :, 0]
    assert aw[-1] > 0.8 and abs(a32[-1] - aw[-1]) < 0.02 and np.abs(a32 - aw).mean() < 0.02
    assert abs(wide.evaluate_global("test")[0]["accuracy"] - f32.evaluate_global("test")[0]["accuracy"]) < 0.02


def test_wide_runs_a_model_the_other_paths_cannot(income):
    shards = [np.arange(c * 512, (c + 1) * 512, dtype=np.int32) for c in range(2)]
    eng = FedEngine([FedProblem.custom(14, 2, shards, hidden=(1024, 1024, 1024), patience=0)], income.X_train[:1024],
                    income.y_train[:1024], income.X_test, income.y_test, device="cuda", max_rounds=16)
    assert eng.precision_name == "bf16_wide"
    h = eng.run(12)
    assert h["global"][0, -1, 0] > 0.78 and np.isfinite(eng.loss.cpu().numpy()).all()
    assert eng.evaluate_global("test")[0]["accuracy"] > 0.75
