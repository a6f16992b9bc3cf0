"""The fused tcgen05 local-round kernel (csrc/fedb_tc.cu) against the fp32 SIMT kernels."""
import numpy as np
import pytest
import torch

from fedb200 import FedEngine, FedProblem, TrainHyper, make_shards

pytestmark = pytest.mark.gpu


def _engines(income, shards, hyper=None, hidden=(50, 200), n=None, **kw):
    n = n or len(income.X_train)
    mk = lambda prec: FedEngine([FedProblem.custom(14, 2, shards, hidden=hidden, hyper=hyper or TrainHyper(), patience=0, **kw)],
                                income.X_train[:n], income.y_train[:n], income.X_test, income.y_test, device="cuda",
                                max_rounds=64, precision=prec)
    return mk("fp32"), mk("bf16")


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def _emulate_grads(spec, slab, X, y):
    """The kernel's arithmetic restated in torch: bf16 rounding exactly where the kernel rounds
    (operands of every MMA), fp32 everywhere else.  Returns a gradient slab."""
    d0, h1, h2, C = spec.dims
    n = len(X)
    W1, b1 = spec.weight_view(slab, 0), spec.bias_view(slab, 0)
    W2, b2 = spec.weight_view(slab, 1), spec.bias_view(slab, 1)
    W3, b3 = spec.weight_view(slab, 2), spec.bias_view(slab, 2)
    Xa = torch.cat([X[:, :d0], torch.ones(n, 1)], 1)
    D1 = _bf(Xa) @ _bf(torch.cat([W1, b1[:, None]], 1)).T
    mask1 = D1 > 0
    H1a = torch.cat([torch.relu(D1), torch.ones(n, 1)], 1)
    D2 = _bf(H1a) @ _bf(torch.cat([W2, b2[:, None]], 1)).T
    H2 = torch.relu(D2)
    z = H2 @ W3.T + b3
    dz3 = (torch.softmax(z, 1) - torch.nn.functional.one_hot(y, C).float()) / n
    loss = torch.nn.functional.cross_entropy(z, y)
    gW3, gb3 = _bf(dz3).T @ _bf(H2), dz3.sum(0)
    dZ2 = (dz3 @ W3) * (D2 > 0)
    g2 = _bf(dZ2).T @ _bf(H1a)                      # [h2, h1+1]: last column is db2
    dH1 = _bf(dZ2) @ _bf(W2)
    dZ1 = dH1 * mask1
    g1 = _bf(dZ1).T @ _bf(Xa)                       # [h1, d0+1]
    out = torch.zeros(spec.slab_size)
    spec.weight_view(out, 0).copy_(g1[:, :d0]); spec.bias_view(out, 0).copy_(g1[:, d0])
    spec.weight_view(out, 1).copy_(g2[:, :h1]); spec.bias_view(out, 1).copy_(g2[:, h1])
    spec.weight_view(out, 2).copy_(gW3); spec.bias_view(out, 2).copy_(gb3)
    return out, float(loss)


@pytest.mark.parametrize("rows,clients,hidden", [(128, 1, (50, 200)), (1000, 2, (50, 200)), (2666, 3, (50, 200)), (300, 2, (50, 100)), (520, 1, (33, 72))])
def test_gradients_match_emulation_and_fp32(income, rows, clients, hidden):
    """Plain SGD with lr = 1, one step: params_before - params_after is the gradient."""
    shards = [np.arange(c * rows, (c + 1) * rows, dtype=np.int32) for c in range(clients)]
    hyper = TrainHyper(opt="sgd", lr=1.0, lr_step_size=0)
    f32, tc = _engines(income, shards, hyper, hidden, n=rows * clients)
    assert tc.precision_name == "bf16" and torch.equal(f32.params, tc.params)
    before = f32.params.clone()
    from fedb200.ops import tc as tcops
    f32.enqueue_local_step(0)
    tcops.local_round_tc(tc.tb, tc.tc_cluster, tc.X, tc.y, tc.idx, tc.params, tc.m, tc.v, tc.step, tc.active, tc.fed_state,
                         tc.loss, tc.conf, tc.scratch, 0, 1, False)
    torch.cuda.synchronize()
    g32, gtc = (before - f32.params).cpu(), (before - tc.params).cpu()
    spec = f32.problems[0].spec
    P = spec.slab_size
    X = torch.tensor(income.X_train[: rows * clients], dtype=torch.float32)
    y = torch.tensor(income.y_train[: rows * clients], dtype=torch.long)
    for c in range(clients):
        emu, emu_loss = _emulate_grads(spec, before[c * P:(c + 1) * P].cpu(), X[shards[c]], y[shards[c]])
        a, b = g32[c * P:(c + 1) * P], gtc[c * P:(c + 1) * P]
        for l in range(3):
            for view in (spec.weight_view, spec.bias_view):
                e, k, f = view(emu, l), view(b, l), view(a, l)
                scale = float(e.abs().max()) + 1e-12
                assert float((e - k).abs().max()) <= 4e-3 * scale, ("vs emulation", c, l, view.__name__, float((e - k).abs().max()), scale)
                cos = float((k * f).sum() / (k.norm() * f.norm() + 1e-20))
                assert cos > 0.995, ("vs fp32", c, l, view.__name__, cos)
        assert float(tc.loss[c]) == pytest.approx(emu_loss, rel=2e-3)
    np.testing.assert_allclose(tc.loss.cpu().numpy(), f32.loss.cpu().numpy(), rtol=1e-2)
    assert torch.equal(tc.step, f32.step)
    pad = ~spec.param_mask().repeat(clients)
    assert float(tc.params.cpu()[pad].abs().max()) == 0.0


def test_eval_counts_match_simt(income):
    shards = make_shards("contiguous", 3000, 3)
    f32, tc = _engines(income, shards, n=3000)
    from fedb200.ops import kernels as K, tc as tcops
    K.round_begin(f32.tb, f32.fed_state, f32.conf)
    f32.enqueue_eval_local()
    tcops.local_round_tc(tc.tb, tc.tc_cluster, tc.X, tc.y, tc.idx, tc.params, tc.m, tc.v, tc.step, tc.active, tc.fed_state,
                         tc.loss, tc.conf, tc.scratch, 0, 0, True)
    torch.cuda.synchronize()
    a, b = f32.confusion()[:3], tc.confusion()[:3]
    assert a.sum() == b.sum() == 3000 and np.abs(a - b).sum() <= 30  # bf16 logits flip a few near-ties


def test_adam_rounds_track_fp32(income):
    shards = make_shards("per_client_shuffle", 8000, 3, seed=3)
    f32, tc = _engines(income, shards)
    h32, htc = f32.run(40), tc.run(40)
    a32, atc = h32["global"][0, :, 0], htc["global"][0, :, 0]
    assert atc[-1] > 0.80 and abs(a32[-1] - atc[-1]) < 0.015
    assert np.abs(a32 - atc).mean() < 0.02
    assert abs(tc.evaluate_global("test")[0]["accuracy"] - f32.evaluate_global("test")[0]["accuracy"]) < 0.02
    # bit-reproducible: same run twice
    _, tc2 = _engines(income, shards)
    tc2.run(40)
    assert torch.equal(tc.params, tc2.params)


def test_multiple_local_steps(income):
    shards = make_shards("contiguous", 2000, 2)
    hyper = TrainHyper(local_epochs=3)
    f32, tc = _engines(income, shards, hyper, n=2000)
    h32, htc = f32.run(10), tc.run(10)
    assert int(tc.step[0]) == 30 == int(f32.step[0])
    assert abs(h32["global"][0, -1, 0] - htc["global"][0, -1, 0]) < 0.02
