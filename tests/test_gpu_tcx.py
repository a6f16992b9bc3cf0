"""The split-operand tcgen05 round kernel (csrc/fedb_tcx.cu: every fp32 operand as two fp16 numbers, three MMAs per
GEMM) against a float64 restatement and the fp32 SIMT kernels: it has to be fp32-grade, not merely "same direction"."""
import numpy as np
import pytest
import torch

from fedb200 import FedEngine, FedProblem, TrainHyper, make_shards

pytestmark = pytest.mark.gpu


def _mk(income, shards, prec, hyper=None, hidden=(50, 200), n=None, patience=0, seed=0, max_rounds=64, **kw):
    n = n or len(income.X_train)
    return FedEngine([FedProblem.custom(14, 2, shards, hidden=hidden, hyper=hyper or TrainHyper(), patience=patience, **kw)],
                     income.X_train[:n], income.y_train[:n], income.X_test, income.y_test, device="cuda",
                     max_rounds=max_rounds, precision=prec, seed=seed)


# ---- hardware conventions the kernel adds to those of tests/test_gpu_tc.py ----------------------------------
def _split(x):
    hi = x.to(torch.float16)
    lo = (x - hi.float()).to(torch.float16)
    return hi.float(), lo.float()


def test_probe_fp16_three_products_reach_fp32_grade():
    """[A_hi | A_lo] side by side in one K-major panel (K = 16 each), same for B: the three products of the layer-1 GEMM."""
    from fedb200.ops import tc
    g = torch.Generator().manual_seed(0)
    A, B = torch.randn(128, 16, generator=g), torch.randn(64, 16, generator=g) * 4
    (ah, al), (bh, bl) = _split(A), _split(B)
    a_img = tc.to_panels(torch.cat([ah, al], 1), torch.float16)
    b_img = tc.to_panels(torch.cat([bh, bl], 1), torch.float16)
    D = tc.probe(a_img, b_img, 64, [32, 0, 0], [0, 32, 0], 0, 1024, 0, 1024, fp16=True)
    want = A.double() @ B.double().T
    err = (D.cpu().double() - want).abs().max().item()
    assert err <= 2e-6 * want.abs().max().item(), err
    # one product alone is only fp16-grade: the test would notice a silently dropped correction term
    D1 = tc.probe(a_img, b_img, 64, [0], [0], 0, 1024, 0, 1024, fp16=True)
    assert (D1.cpu().double() - want).abs().max().item() > 1e-4 * want.abs().max().item()


def test_probe_64_row_panels_upper_half_is_dont_care():
    """M = 128 MMA on a 64-row A panel: lanes 0..63 are exact whatever follows the panel in shared memory; MN-major operands
    with 8 KB panels (LBO = 8192) and K = 64 rows; N = 32 against a [hi | lo] panel."""
    from fedb200.ops import tc
    g = torch.Generator().manual_seed(1)
    A = torch.randn(64, 64, generator=g)
    junk = torch.full((64, 64), float("nan"))
    B = torch.randn(208, 64, generator=g)
    a_img = torch.cat([tc.to_panels(A, torch.float16), tc.to_panels(junk, torch.float16)])
    D = tc.probe(a_img, tc.to_panels(B, torch.float16), 208, [0, 32, 64, 96], [0, 32, 64, 96], 0, 1024, 0, 1024, fp16=True)
    want = A.half().float() @ B.half().float().T
    assert (D.cpu()[:64] - want).abs().max().item() <= 1e-3 * want.abs().max().item()
    # wgrad: dZ [64 rows][128 feats] as two 8 KB panels, X [64 rows][hi 16 | lo 16] -> D[feat][0..31]
    dZ = torch.randn(64, 128, generator=g)
    X = torch.randn(64, 32, generator=g)
    steps = [k * 2048 for k in range(4)]
    D = tc.probe(tc.to_panels(dZ, torch.float16), tc.to_panels(X, torch.float16), 32, steps, steps, 8192, 1024, 0, 1024,
                 a_mn=True, b_mn=True, fp16=True)
    want = dZ.half().float().T @ X.half().float()
    assert (D.cpu() - want).abs().max().item() <= 1e-3 * want.abs().max().item()


def test_probe_mn_major_operand_starting_inside_the_swizzle_atom():
    """dZ3 lives in columns 32..47 of the X panel: an MN-major B operand whose start address is 64 bytes into the 128-byte
    swizzle row (the hardware applies the XOR to the absolute address, as it does for K-major sub-tile steps)."""
    from fedb200.ops import tc
    g = torch.Generator().manual_seed(2)
    H = torch.randn(64, 128, generator=g)          # A: [K = 64 rows][M = 128 feats] -> 2 panels of 8 KB
    X = torch.randn(64, 64, generator=g)           # B panel; the operand is its columns 32..47
    D = tc.probe(tc.to_panels(H, torch.float16), tc.to_panels(X, torch.float16), 16, [k * 2048 for k in range(4)],
                 [64 + k * 2048 for k in range(4)], 8192, 1024, 0, 1024, a_mn=True, b_mn=True, fp16=True)
    want = H.half().float().T @ X[:, 32:48].half().float()
    assert (D.cpu() - want).abs().max().item() <= 1e-3 * want.abs().max().item()


# ---- gradients ------------------------------------------------------------------------------------------------
def _grads64(spec, slab, X, y):
    """Exact (float64) gradient slab and loss of the custom model."""
    d0, h1, h2, C = spec.dims
    s = slab.double()
    W = [spec.weight_view(s, l).clone().requires_grad_(True) for l in range(3)]
    b = [spec.bias_view(s, l).clone().requires_grad_(True) for l in range(3)]
    a = X[:, :d0].double()
    for l in range(3):
        a = a @ W[l].T + b[l]
        if l < 2:
            a = torch.relu(a)
    loss = torch.nn.functional.cross_entropy(a, y)
    loss.backward()
    out = torch.zeros(spec.slab_size, dtype=torch.float64)
    for l in range(3):
        spec.weight_view(out, l).copy_(W[l].grad)
        spec.bias_view(out, l).copy_(b[l].grad)
    return out, float(loss)


@pytest.mark.parametrize("rows,clients,hidden", [(64, 1, (50, 200)), (1000, 2, (50, 200)), (2666, 3, (50, 200)),
                                                   (300, 2, (50, 100)), (520, 1, (33, 72)), (5000, 1, (50, 200))])
def test_gradients_are_fp32_grade(income, rows, clients, hidden):
    """Plain SGD with lr = 1, one step: params_before - params_after is the gradient.  Bound per layer: 1e-5 of the
    layer's largest gradient entry, against float64 AND against the fp32 SIMT kernels."""
    shards = [np.arange(c * rows, (c + 1) * rows, dtype=np.int32) for c in range(clients)]
    hyper = TrainHyper(opt="sgd", lr=1.0, lr_step_size=0)
    f32, tc = _mk(income, shards, "fp32", hyper, hidden, n=rows * clients), _mk(income, shards, "tc32", hyper, hidden, n=rows * clients)
    assert tc.precision_name == "tc32" and tc.tc_variant == "x2" and torch.equal(f32.params, tc.params)
    before = f32.params.clone()
    from fedb200.ops import tc as tcops
    f32.enqueue_local_step(0)
    tcops.local_round_tc(tc.tb, tc.tc_cluster, tc.X, tc.y, tc.idx, tc.params, tc.m, tc.v, tc.step, tc.active, tc.fed_state,
                         tc.loss, tc.conf, tc.scratch, 0, 1, False, variant="x2")
    torch.cuda.synchronize()
    g32, gtc = (before - f32.params).cpu(), (before - tc.params).cpu()
    spec = f32.problems[0].spec
    P = spec.slab_size
    X = torch.tensor(income.X_train[: rows * clients], dtype=torch.float32)
    y = torch.tensor(income.y_train[: rows * clients], dtype=torch.long)
    worst = 0.0
    for c in range(clients):
        exact, exact_loss = _grads64(spec, before[c * P:(c + 1) * P].cpu(), X[shards[c]], y[shards[c]])
        a, b = g32[c * P:(c + 1) * P].double(), gtc[c * P:(c + 1) * P].double()
        for l in range(3):
            for view in (spec.weight_view, spec.bias_view):
                e, k, f = view(exact, l), view(b, l), view(a, l)
                scale = float(e.abs().max()) + 1e-30
                err64, err32 = float((e - k).abs().max()) / scale, float((f - k).abs().max()) / scale
                worst = max(worst, err64)
                assert err64 <= 1e-5, ("vs float64", c, l, view.__name__, err64, scale)
                assert err32 <= 1e-5, ("vs fp32 SIMT", c, l, view.__name__, err32, scale)
        assert float(tc.loss[c]) == pytest.approx(exact_loss, rel=2e-6)
    print(f"[tcx] rows={rows} hidden={hidden}: worst |g - g64| / scale = {worst:.2e}")
    assert torch.equal(tc.step, f32.step)
    pad = ~spec.param_mask().repeat(clients)
    assert float(tc.params.cpu()[pad].abs().max()) == 0.0


def test_eval_counts_match_simt(income):
    shards = make_shards("contiguous", 3000, 3)
    f32, tc = _mk(income, shards, "fp32", n=3000), _mk(income, shards, "tc32", n=3000)
    from fedb200.ops import kernels as K, tc as tcops
    K.round_begin(f32.tb, f32.fed_state, f32.conf)
    f32.enqueue_eval_local()
    tcops.local_round_tc(tc.tb, tc.tc_cluster, tc.X, tc.y, tc.idx, tc.params, tc.m, tc.v, tc.step, tc.active, tc.fed_state,
                         tc.loss, tc.conf, tc.scratch, 0, 0, True, variant="x2")
    torch.cuda.synchronize()
    a, b = f32.confusion()[:3], tc.confusion()[:3]
    assert a.sum() == b.sum() == 3000 and np.abs(a - b).sum() <= 2   # a sign flip needs a logit tie at the 1e-6 level


def test_adam_rounds_track_fp32(income):
    """60 Adam rounds: the per-round accuracy of the split-operand path follows the fp32 SIMT path (the bf16 kernel's
    test allows 0.015 / 0.02 here); the run is bit-reproducible."""
    shards = make_shards("per_client_shuffle", 8000, 3, seed=3)
    f32, tc = _mk(income, shards, "fp32"), _mk(income, shards, "tc32")
    h32, htc = f32.run(60), tc.run(60)
    a32, atc = h32["global"][0, :, 0], htc["global"][0, :, 0]
    assert atc[-1] > 0.80 and abs(a32[-1] - atc[-1]) < 2e-3
    assert np.abs(a32 - atc).max() < 3e-3 and np.abs(a32 - atc).mean() < 5e-4
    # (weights are compared loosely: Adam normalises every update to ~lr, so a rounding-level difference in a near-zero
    # gradient entry moves that weight by up to 2 lr per round in either path — trajectories agree, individual entries drift)
    rel = float((f32.global_slab(0) - tc.global_slab(0)).abs().max() / f32.global_slab(0).abs().max())
    print(f"[tcx] 60 rounds: max |acc32 - acc| = {np.abs(a32 - atc).max():.2e}, weights differ by {rel:.2e} of scale")
    assert rel < 0.15
    tc2 = _mk(income, shards, "tc32")
    tc2.run(60)
    assert torch.equal(tc.params, tc2.params)


def test_multiple_local_steps_and_small_clusters(income):
    for n, k in ((2000, 2), (384, 3), (100, 1)):
        shards = make_shards("contiguous", n, k)
        hyper = TrainHyper(local_epochs=3)
        f32, tc = _mk(income, shards, "fp32", hyper, n=n), _mk(income, shards, "tc32", hyper, n=n)
        h32, htc = f32.run(10), tc.run(10)
        assert int(tc.step[0]) == 30 == int(f32.step[0])
        assert abs(h32["global"][0, -1, 0] - htc["global"][0, -1, 0]) < 5e-3, (n, k)
        assert float((f32.params - tc.params).abs().max() / f32.params.abs().max()) < 0.3, (n, k)


def test_sigmoid_head_sklearn_adam_l2(income_nomean):
    """The other model family on the same kernel: one logistic output, log-loss, L2, sklearn's Adam (full batch)."""
    shards = make_shards("contiguous", 1500, 2)
    hyper = TrainHyper(opt="adam_sklearn", lr=0.004, lr_step_size=0, alpha=1e-4)
    mk = lambda prec: FedEngine([FedProblem.sklearn_like(14, shards, hidden=(50, 100), hyper=hyper, init="torch_default", init_per_client=True)],
                                income_nomean.X_train[:1500], income_nomean.y_train[:1500], device="cuda", max_rounds=32, precision=prec)
    f32, tc = mk("fp32"), mk("tc32")
    assert tc.tc_variant == "x2"
    h32, htc = f32.run(20), tc.run(20)
    assert np.abs(h32["global"][0, :, 0] - htc["global"][0, :, 0]).max() < 5e-3
    assert float((f32.params - tc.params).abs().max() / f32.params.abs().max()) < 0.1


def test_early_stop_lands_in_the_reference_band(income):
    """`custom --clients 3` on the default precision: the reference stops after 161-226 rounds (four unseeded
    runs) with held-out accuracy 0.81-0.82; runs differ by init and shard draw, so the band is widened a little."""
    from fedb200 import prepare
    stops, accs = [], []
    for seed in range(6):
        shards = make_shards("per_client_shuffle", 8000, 3, seed=seed)
        eng = FedEngine([FedProblem.custom(14, 2, shards, patience=10, init_seed=seed)], income.X_train, income.y_train,
                        income.X_test, income.y_test, device="cuda", max_rounds=300)
        assert eng.precision_name == "tc32"
        h = eng.run(300)
        stops.append(int(h["rounds"][0]))
        accs.append(eng.evaluate_global("test")[0]["accuracy"])
    print("[tcx] early stop rounds", stops, "held-out accuracy", np.round(accs, 4).tolist())
    assert sum(120 <= r <= 260 for r in stops) >= 5, stops
    assert all(0.80 <= a <= 0.83 for a in accs), accs
