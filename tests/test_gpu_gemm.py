"""Grouped tcgen05 GEMM (csrc/fedb_gemm.cu) vs torch, for the three operand-layout combinations a
Linear layer needs and every epilogue."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["auto", "tma", "cpasync"], autouse=True)
def _variant(request, monkeypatch):
    """Every case runs on the default dispatch (CTA-pair kernel where the shape allows), the single-CTA TMA kernel and the
    cp.async fallback."""
    monkeypatch.setenv("FEDB200_GEMM", request.param)


def _bf(x):
    return x.to(torch.bfloat16)


def _close(got, want, tol=1.5e-2):
    err = (got.float() - want.float()).abs().max().item()
    assert err <= tol * max(1.0, want.float().abs().max().item()), err


@pytest.mark.parametrize("G,B,IN,OUT", [(2, 128, 64, 128), (3, 256, 192, 320), (1, 128, 16, 256), (2, 384, 4096, 128)])
def test_forward_bias_relu(G, B, IN, OUT):
    from fedb200.ops import gemm as gm
    g = torch.Generator(device="cuda").manual_seed(0)
    X = _bf(torch.randn(G, B, IN, device="cuda", generator=g))
    W = _bf(torch.randn(G, OUT, IN, device="cuda", generator=g) / IN ** 0.5)
    b = torch.randn(G, OUT, device="cuda", generator=g)
    Y = torch.empty(G, B, OUT, device="cuda", dtype=torch.bfloat16)
    gm.gemm(G, B, OUT, IN, X, B * IN, IN, False, W, OUT * IN, IN, False, gm.FWD_BIAS_RELU, C=Y, c_stride=B * OUT, ldc=OUT,
            bias=b, bias_stride=OUT)
    torch.cuda.synchronize()
    want = torch.relu(torch.einsum("gbi,goi->gbo", X.float(), W.float()) + b[:, None, :])
    _close(Y, want)


@pytest.mark.parametrize("G,B,IN,OUT", [(2, 128, 128, 64), (2, 256, 320, 192), (1, 128, 4096, 256)])
def test_dgrad_mask(G, B, IN, OUT):
    """dX = (dZ @ W) * (act > 0): A = dZ [B][OUT] K-major, B = W [OUT][IN] as an MN-major operand."""
    from fedb200.ops import gemm as gm
    g = torch.Generator(device="cuda").manual_seed(1)
    dZ = _bf(torch.randn(G, B, OUT, device="cuda", generator=g))
    W = _bf(torch.randn(G, OUT, IN, device="cuda", generator=g) / OUT ** 0.5)
    act = _bf(torch.randn(G, B, IN, device="cuda", generator=g))
    dX = torch.empty(G, B, IN, device="cuda", dtype=torch.bfloat16)
    gm.gemm(G, B, IN, OUT, dZ, B * OUT, OUT, False, W, OUT * IN, IN, True, gm.DGRAD_MASK, C=dX, c_stride=B * IN, ldc=IN,
            act=act, act_stride=B * IN, ldact=IN)
    torch.cuda.synchronize()
    want = torch.einsum("gbo,goi->gbi", dZ.float(), W.float()) * (act.float() > 0)
    _close(dX, want)


@pytest.mark.parametrize("G,B,IN,OUT", [(2, 128, 64, 128), (2, 256, 192, 384), (1, 1024, 16, 128)])
def test_wgrad_store_and_fused_sgd(G, B, IN, OUT):
    """dW = dZ^T @ A with both operands MN-major; then the same GEMM with the optimizer in its epilogue."""
    from fedb200.ops import gemm as gm
    from fedb200.ops import _native as nat
    g = torch.Generator(device="cuda").manual_seed(2)
    dZ = _bf(torch.randn(G, B, OUT, device="cuda", generator=g))
    A = _bf(torch.randn(G, B, IN, device="cuda", generator=g))
    dW = torch.empty(G, OUT, IN, device="cuda", dtype=torch.float32)
    gm.gemm(G, OUT, IN, B, dZ, B * OUT, OUT, True, A, B * IN, IN, True, gm.STORE_F32, C=dW, c_stride=OUT * IN, ldc=IN)
    torch.cuda.synchronize()
    want = torch.einsum("gbo,gbi->goi", dZ.float(), A.float())
    _close(dW, want, 5e-3)
    # fused SGD(lr=0.5, momentum=0): W <- W - 0.5 * dW, bf16 shadow refreshed
    groups = np.zeros(G, dtype=nat.GROUP_DTYPE)
    groups["opt"], groups["lr"], groups["lr_gamma"] = nat.OPT_IDS["sgd"], 0.5, 1.0
    d_groups = torch.from_numpy(groups.view(np.uint8).copy()).cuda()
    step = torch.ones(G, dtype=torch.int32, device="cuda")
    W = torch.randn(G, OUT, IN, device="cuda", generator=g)
    W0 = W.clone()
    Mo, Ve = torch.zeros_like(W), torch.zeros_like(W)
    Wsh = torch.zeros(G, OUT, IN, device="cuda", dtype=torch.bfloat16)
    gm.gemm(G, OUT, IN, B, dZ, B * OUT, OUT, True, A, B * IN, IN, True, gm.WGRAD_OPT, W=W, Mom=Mo, Vel=Ve, Wsh=Wsh,
            w_stride=OUT * IN, ldw=IN, groups=d_groups, step=step)
    torch.cuda.synchronize()
    _close(W, W0 - 0.5 * want, 5e-3)
    assert torch.equal(Wsh, W.to(torch.bfloat16))


# ---------------------------------------------------------------------------------------------------
# CTA-pair kernel (csrc/fedb_gemm_tma2.cu): 256 x 256 tiles on tcgen05 cta_group::2
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode", ["tma2", "tma2p"])
@pytest.mark.parametrize("G,B,IN,OUT", [(1, 256, 64, 256), (2, 512, 192, 256), (2, 256, 4096, 512), (5, 1024, 1024, 2048)])
def test_pair_forward_dgrad_wgrad(G, B, IN, OUT, mode, monkeypatch):
    from fedb200.ops import gemm as gm
    monkeypatch.setenv("FEDB200_GEMM", mode)
    g = torch.Generator(device="cuda").manual_seed(5)
    X = _bf(torch.randn(G, B, IN, device="cuda", generator=g))
    W = _bf(torch.randn(G, OUT, IN, device="cuda", generator=g) / IN ** 0.5)
    b = torch.randn(G, OUT, device="cuda", generator=g)
    Y = torch.empty(G, B, OUT, device="cuda", dtype=torch.bfloat16)
    gm.gemm(G, B, OUT, IN, X, B * IN, IN, False, W, OUT * IN, IN, False, gm.FWD_BIAS_RELU, C=Y, c_stride=B * OUT, ldc=OUT,
            bias=b, bias_stride=OUT)
    torch.cuda.synchronize()
    assert gm.pair_status() == 0
    _close(Y, torch.relu(torch.einsum("gbi,goi->gbo", X.float(), W.float()) + b[:, None, :]))
    if IN % 256 == 0:
        # dgrad: dX[B][IN] = dZ[B][OUT] @ W[OUT][IN], W as an MN-major operand
        dZ = _bf(torch.randn(G, B, OUT, device="cuda", generator=g))
        dX = torch.empty(G, B, IN, device="cuda", dtype=torch.bfloat16)
        gm.gemm(G, B, IN, OUT, dZ, B * OUT, OUT, False, W, OUT * IN, IN, True, gm.DGRAD_MASK, C=dX, c_stride=B * IN, ldc=IN,
                act=X, act_stride=B * IN, ldact=IN)
        torch.cuda.synchronize()
        assert gm.pair_status() == 0
        _close(dX, torch.einsum("gbo,goi->gbi", dZ.float(), W.float()) * (X.float() > 0))
        # wgrad: dW[OUT][IN] = dZ^T @ X, both operands MN-major
        dW = torch.empty(G, OUT, IN, device="cuda", dtype=torch.float32)
        gm.gemm(G, OUT, IN, B, dZ, B * OUT, OUT, True, X, B * IN, IN, True, gm.STORE_F32, C=dW, c_stride=OUT * IN, ldc=IN)
        torch.cuda.synchronize()
        assert gm.pair_status() == 0
        _close(dW, torch.einsum("gbo,gbi->goi", dZ.float(), X.float()), 5e-3)


@pytest.mark.parametrize("mode", ["auto", "tma", "tma2", "tma2p"])
@pytest.mark.parametrize("opt", ["adam", "sgd_momentum", "sgd"])
def test_wgrad_fused_optimizer_matches_torch(mode, opt, monkeypatch):
    """dW = dZ^T @ A with the optimizer in the epilogue, several groups with different steps, against fp32 torch math
    (torch-style Adam: m, v, bias correction; SGD with sklearn-style momentum buffer)."""
    from fedb200.ops import gemm as gm
    from fedb200.ops import _native as nat
    monkeypatch.setenv("FEDB200_GEMM", mode)
    G, B, IN, OUT = 3, 512, 512, 256
    g = torch.Generator(device="cuda").manual_seed(7)
    dZ = _bf(torch.randn(G, B, OUT, device="cuda", generator=g) / B ** 0.5)
    A = _bf(torch.randn(G, B, IN, device="cuda", generator=g))
    groups = np.zeros(G, dtype=nat.GROUP_DTYPE)
    lr, b1, b2, eps, mom = 1e-2, 0.9, 0.999, 1e-8, 0.9
    if opt == "adam":
        groups["opt"], groups["beta1"], groups["beta2"], groups["eps"] = nat.OPT_IDS["adam_torch"], b1, b2, eps
    else:
        groups["opt"], groups["momentum"] = nat.OPT_IDS["sgd"], (mom if opt == "sgd_momentum" else 0.0)
    groups["lr"], groups["lr_gamma"] = lr, 1.0
    d_groups = torch.from_numpy(groups.view(np.uint8).copy()).cuda()
    steps = torch.tensor([1, 2, 5], dtype=torch.int32, device="cuda")
    W = torch.randn(G, OUT, IN, device="cuda", generator=g)
    Mo = torch.randn(G, OUT, IN, device="cuda", generator=g) * 0.01
    Ve = torch.rand(G, OUT, IN, device="cuda", generator=g) * 1e-3
    W0, M0, V0 = W.clone(), Mo.clone(), Ve.clone()
    Wsh = torch.zeros(G, OUT, IN, device="cuda", dtype=torch.bfloat16)
    gm.gemm(G, OUT, IN, B, dZ, B * OUT, OUT, True, A, B * IN, IN, True, gm.WGRAD_OPT, W=W, Mom=Mo, Vel=Ve, Wsh=Wsh,
            w_stride=OUT * IN, ldw=IN, groups=d_groups, step=steps)
    torch.cuda.synchronize()
    assert gm.pair_status() == 0
    dW = torch.einsum("gbo,gbi->goi", dZ.float(), A.float())
    t = steps.float()[:, None, None]
    if opt == "adam":
        m = b1 * M0 + (1 - b1) * dW
        v = b2 * V0 + (1 - b2) * dW * dW
        want = W0 - (lr / (1 - b1 ** t)) * m / (v.sqrt() / (1 - b2 ** t).sqrt() + eps)
        assert torch.allclose(Mo, m, rtol=1e-3, atol=1e-5) and torch.allclose(Ve, v, rtol=1e-3, atol=1e-7)
    elif opt == "sgd_momentum":
        m = mom * M0 - lr * dW
        want = W0 + m
        assert torch.allclose(Mo, m, rtol=1e-3, atol=1e-5)
    else:
        want = W0 - lr * dW
        assert torch.equal(Mo, M0) and torch.equal(Ve, V0)          # plain SGD leaves the state arrays alone
    assert torch.allclose(W, want, rtol=1e-3, atol=2e-4), (W - want).abs().max().item()
    assert torch.allclose(Wsh.float(), W, rtol=1e-2, atol=1e-2)
