"""Hardware validation of the tcgen05 operand conventions (csrc/tc_common.cuh): every case
lays operands out on the host with the documented panel formula and checks D = A @ B^T."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def _check(D, want, tol=2e-2):
    err = (D.cpu() - want).abs().max().item()
    scale = want.abs().max().item()
    assert err <= tol * max(scale, 1.0), f"max err {err} vs scale {scale}"


def test_kmajor_single_panel():
    from fedb200.ops import tc
    g = torch.Generator().manual_seed(0)
    A, B = torch.randn(128, 64, generator=g), torch.randn(64, 64, generator=g)
    D = tc.probe(tc.to_panels(A), tc.to_panels(B), 64, [0, 32, 64, 96], [0, 32, 64, 96], 0, 1024, 0, 1024)
    _check(D, _bf(A) @ _bf(B).T)


def test_kmajor_n208_and_partial_k():
    from fedb200.ops import tc
    g = torch.Generator().manual_seed(1)
    A, B = torch.randn(128, 64, generator=g), torch.randn(208, 64, generator=g)
    D = tc.probe(tc.to_panels(A), tc.to_panels(B), 208, [0, 32, 64, 96], [0, 32, 64, 96], 0, 1024, 0, 1024)
    _check(D, _bf(A) @ _bf(B).T)
    # only the first 16 of the 64 K columns (layer 1: K = 16 features)
    D = tc.probe(tc.to_panels(A), tc.to_panels(B), 208, [0], [0], 0, 1024, 0, 1024)
    _check(D, _bf(A[:, :16]) @ _bf(B[:, :16]).T)


def test_kmajor_multi_panel_k256_and_accumulate():
    from fedb200.ops import tc
    g = torch.Generator().manual_seed(2)
    A, B = torch.randn(128, 256, generator=g), torch.randn(64, 256, generator=g)
    a_steps = [p * 128 * 128 + k * 32 for p in range(4) for k in range(4)]
    b_steps = [p * 64 * 128 + k * 32 for p in range(4) for k in range(4)]
    D = tc.probe(tc.to_panels(A), tc.to_panels(B), 64, a_steps, b_steps, 0, 1024, 0, 1024)
    _check(D, _bf(A) @ _bf(B).T)
    D2 = tc.probe(tc.to_panels(A), tc.to_panels(B), 64, a_steps, b_steps, 0, 1024, 0, 1024, repeat=2)
    _check(D2, 2 * (_bf(A) @ _bf(B).T))


def test_b_mn_major_shares_the_kmajor_panel():
    """dgrad: B is W2 stored [h2 rows][h1 cols]; as an MN-major operand N = h1 (cols), K = h2 (rows)."""
    from fedb200.ops import tc
    g = torch.Generator().manual_seed(3)
    A = torch.randn(128, 128, generator=g)         # [rows, K=128]
    Bt = torch.randn(128, 64, generator=g)         # memory: [K=128 rows][N=64 cols]
    a_steps = [p * 128 * 128 + k * 32 for p in range(2) for k in range(4)]
    b_steps = [k * 2048 for k in range(8)]         # 16 K-rows per instruction = 2 swizzle atoms
    D = tc.probe(tc.to_panels(A), tc.to_panels(Bt), 64, a_steps, b_steps, 0, 1024, 0, 1024, b_mn=True)
    _check(D, _bf(A) @ _bf(Bt))


def test_a_and_b_mn_major_wgrad():
    """wgrad: dW[M = feats of dZ][N = feats of H] = dZ^T @ H, both stored [rows][feats]."""
    from fedb200.ops import tc
    g = torch.Generator().manual_seed(4)
    dZ = torch.randn(128, 128, generator=g)        # [K = 128 rows][M = 128 feats] -> 2 panels
    H = torch.randn(128, 64, generator=g)          # [K = 128 rows][N = 64 feats]
    steps = [k * 2048 for k in range(8)]
    D = tc.probe(tc.to_panels(dZ), tc.to_panels(H), 64, steps, steps, 128 * 128, 1024, 0, 1024, a_mn=True, b_mn=True)
    _check(D, _bf(dZ).T @ _bf(H))
    # N = 16: the layer-1 weight gradient against the 16 feature columns
    D = tc.probe(tc.to_panels(dZ), tc.to_panels(H), 16, steps, steps, 128 * 128, 1024, 0, 1024, a_mn=True, b_mn=True)
    _check(D, _bf(dZ).T @ _bf(H[:, :16]))
