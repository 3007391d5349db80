"""tcgen05 / TMA member-GEMM pipeline in isolation: C = A W^T + sigma * A f16(E)^T against float64 torch."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(A, W, E, sigma):
    A64, W64 = A.double(), W.double()
    out = A64 @ W64.T
    if E is not None:
        out = out + float(np.float32(sigma)) * (A64 @ E.half().double().T)
    return out


@pytest.mark.parametrize('M,N,K', [(128, 128, 64), (128, 256, 512), (77, 200, 128), (256, 512, 2048), (5, 9488, 512)])
@pytest.mark.parametrize('prec', [1, 0])
def test_gemm_selftest(M, N, K, prec):
    import nes_b200
    g = torch.Generator().manual_seed(M * 1000 + N + K)
    A = torch.randn(M, K, generator=g) * 0.5
    W = torch.randn(N, K, generator=g) * 0.03
    E = torch.randn(N, K, generator=g)
    C = nes_b200.gemm_selftest(A, W, E, sigma=0.01, precision=prec).double()
    ref = _ref(A, W, E, 0.01)
    scale = ref.abs().max().item()
    err = (C - ref).abs().max().item() / scale
    # EXACT: split-fp16 operands carry ~22 mantissa bits, fp32 accumulation over K; FAST: one fp16 rounding (2^-11) per operand
    assert err < (2e-5 if prec == 1 else 2e-3), err


def test_gemm_selftest_theta_only_and_identity():
    import nes_b200
    K = 128
    A = torch.eye(128, K)
    W = torch.arange(128 * K, dtype=torch.float32).reshape(128, K) / 1024.0    # exactly representable after the split
    C = nes_b200.gemm_selftest(A, W, None, 0.0, 1)
    assert torch.equal(C, W[:, :128].T.contiguous())
