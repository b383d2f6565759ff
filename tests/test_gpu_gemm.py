"""mst_model_gemm (the dense building block of the model path) against torch float64 on the same operands: both kernels
(fp32 SIMT tiles, tensor-core TF32 with the 3-term split), the four operand layouts the model uses (nn.Linear forward
X.W^T, input gradient dY.W, weight gradient dY^T.X with split-K, transposed A), ragged sizes, bias / LeakyReLU /
accumulate epilogues.  Tolerance: 2e-6 of the row scale sum|a||b| — fp32 accumulation order only; a single-pass TF32
product would miss it by two orders of magnitude."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def torch():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('GPU tests need a CUDA device (no CPU fallback exists)')
    return torch


CASES = [  # (M, N, K, layout, split_k, bias, act, accumulate)
    (1000, 450, 32, 'linear', 1, True, 1, 0),
    (4099, 17, 450, 'dx', 1, False, 0, 0),
    (450, 32, 20011, 'dw', 37, False, 0, 2),
    (376, 17, 9000, 'dw', 8, False, 0, 2),
    (129, 64, 57, 'linear', 1, True, 0, 1),
    (300, 33, 100, 'at', 1, False, 0, 0),
    (96, 512, 128, 'linear', 1, True, 1, 0),
    (2048, 256, 192, 'dx', 1, False, 0, 1),
    (5, 7, 3, 'linear', 1, True, 0, 0),
]
# short contractions over many rows: kernel 0 (auto) routes them to the row-streaming k_gemm_smallk
SMALLK_CASES = [
    (5003, 450, 10, 'linear', 1, True, 0, 0),
    (4100, 376, 17, 'linear', 1, True, 1, 0),
    (4099, 130, 24, 'linear', 1, False, 0, 1),
    (4160, 512, 4, 'dx', 1, False, 0, 0),
]


@pytest.mark.parametrize('case,kernel', [(c, k) for c in CASES for k in (1, 2, 3)] + [(c, 0) for c in CASES + SMALLK_CASES])
def test_gemm_matches_float64(torch, case, kernel):
    from mst_native import model as mm
    M, N, K, layout, split, use_bias, act, accumulate = case
    gen = torch.Generator(device='cuda').manual_seed(M * 31 + N * 7 + K)
    r = lambda *s: torch.randn(*s, device='cuda', generator=gen)
    pad = 3                                   # leading dimensions wider than the logical width (column blocks)
    if layout == 'linear':                    # C = X (M,K) . W (N,K)^T
        X, W = r(M, K + pad), r(N, K + pad)
        A, a_si, a_sk, B, b_sk, b_sj = X, K + pad, 1, W, 1, K + pad
        A64, B64 = X[:, :K].double(), W[:, :K].double().t()
    elif layout == 'dx':                      # C = dY (M,K) . W (K,N)
        dY, W = r(M, K + pad), r(K, N + pad)
        A, a_si, a_sk, B, b_sk, b_sj = dY, K + pad, 1, W, N + pad, 1
        A64, B64 = dY[:, :K].double(), W[:, :N].double()
    elif layout == 'dw':                      # C (M,N) = dY (K,M)^T . X (K,N)
        dY, X = r(K, M + pad), r(K, N + pad)
        A, a_si, a_sk, B, b_sk, b_sj = dY, 1, M + pad, X, N + pad, 1
        A64, B64 = dY[:, :M].double().t(), X[:, :N].double()
    else:                                     # 'at': A transposed, B k-contiguous
        At, W = r(K, M + pad), r(N, K + pad)
        A, a_si, a_sk, B, b_sk, b_sj = At, 1, M + pad, W, 1, K + pad
        A64, B64 = At[:, :M].double().t(), W[:, :K].double().t()
    bias = r(N + 1)[1:] if use_bias else None           # parameters sit at arbitrary float offsets of the flat buffer
    C = r(M, N + pad)
    C0 = C.clone()
    mm.gemm(A, a_si, a_sk, B, b_sk, b_sj, C, N + pad, M, N, K, bias=bias, act=act, accumulate=accumulate, split_k=split,
            kernel=kernel)
    want = A64 @ B64
    if use_bias:
        want = want + bias.double()
    if act:
        want = torch.where(want > 0, want, 0.01 * want)
    if accumulate:
        want = want + C0[:, :N].double()
    scale = (A64.abs() @ B64.abs()) + 1.0
    err = ((C[:, :N].double() - want).abs() / scale).max().item()
    assert err < 2e-6, (case, kernel, err)
    assert torch.equal(C[:, N:], C0[:, N:])                      # nothing written outside the logical width
