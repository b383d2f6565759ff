"""Times mst_model_gemm on the tall-skinny shapes of the training step (development probe)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))
import torch
from mst_native import model as mm

R = 512000
SHAPES = [  # name, layout, M, N, K, lda_pad, split
    ('psa zr fwd', 'linear', R, 450, 10, 88, 1),
    ('usa zr fwd', 'linear', R, 376, 17, 82, 1),
    ('psa dX', 'dx', R, 10, 450, 88, 1),
    ('usa dX', 'dx', R, 17, 376, 82, 1),
    ('psa dW', 'dw', 450, 10, R, 0, 74),
    ('usa dW', 'dw', 376, 17, R, 0, 99),
    ('wide dW', 'dw', 64, 456, 204800, 0, 37),
    ('sq dW', 'dw', 64, 64, 204800, 0, 296),
    ('lin 64x456', 'linear', 204800, 64, 456, 456, 1),
    ('lin 128x64', 'linear', 204800, 128, 64, 64, 1),
    ('lin 456x64', 'linear', 204800, 456, 64, 64, 1),
    ('lin 256x64', 'linear', 204800, 256, 64, 64, 1),
]


def run(name, layout, M, N, K, ld, split, kernel, reps=5):
    dev = 'cuda'
    if layout == 'linear':
        X = torch.randn(M, K, device=dev); W = torch.randn(N, max(ld, K), device=dev)
        args = (X, K, 1, W, 1, W.shape[1])
        byts = 4 * (M * K + M * N)
    elif layout == 'dx':
        dY = torch.randn(M, K, device=dev); W = torch.randn(K, max(ld, N), device=dev)
        args = (dY, K, 1, W, W.shape[1], 1)
        byts = 4 * (M * K + M * N)
    else:
        dY = torch.randn(K, M, device=dev); X = torch.randn(K, N, device=dev)
        args = (dY, 1, M, X, N, 1)
        byts = 4 * (K * M + K * N)
    C = torch.zeros(M, N, device=dev)
    acc = 2 if split > 1 else 0
    for _ in range(2):
        mm.gemm(*args, C, N, M, N, K, act=int(os.environ.get('GEMM_ACT', '0')), accumulate=acc, split_k=split, kernel=kernel)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        mm.gemm(*args, C, N, M, N, K, act=int(os.environ.get('GEMM_ACT', '0')), accumulate=acc, split_k=split, kernel=kernel)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    return ms, byts / ms / 1e6


if __name__ == '__main__':
    for s in SHAPES:
        out = []
        for kernel in (2, 3):
            ms, gbs = run(*s, kernel)
            out.append(f'k{kernel}: {ms:7.3f} ms {gbs:7.0f} GB/s')
        print(f'{s[0]:12s} M={s[2]:7d} N={s[3]:4d} K={s[4]:7d}  ' + '   '.join(out), flush=True)
