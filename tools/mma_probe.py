"""Is k_gemm_tc bound by the mma.sync rate?  L2-resident operands, varying K (development probe)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))
import torch
from mst_native import model as mm
for (M, N, K, split) in [(16384, 64, 512, 1), (16384, 64, 1024, 1), (16384, 64, 2048, 1), (37888, 64, 512, 1), (37888, 32, 512, 1), (37888, 64, 1024, 1)]:
    X = torch.randn(M, K, device='cuda'); W = torch.randn(N, K, device='cuda'); C = torch.zeros(M, N, device='cuda')
    for kernel in (1, 2):
        for _ in range(3):
            mm.gemm(X, K, 1, W, 1, K, C, N, M, N, K, kernel=kernel)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            mm.gemm(X, K, 1, W, 1, K, C, N, M, N, K, kernel=kernel)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        flop = 2.0 * M * N * K
        print(f'M={M} N={N} K={K} kernel={kernel}: {ms*1e3:8.1f} us  {flop/ms/1e9:7.1f} TFLOP/s(fp32-equivalent)  A={M*K*4/1e6:.0f} MB  {M*K*4/ms/1e6:.0f} GB/s', flush=True)
