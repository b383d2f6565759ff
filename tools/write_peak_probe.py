"""Write-only HBM bandwidth on this GPU (torch fill_ / cudaMemsetAsync of a buffer far larger than L2): the ceiling a
grid-writing kernel such as k_grid can reach, next to MEASURED_PEAKS.json's copy figure (development probe)."""
import torch

x = torch.empty(8 << 30, dtype=torch.uint8, device='cuda')
y = torch.empty(4 << 30, dtype=torch.uint8, device='cuda')
f = x.view(torch.float32)


def timed(fn, byts, name, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f'{name:40s} {byts / ms / 1e6:8.0f} GB/s  ({ms:.3f} ms)', flush=True)


timed(lambda: x.zero_(), x.numel(), 'memset 8 GiB (zero_)')
timed(lambda: f.fill_(1.5), x.numel(), 'fill_ f32 8 GiB')
timed(lambda: y.copy_(x[:y.numel()]), 2 * y.numel(), 'copy 4 GiB (read + write bytes)')
