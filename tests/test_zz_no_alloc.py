"""GPU test of mst_model_loss_noalloc (include/mst_b200.h): the allocation-free variant of mst_model_loss whose kernel scratch
is a static device array inside the library.  It has no counterpart in the reference,
so the check is against the EXISTING entry point mst_model_loss on the same inputs, within the loss tolerance of
tests/test_gpu_train.py (2e-5 relative: both paths accumulate their per-sample sums with double-precision atomics, whose
order differs from launch to launch), on two calls in a row that share the scratch (different batch sizes and inputs, so
stale contents of the first call would show in the second).

Named test_zz_* so that pytest collects it last."""
import ctypes

import pytest

pytestmark = pytest.mark.gpu

LOSS_RTOL = 2e-5


@pytest.fixture(scope='module')
def torch():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('GPU tests need a CUDA device (no CPU fallback exists)')
    return torch


def make_inputs(torch, seed, N, C, B, nb, drums):
    """Predictions in (0, 1) (durations up to 6), sparse 0/1-ish targets: the shapes get_total_loss takes, on the device."""
    g = torch.Generator().manual_seed(seed)
    pp = torch.rand(N, C, B, nb, 10, 56, 5, generator=g) * 0.98 + 0.01
    pp[..., 0] *= 6
    from mst_native import synth
    samples = [synth.synthetic_sample(seed * 100 + i, C=C, B=B, nb=nb, drums=drums, density=0.02) for i in range(N)]
    tp = torch.cat([s['pitched'] for s in samples])
    pu = tu = None
    if drums:
        pu = torch.rand(N, 1, B, nb, 10, 47, 2, generator=g) * 0.98 + 0.01
        tu = torch.cat([s['unpitched'] for s in samples])
    il, ml = torch.randn(N, 41, generator=g), torch.randn(N, 2, generator=g)
    it = (torch.rand(N, 41, generator=g) < 0.2).float()
    mt = torch.zeros(N, 2)
    mt[torch.arange(N), torch.randint(0, 2, (N,), generator=g)] = 1
    bp, bt = torch.rand(N, generator=g) * 150 + 50, torch.rand(N, generator=g) * 120 + 60
    dev = lambda t: None if t is None else t.cuda().contiguous()
    return [dev(t) for t in (pp, tp, pu, tu, il, it, ml, mt, bp, bt)]


def assert_same(torch, got, ref):
    got, ref = got.double().cpu(), ref.double().cpu()
    assert torch.isfinite(ref).all()
    bad = (got - ref).abs() > LOSS_RTOL * ref.abs() + 1e-9
    assert not bool(bad.any()), (got[bad][:8], ref[bad][:8])


def test_noalloc_loss_matches_existing_entry_twice_in_a_row(torch):
    from mst_native import model as mm
    cases = [dict(seed=11, N=5, C=3, B=4, nb=4, drums=True, normalize=True),
             dict(seed=12, N=2, C=2, B=6, nb=4, drums=False, normalize=False),      # fewer samples: stale sums must not leak
             dict(seed=13, N=5, C=3, B=4, nb=4, drums=True, normalize=False),
             dict(seed=14, N=256, C=1, B=1, nb=4, drums=True, normalize=True)]       # MST_LOSS_NOALLOC_MAX_SAMPLES exactly
    prepared = []
    for c in cases:
        t = make_inputs(torch, c['seed'], c['N'], c['C'], c['B'], c['nb'], c['drums'])
        ref = mm.total_loss(*t, normalize=c['normalize'])                               # existing entry point (caller workspace)
        out = torch.full((c['N'], mm.N_LOSS_LEAVES), float('nan'), device='cuda')        # caller-owned output
        prepared.append((c, t, ref, out))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free_before = torch.cuda.mem_get_info()[0]
    for c, t, ref, out in prepared:                  # back to back on one stream, sharing the static scratch
        mm.total_loss_noalloc(out, *t, normalize=c['normalize'])
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert free_before - torch.cuda.mem_get_info()[0] < (1 << 20)      # nothing was taken from the device behind torch's back either
    for c, t, ref, out in prepared:
        assert_same(torch, out, ref)


def test_noalloc_loss_through_raw_c_abi_and_limits(torch):
    """The plain ctypes call a C caller would make, then the size limit and null checks."""
    from mst_native import lib, model as mm
    L = lib.require_gpu()
    mm._L()                                           # installs argtypes
    N, C, B, nb = 3, 2, 4, 4
    t = make_inputs(torch, 21, N, C, B, nb, True)
    ref = mm.total_loss(*t, normalize=True)
    out = torch.zeros(N, mm.N_LOSS_LEAVES, device='cuda')
    shape = mm.ModelShape(N, C, B, nb, 1, 0)
    li = mm.LossIn(*[x.data_ptr() for x in t], 1, 0)
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(2):                                # twice in a row into the same output
        out.fill_(float('nan'))
        assert L.mst_model_loss_noalloc(ctypes.byref(shape), ctypes.byref(li), out.data_ptr(), stream) == 0
        torch.cuda.synchronize()
        assert_same(torch, out, ref)
    big = mm.ModelShape(mm.LOSS_NOALLOC_MAX_SAMPLES + 1, C, B, nb, 1, 0)
    rc = L.mst_model_loss_noalloc(ctypes.byref(big), ctypes.byref(li), out.data_ptr(), stream)
    assert rc != 0 and b'MST_LOSS_NOALLOC_MAX_SAMPLES' in L.mst_last_error()
    zero = mm.ModelShape(0, C, B, nb, 1, 0)
    assert L.mst_model_loss_noalloc(ctypes.byref(zero), ctypes.byref(li), out.data_ptr(), stream) != 0
    assert L.mst_model_loss_noalloc(ctypes.byref(shape), ctypes.byref(li), None, stream) != 0
    torch.cuda.synchronize()
    with pytest.raises(lib.MstError):
        mm.total_loss_noalloc(torch.zeros(mm.LOSS_NOALLOC_MAX_SAMPLES + 1, mm.N_LOSS_LEAVES, device='cuda'),
                              *[None if x is None else x[:1].expand(mm.LOSS_NOALLOC_MAX_SAMPLES + 1, *x.shape[1:]).contiguous()
                                for x in make_inputs(torch, 22, 1, 1, 1, 4, False)])
