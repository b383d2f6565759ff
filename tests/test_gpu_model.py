"""GPU parity tests of the model path through the C ABI (mst_model_*) against the plain-torch fp32 oracle
(oracle/model_oracle.py, itself pinned to the reference) and the reference goldens.
Tolerances: fp32 path 1e-5 relative on predictions/loss leaves, 1e-4 on gradients."""
import ast
import os

import numpy as np
import pytest

import golden_util as gu

pytestmark = pytest.mark.gpu

FWD_RTOL, FWD_ATOL = 2e-5, 2e-6


@pytest.fixture(scope='module')
def torch():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('GPU tests need a CUDA device (no CPU fallback exists)')
    return torch


@pytest.fixture(scope='module')
def weights(torch):
    import model_oracle as mo
    from mst_native import model as mm
    sd = mo.make_state_dict(108)
    assert mm.total_params() == 980325 == sum(v.numel() for v in sd.values())
    assert {n for n, _, _ in mm.param_table()} == set(sd)
    return sd, mm.flatten_state_dict(sd)


def _batch(torch, samples):
    cat = lambda k: torch.cat([s[k] for s in samples]).cuda().contiguous()
    u = None
    if samples[0]['unpitched'] is not None:
        u = torch.cat([s['unpitched'] for s in samples]).cuda().contiguous()
    return cat('pitched'), u, cat('feats'), cat('mode'), cat('bpm')


def _close(torch, got, want, name, rtol=FWD_RTOL, atol=FWD_ATOL):
    got, want = got.detach().float().cpu(), want.detach().float().cpu()
    assert got.shape == want.shape, (name, got.shape, want.shape)
    err = (got - want).abs()
    tol = atol + rtol * want.abs()
    assert bool((err <= tol).all()), f'{name}: max err {float(err.max()):.3e} (max |ref| {float(want.abs().max()):.3e})'


@pytest.mark.parametrize('case', [dict(C=3, B=12, nb=4, drums=True), dict(C=3, B=12, nb=4, drums=False),
                                  dict(C=2, B=5, nb=3, drums=True), dict(C=1, B=7, nb=4, drums=True),
                                  dict(C=4, B=40, nb=4, drums=True)])
def test_forward_matches_oracle(torch, weights, case):
    import model_oracle as mo
    from mst_native import model as mm
    sd, flat = weights
    samples = [mo.synthetic_sample(seed=100 + i, **case) for i in range(2)]      # N = 2 independent samples
    p, u, f, m, b = _batch(torch, samples)
    ws = mm.Workspace(2, case['C'], case['B'], case['nb'], case['drums'])
    out = ws.forward(flat, p, u, f, m, b)
    torch.cuda.synchronize()
    assert ws.launches > 50
    for i, s in enumerate(samples):
        with torch.no_grad():
            style, melody, rhythm = mo.extract_style(sd, s['mode'], s['bpm'], s['pitched'], s['feats'], s['unpitched'])
            (ip, mp, bp), xp, xu = mo.forward(sd, s['mode'], s['bpm'], s['pitched'], s['feats'], s['unpitched'])
        _close(torch, out['style'][i:i + 1], style, 'style')
        _close(torch, out['rhythm'][i:i + 1], rhythm, 'rhythm')
        _close(torch, out['melody'][i:i + 1], melody, 'melody')
        _close(torch, out['instruments'][i:i + 1], ip, 'instruments')
        _close(torch, out['mode'][i:i + 1], mp, 'mode')
        _close(torch, out['bpm'][i:i + 1], bp, 'bpm', rtol=1e-6)
        _close(torch, out['pitched'][i:i + 1], xp, 'pitched')
        if xu is not None:
            _close(torch, out['unpitched'][i:i + 1], xu, 'unpitched')


def test_forward_matches_reference_goldens(torch, weights):
    """Same inputs as oracle/gen_golden_model.py ran through the unmodified reference model."""
    import model_oracle as mo
    from mst_native import model as mm
    sd, flat = weights
    Z = np.load(os.path.join(gu.GOLDEN_DIR, 'model_small.npz'))
    for i in range(int(Z['n_cases'])):
        pre = f'{i:02d}/'
        case = ast.literal_eval(str(Z[pre + 'case']))
        s = mo.synthetic_sample(**case)
        p, u, f, m, b = _batch(torch, [s])
        ws = mm.Workspace(1, case['C'], case['B'], case['nb'], case.get('drums', True))
        out = ws.forward(flat, p, u, f, m, b)
        _close(torch, out['instruments'], torch.from_numpy(Z[pre + 'instruments_pred']), 'instruments')
        _close(torch, out['bpm'], torch.from_numpy(Z[pre + 'bpm_pred']), 'bpm', rtol=1e-6)
        _close(torch, out['style'], torch.from_numpy(Z[pre + 'style']), 'style')
        _close(torch, out['pitched'][0, :, 0, 0], torch.from_numpy(Z[pre + 'pitched_pred_head']), 'pitched head')
        _close(torch, out['melody'][0, 0, 0], torch.from_numpy(Z[pre + 'melody_head']), 'melody head')
        _close(torch, out['rhythm'][0, 0], torch.from_numpy(Z[pre + 'rhythm_head']), 'rhythm head')
