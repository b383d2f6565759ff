"""GPU parity tests of the training step through the C ABI (mst_model_train_step / mst_model_loss / mst_adam_step)
against the plain-torch fp32 oracle (oracle/model_oracle.py, pinned to the unmodified reference by
tests/test_oracle_model.py) and against the reference goldens (tests/golden/model_small.npz).

Tolerances (SURVEY.md 8(d), fp32 path): 1e-5 relative on loss leaves (2e-5 allowed for the sums' reduction order),
1e-4 on gradients measured as ||g - g_ref|| / ||g_ref|| per parameter (plus element-wise on the golden heads)."""
import ast
import os

import numpy as np
import pytest

import golden_util as gu

pytestmark = pytest.mark.gpu

LOSS_RTOL = 2e-5
GRAD_RTOL = 2e-4


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
    return sd, mm.flatten_state_dict(sd)


def flat_losses(d, prefix=''):
    out = {}
    for k, v in d.items():
        if isinstance(v, dict):
            out.update(flat_losses(v, prefix + k + '/'))
        elif v is not None:
            out[prefix + k] = v
    return out


def _batch(torch, samples):
    cat = lambda k: torch.cat([s[k] for s in samples]).cuda().contiguous()
    u = None
    if samples[0]['unpitched'] is not None:
        u = torch.cat([s['unpitched'] for s in samples]).cuda().contiguous()
    return cat('pitched'), u, cat('feats'), cat('mode'), cat('bpm')


def oracle_step(torch, sd, samples):
    """Oracle losses per sample and summed parameter gradients (iter_size-style accumulation)."""
    import model_oracle as mo
    sdg = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    leaves = []
    for s in samples:
        losses, _ = mo.training_losses(sdg, s['mode'], s['bpm'], s['pitched'], s['feats'], s['unpitched'], s['bpm_target'])
        losses['total'].sum().backward()
        leaves.append({k: float(v) for k, v in flat_losses(losses).items()})
    grads = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in sdg.items()}
    return leaves, grads


def run_ours(torch, flat, samples, case, normalize=True):
    import model_oracle as mo
    from mst_native import model as mm
    N = len(samples)
    p, u, f, m, b = _batch(torch, samples)
    it = torch.cat([mo.used_instruments(s['feats'], s['unpitched']) for s in samples]).cuda()
    bt = torch.tensor([float(s['bpm_target']) for s in samples], device='cuda')
    ws = mm.Workspace(N, case['C'], case['B'], case['nb'], case.get('drums', True), training=True)
    grads = torch.zeros_like(flat)
    losses, outs = ws.train_step(flat, grads, p, u, f, m, b, it, bt, normalize=normalize)
    torch.cuda.synchronize()
    return losses.cpu().numpy(), grads, ws


def grad_report(torch, grads_flat, ref_grads):
    from mst_native import model as mm
    rows = []
    for name, off, numel in mm.param_table():
        g = grads_flat[off:off + numel].detach().cpu().double()
        r = ref_grads[name].detach().reshape(-1).double()
        rn = float(r.norm())
        err = float((g - r).norm())
        rows.append((name, err / max(rn, 1e-12) if rn > 0 else err, rn, float(g.norm())))
    return rows


CASES = [dict(C=3, B=12, nb=4, drums=True), dict(C=3, B=12, nb=4, drums=False), dict(C=2, B=5, nb=3, drums=True),
         dict(C=1, B=7, nb=4, drums=True), dict(C=6, B=9, nb=4, drums=True, density=0.01)]


@pytest.mark.parametrize('case', CASES)
def test_losses_and_gradients_match_oracle(torch, weights, case):
    import model_oracle as mo
    from mst_native import model as mm
    sd, flat = weights
    samples = [mo.synthetic_sample(seed=200 + i, **case) for i in range(2)]
    ref_leaves, ref_grads = oracle_step(torch, sd, samples)
    losses, grads, ws = run_ours(torch, flat, samples, case)
    assert ws.launches > 100
    bad = []
    for i, ref in enumerate(ref_leaves):
        for k, v in ref.items():
            got = float(losses[i, mm.LOSS_LEAVES[k]])
            if abs(got - v) > LOSS_RTOL * abs(v) + 1e-9:
                bad.append(f'loss[{i}] {k}: got {got!r} want {v!r}')
    rows = grad_report(torch, grads, ref_grads)
    report = '\n'.join(f'{n:70s} rel_err {e:9.2e}  |ref| {rn:9.3e} |got| {gn:9.3e}' for n, e, rn, gn in rows)
    os.makedirs(os.path.join(gu.ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(gu.ROOT, 'gpurun_out', f'grad_report_C{case["C"]}_B{case["B"]}_{int(case["drums"])}.txt'), 'w') as fh:
        fh.write('\n'.join(bad) + '\n' + report + '\n')
    for n, e, rn, gn in rows:
        if rn > 0 and e > GRAD_RTOL:
            bad.append(f'grad {n}: rel err {e:.3e} (|ref| {rn:.3e}, |got| {gn:.3e})')
        if rn == 0 and gn > 1e-8:
            bad.append(f'grad {n}: expected zero, got norm {gn:.3e}')
    assert not bad, '\n'.join(bad[:60])


FULL_SIZE = [dict(C=4, B=200, nb=4, drums=True), dict(C=8, B=800, nb=4, drums=True)]


@pytest.mark.parametrize('case', FULL_SIZE, ids=['config4_C4_B200', 'config5_long_C8_B800'])
def test_full_size_shapes_match_oracle(torch, weights, case):
    """BASELINE.json configs[3] / configs[4] at their full per-sample sizes (800 and 6400 channel-bars, the latter
    1.4e7 note rows and 800-step bar LSTMs): every loss leaf and every parameter gradient against the CPU oracle."""
    import model_oracle as mo
    from mst_native import model as mm
    sd, flat = weights
    samples = [mo.synthetic_sample(seed=7, **case)]
    ref_leaves, ref_grads = oracle_step(torch, sd, samples)
    losses, grads, _ = run_ours(torch, flat, samples, case)
    for k, v in ref_leaves[0].items():
        got = float(losses[0, mm.LOSS_LEAVES[k]])
        assert abs(got - v) <= LOSS_RTOL * abs(v) + 1e-9, (case, k, got, v)
    bad = [f'grad {n}: rel err {e:.3e} (|ref| {rn:.3e}, |got| {gn:.3e})'
           for n, e, rn, gn in grad_report(torch, grads, ref_grads) if (rn > 0 and e > GRAD_RTOL) or (rn == 0 and gn > 1e-8)]
    assert not bad, '\n'.join(bad[:60])


def test_training_matches_reference_goldens(torch, weights):
    """Loss leaves and per-parameter gradient norms / heads of the UNMODIFIED reference (oracle/gen_golden_model.py)."""
    import model_oracle as mo
    from mst_native import model as mm
    sd, flat = weights
    Z = np.load(os.path.join(gu.GOLDEN_DIR, 'model_small.npz'))
    table = {n: (o, k) for n, o, k in mm.param_table()}
    names = [str(x) for x in Z['param_names']]
    for i in range(int(Z['n_cases'])):
        pre = f'{i:02d}/'
        case = ast.literal_eval(str(Z[pre + 'case']))
        s = mo.synthetic_sample(**case)
        losses, grads, _ = run_ours(torch, flat, [s], case)
        for k, v in zip([str(x) for x in Z[pre + 'loss_names']], Z[pre + 'loss_values']):
            got = float(losses[0, mm.LOSS_LEAVES[k]])
            assert abs(got - v) <= LOSS_RTOL * abs(v) + 1e-9, (case, k, got, v)
        for j, name in enumerate(names):
            off, numel = table[name]
            g = grads[off:off + numel].cpu()
            ref_norm = float(Z[pre + 'grad_norms'][j])
            assert abs(float(g.double().norm()) - ref_norm) <= GRAD_RTOL * max(ref_norm, 1e-8), (case, name)
            head = g[:8].numpy()
            np.testing.assert_allclose(head, Z[pre + 'grad_heads'][j][:len(head)], rtol=1e-3,
                                       atol=GRAD_RTOL * max(ref_norm, 1e-8), err_msg=f'{case} {name}')


def test_gradients_accumulate_across_calls(torch, weights):
    """Two calls into the same gradient buffer = the sum (train-model.py's iter_size accumulation)."""
    import model_oracle as mo
    from mst_native import model as mm
    sd, flat = weights
    case = dict(C=2, B=6, nb=4, drums=True)
    s0, s1 = mo.synthetic_sample(seed=31, **case), mo.synthetic_sample(seed=32, **case)
    _, g_both, _ = run_ours(torch, flat, [s0, s1], case)
    ws = mm.Workspace(1, 2, 6, 4, True, training=True)
    grads = torch.zeros_like(flat)
    for s in (s0, s1):
        p, u, f, m, b = _batch(torch, [s])
        it = mo.used_instruments(s['feats'], s['unpitched']).cuda()
        ws.train_step(flat, grads, p, u, f, m, b, it, torch.tensor([120.], device='cuda'))
    torch.cuda.synchronize()
    err = float((grads - g_both).norm() / g_both.norm())
    assert err < 1e-5, err


@pytest.mark.parametrize('normalize', [False, True])
def test_loss_entry_matches_oracle(torch, normalize):
    import model_oracle as mo
    from mst_native import model as mm
    g = torch.Generator().manual_seed(5)
    N, C, B, nb = 3, 2, 4, 4
    samples = [mo.synthetic_sample(seed=60 + i, C=C, B=B, nb=nb, drums=True, density=0.02) for i in range(N)]
    pp = torch.rand(N, C, B, nb, 10, 56, 5, generator=g) * 0.98 + 0.01
    pp[..., 0] *= 6
    pu = torch.rand(N, 1, B, nb, 10, 47, 2, generator=g) * 0.98 + 0.01
    il, ml = torch.randn(N, 41, generator=g), torch.randn(N, 2, generator=g)
    bp = torch.rand(N, generator=g) * 150 + 50
    it = (torch.rand(N, 41, generator=g) < 0.2).float()
    mt = torch.tensor([[1., 0.], [0., 1.], [1., 0.]])
    bt = torch.tensor([120., 90., 171.])
    tp = torch.cat([s['pitched'] for s in samples])
    tu = torch.cat([s['unpitched'] for s in samples])
    got = mm.total_loss(*[t.cuda() for t in (pp, tp, pu, tu, il, it, ml, mt, bp, bt)], normalize=normalize).cpu().numpy()
    for i in range(N):
        ref = mo.total_loss(il[i:i + 1], it[i:i + 1], bp[i:i + 1], bt[i:i + 1], ml[i:i + 1], mt[i:i + 1], pp[i:i + 1],
                            tp[i:i + 1], pu[i:i + 1], tu[i:i + 1], normalize=normalize)
        for k, v in flat_losses(ref).items():
            v = float(v)
            assert abs(float(got[i, mm.LOSS_LEAVES[k]]) - v) <= LOSS_RTOL * abs(v) + 1e-9, (i, k, got[i], v)


def test_adam_matches_torch(torch):
    from mst_native import model as mm
    g = torch.Generator().manual_seed(3)
    n = 100003
    p0 = torch.randn(n, generator=g)
    ref_p = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref_p], lr=0.01)
    sched = torch.optim.lr_scheduler.StepLR(opt, 2, gamma=0.9)
    p = p0.clone().cuda()
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    lr = 0.01
    for step in range(1, 6):
        grad = torch.randn(n, generator=g) * (1 + step)
        ref_p.grad = grad.clone()
        opt.step(); sched.step()
        gbuf = grad.clone().cuda()
        mm.adam_step(p, gbuf, m, v, step, lr)
        lr = 0.01 * 0.9 ** (step // 2)
        assert float(gbuf.abs().max()) == 0.0          # zero_grad fused
        torch.testing.assert_close(p.cpu(), ref_p.detach(), rtol=1e-5, atol=1e-6)
