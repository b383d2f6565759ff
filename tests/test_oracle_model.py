"""CPU tests: the plain-torch model oracle (oracle/model_oracle.py) against golden vectors produced by the
unmodified reference model (oracle/gen_golden_model.py): seed-108 weights bit-identical, every loss leaf,
predictions and every parameter gradient within fp32 tolerance."""
import ast
import hashlib
import os

import numpy as np
import pytest
import torch

import golden_util as gu
import model_oracle as mo

Z = np.load(os.path.join(gu.GOLDEN_DIR, 'model_small.npz'))
LOSS_RTOL, GRAD_RTOL = 1e-5, 1e-4     # fp32 tolerances on loss leaves / gradients


def flat_losses(d, prefix=''):
    out = {}
    for k, v in d.items():
        if isinstance(v, dict):
            out.update(flat_losses(v, prefix + k + '/'))
        elif v is not None:
            out[prefix + k] = v
    return out


def test_seed_108_weights_identical_to_reference():
    sd = mo.make_state_dict(108)
    assert sum(v.numel() for v in sd.values()) == 980325
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode()); h.update(sd[k].numpy().tobytes())
    assert h.digest() == bytes(Z['weights_sha256'])


@pytest.mark.parametrize('i', range(int(Z['n_cases'])))
def test_model_oracle_matches_reference(i):
    torch.set_num_threads(4)
    pre = f'{i:02d}/'
    case = ast.literal_eval(str(Z[pre + 'case']))
    s = mo.synthetic_sample(**case)
    sd = {k: v.clone().requires_grad_(True) for k, v in mo.make_state_dict(108).items()}
    losses, (ip, mp, bp, xp, xu) = mo.training_losses(sd, s['mode'], s['bpm'], s['pitched'], s['feats'],
                                                       s['unpitched'], s['bpm_target'])
    fl = flat_losses(losses)
    assert list(fl) == [str(x) for x in Z[pre + 'loss_names']]
    np.testing.assert_allclose([float(v) for v in fl.values()], Z[pre + 'loss_values'], rtol=LOSS_RTOL)
    np.testing.assert_allclose(ip.detach().numpy(), Z[pre + 'instruments_pred'], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(mp.detach().numpy(), Z[pre + 'mode_pred'], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(bp.detach().numpy(), Z[pre + 'bpm_pred'], rtol=1e-6)
    np.testing.assert_allclose(xp.detach()[0, :, 0, 0].numpy(), Z[pre + 'pitched_pred_head'], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(xp.detach().double().sum((0, 2, 3, 4, 5)).numpy(), Z[pre + 'pitched_pred_sum'], rtol=1e-6)
    if xu is not None:
        np.testing.assert_allclose(xu.detach()[0, 0, 0, 0].numpy(), Z[pre + 'unpitched_pred_head'], rtol=1e-5, atol=1e-6)
    losses['total'].backward()
    names = [str(x) for x in Z['param_names']]
    for j, k in enumerate(names):
        g = sd[k].grad if sd[k].grad is not None else torch.zeros_like(sd[k])
        ref_norm = float(Z[pre + 'grad_norms'][j])
        assert abs(float(g.double().norm()) - ref_norm) <= GRAD_RTOL * max(ref_norm, 1e-8), k
        head = g.reshape(-1)[:8].numpy()
        np.testing.assert_allclose(head, Z[pre + 'grad_heads'][j][:len(head)], rtol=GRAD_RTOL,
                                   atol=GRAD_RTOL * max(ref_norm, 1e-8), err_msg=k)
