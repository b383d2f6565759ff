"""Golden vectors for the model / loss / gradient path, produced by the UNMODIFIED reference
(style/model.py imported from /root/reference) with the seed-108 weights of train-model.py:52.

    python oracle/gen_golden_model.py        # writes tests/golden/model_small.npz

Per case (synthetic sample from model_oracle.synthetic_sample): every loss leaf of get_total_loss
(normalize=True, the positional call of train-model.py:115-122), the predictions, and for every parameter
the gradient's L2 norm plus its first 8 values.  TEST INFRASTRUCTURE.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness  # noqa: E402

ref = ref_harness.load()
import torch  # noqa: E402
import model_oracle as mo  # noqa: E402

CASES = [dict(seed=7, C=3, B=12, nb=4, drums=True), dict(seed=8, C=3, B=12, nb=4, drums=False),
         dict(seed=9, C=2, B=5, nb=3, drums=True), dict(seed=10, C=1, B=7, nb=4, drums=True, density=0.01)]


def flat_losses(d, prefix=''):
    out = {}
    for k, v in d.items():
        if isinstance(v, dict):
            out.update(flat_losses(v, prefix + k + '/'))
        elif v is not None:
            out[prefix + k] = float(v)
    return out


def main():
    torch.set_num_threads(8)
    model = ref_harness.build_model(108)
    names = [k for k, _ in model.named_parameters()]
    h = hashlib.sha256()
    sd = model.state_dict()
    for k in sorted(sd):
        h.update(k.encode()); h.update(sd[k].numpy().tobytes())
    flat = {'weights_sha256': np.frombuffer(h.digest(), dtype=np.uint8), 'n_cases': np.int32(len(CASES)),
            'param_names': np.array(names)}
    for i, case in enumerate(CASES):
        s = mo.synthetic_sample(**case)
        u = s['unpitched']
        model.zero_grad()
        (ip, mp, bp), xp, xu = model(s['mode'], s['bpm'], s['pitched'], s['feats'], u)
        used = ref.data.get_used_instruments(s['feats'], u)
        losses = ref.model.get_total_loss(ip, used, bp, s['bpm_target'], mp, s['mode'], xp, s['pitched'], xu, u,
                                          normalize=True)
        losses['total'].backward()
        fl = flat_losses(losses)
        pre = f'{i:02d}/'
        flat[pre + 'case'] = np.array(repr(case))
        flat[pre + 'loss_names'] = np.array(list(fl))
        flat[pre + 'loss_values'] = np.array(list(fl.values()), dtype=np.float64)
        flat[pre + 'instruments_pred'] = ip.detach().numpy()
        flat[pre + 'mode_pred'] = mp.detach().numpy()
        flat[pre + 'bpm_pred'] = bp.detach().numpy()
        flat[pre + 'pitched_pred_sum'] = xp.detach().double().sum((0, 2, 3, 4, 5)).numpy()      # (C, 5)
        flat[pre + 'pitched_pred_head'] = xp.detach()[0, :, 0, 0].numpy()                        # (C, 10, 56, 5)
        if xu is not None:
            flat[pre + 'unpitched_pred_head'] = xu.detach()[0, 0, 0, 0].numpy()
        style, melody, rhythm = model.extract_style(s['mode'], s['bpm'], s['pitched'], s['feats'], u)
        flat[pre + 'style'] = style.detach().numpy()
        flat[pre + 'rhythm_head'] = rhythm.detach()[0, 0].numpy()
        flat[pre + 'melody_head'] = melody.detach()[0, 0, 0].numpy()
        norms, heads = [], []
        for k, prm in model.named_parameters():
            g = prm.grad if prm.grad is not None else torch.zeros_like(prm)
            norms.append(float(g.double().norm()))
            hd = np.zeros(8, np.float32)
            v = g.reshape(-1)[:8].numpy()
            hd[:len(v)] = v
            heads.append(hd)
        flat[pre + 'grad_norms'] = np.array(norms)
        flat[pre + 'grad_heads'] = np.stack(heads)
        print(case, fl['total'])
    path = os.path.join(os.path.dirname(HERE), 'tests', 'golden', 'model_small.npz')
    np.savez_compressed(path, **flat)
    print(path, os.path.getsize(path) // 1024, 'KiB')


if __name__ == '__main__':
    main()
