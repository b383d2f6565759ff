"""Golden vectors for the drop-in package (tests/golden/dropin.npz).  TEST INFRASTRUCTURE ONLY.

Runs the UNMODIFIED reference in this container (oracle/ref_harness.py) on small MIDI files and records
  * style_transfer.transfer_style(model, A, [B, C], out) with the seed-108 model: every file it writes;
  * the first iterations of train-model.py's loop (:92-160, iter_size=2, Adam lr=.01, StepLR(200, .9)) over the same
    files: the flattened loss dict of every iteration and the norms / heads of the weights afterwards.
The input files are SMF blobs the reference itself wrote for the synthetic songs of codec_synth.npz ('decoded_smf'),
so the fixture needs nothing from /root/reference at test time.

    python oracle/gen_golden_dropin.py
"""
import glob
import math
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import ref_harness                     # noqa: E402

ref = ref_harness.load()

INPUT_SONGS = {'songA': 0, 'songB': 7, 'songC': 2}     # indices into codec_synth.npz
N_TRAIN_ITERATIONS = 6


def golden_inputs(tmp):
    z = np.load(os.path.join(ROOT, 'tests', 'golden', 'codec_synth.npz'))
    paths = {}
    for name, i in INPUT_SONGS.items():
        paths[name] = os.path.join(tmp, name + '.mid')
        with open(paths[name], 'wb') as f:
            f.write(bytes(z[f'{i:03d}/decoded_smf']))
    return paths


def run_transfer(paths, out):
    model = ref_harness.build_model(108)
    with torch.no_grad():
        ref.style_transfer.transfer_style(model, paths['songA'], [paths['songB'], paths['songC']], out)
    files = {}
    for f in sorted(glob.glob(os.path.join(out, '**', '*.mid'), recursive=True)):
        files[os.path.relpath(f, out)] = open(f, 'rb').read()
    # intermediate values for diagnosis
    inp = ref.style_transfer.get_model_input(paths['songA'])
    style, melody, rhythm = ref.style_transfer.extract_style(model, inp)
    with torch.no_grad():
        ins, mode, bpm = model.predict_song_info(style, rhythm)
    extra = dict(style=style.numpy(), melody_head=melody.numpy().ravel()[:256], rhythm_head=rhythm.numpy().ravel()[:256],
                 instruments_pred=ins.numpy(), mode_pred=mode.numpy(), bpm_pred=np.float32(float(bpm)))
    return files, extra


def run_training(paths):
    d, m = ref.data, ref.model
    model = ref_harness.build_model(108)
    optimizer = torch.optim.Adam(model.parameters(), lr=.01)
    scheduler = torch.optim.lr_scheduler.StepLR(optimizer, step_size=200, gamma=.9)
    files = [paths['songA'], paths['songB'], paths['songC']]
    inputs = d.iter_inputs(files, d.included_instruments, min_n_messages=0, shuffle=False, looped=True)
    iter_size = 2
    rows, names = [], None
    optimizer.zero_grad()
    for iteration in range(N_TRAIN_ITERATIONS):
        input = next(inputs)
        filename, (info, pitched_channels, instruments_features, instruments, unpitched_channels) = input
        max_n_bars = 800 // pitched_channels.shape[0]
        mode, bpm, pitched_channels, instruments_features, unpitched_channels = d.prepare_input(input, max_n_bars)
        if unpitched_channels is not None and unpitched_channels.sum() == 0:
            unpitched_channels = None
        used = d.get_used_instruments(instruments_features, unpitched_channels)
        (ip, mp, bp), pp, up = model(mode, bpm, pitched_channels, instruments_features, unpitched_channels)
        losses = m.get_total_loss(ip, used, bp, info['bpm'], mp, mode, pp, pitched_channels, up, unpitched_channels,
                                  normalize=True)
        loss = losses['total']
        assert not math.isnan(loss)
        loss.backward()
        row = {}

        def walk(prefix, x):
            if isinstance(x, dict):
                for k, v in x.items():
                    walk(prefix + [k], v)
            else:
                row['/'.join(prefix)] = float('nan') if x is None else float(x)
        walk([], losses)
        if names is None:
            names = sorted(row)
        rows.append([row.get(k, float('nan')) for k in names])
        print(iteration, os.path.basename(filename), row['total'])
        if (iteration + 1) % iter_size == 0:
            optimizer.step()
            optimizer.zero_grad()
            scheduler.step()
    sd = model.state_dict()
    pnames = sorted(sd)
    return dict(loss_names=np.array(names), losses=np.array(rows, dtype=np.float64),
                param_names=np.array(pnames),
                param_norms=np.array([float(sd[k].double().norm()) for k in pnames]),
                param_heads=np.stack([np.pad(sd[k].numpy().ravel()[:8], (0, max(0, 8 - sd[k].numel()))) for k in pnames]))


def main():
    flat = {}
    with tempfile.TemporaryDirectory() as tmp:
        paths = golden_inputs(tmp)
        out = os.path.join(tmp, 'out')
        files, extra = run_transfer(paths, out)
        names = sorted(files)
        flat['transfer/names'] = np.array(names)
        for i, n in enumerate(names):
            flat[f'transfer/file{i:02d}'] = np.frombuffer(files[n], dtype=np.uint8)
            print(n, len(files[n]))
        for k, v in extra.items():
            flat['transfer/' + k] = v
        for k, v in run_training(paths).items():
            flat['train/' + k] = v
    flat['input_songs'] = np.array([f'{k}={v}' for k, v in INPUT_SONGS.items()])
    path = os.path.join(ROOT, 'tests', 'golden', 'dropin.npz')
    np.savez_compressed(path, **flat)
    print(path, os.path.getsize(path) // 1024, 'KiB')


if __name__ == '__main__':
    main()
