"""Golden vectors for BASELINE.json configs[0] and configs[1] on the BUNDLED example files.  TEST INFRASTRUCTURE ONLY.

Runs the UNMODIFIED reference in this container (oracle/ref_harness.py) and writes

  tests/golden/bundled_inputs.npz   the raw bytes of the 18 bundled input files the two configs read (data, not source:
                                    `examples/drums - midi/*.mid` and `examples/style transfer - midi/<Song>/original/*.mid`),
                                    because /root/reference does not exist on the GPU box;
  tests/golden/config0.npz          configs[0]: train-model.py's loop (:92-160) over the two drums files for one epoch
                                    (2 iterations, iter_size 2 => ONE Adam step, seed-108 weights, >= 100-message filter,
                                    crop 800 // C): every loss leaf of both iterations, the norm / head / SHA of every
                                    parameter after the step (BASELINE.md section 2 records totals 0.5969 / 0.6109);
  tests/golden/config1.npz          configs[1]: style_transfer.transfer_style (:22-54) for the four bundled folders with
                                    seed-108 weights (no trained snapshot exists): for every written file its size,
                                    SHA-256, number of note messages, one 64-bit digest + message count per
                                    (channel, bar) chunk, the full bytes of the `original/` files, and - because a file
                                    decoded from an untrained model's grid holds millions of messages - the exact
                                    message tuples of every SAMPLE_EVERY-th bar only (event-by-event comparison on
                                    that sample, chunk digests on the rest); plus the chosen instruments / mode / tempo.

    python oracle/gen_golden_configs.py [--config0] [--config1] [--inputs]
"""
import glob
import hashlib
import math
import os
import sys
import tempfile
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import ref_harness                     # noqa: E402

EXAMPLES = os.path.join(ref_harness.REFERENCE_ROOT, 'examples')
DRUMS = os.path.join(EXAMPLES, 'drums - midi')
TRANSFER = os.path.join(EXAMPLES, 'style transfer - midi')
GOLDEN = os.path.join(ROOT, 'tests', 'golden')
SAMPLE_EVERY = 16          # bars whose messages are stored verbatim: bar % SAMPLE_EVERY == 0


def transfer_folders():
    """[(song folder name, composition path, [style paths])] per SURVEY.md 8(d) config 2: the composition is the
    original/*.mid whose name starts with the folder name, the styles are the other three (sorted)."""
    out = []
    for folder in sorted(os.listdir(TRANSFER)):
        orig = sorted(glob.glob(os.path.join(glob.escape(os.path.join(TRANSFER, folder, 'original')), '*.mid')))
        comp = [p for p in orig if os.path.basename(p).startswith(folder)]
        assert len(comp) == 1, (folder, orig)
        out.append((folder, comp[0], [p for p in orig if p != comp[0]]))
    return out


def write_inputs():
    flat = {}
    names = []
    for p in sorted(glob.glob(os.path.join(glob.escape(DRUMS), '*.mid'))):
        names.append('drums/' + os.path.basename(p))
        flat[f'file{len(names) - 1:02d}'] = np.frombuffer(open(p, 'rb').read(), dtype=np.uint8)
    for folder, comp, styles in transfer_folders():
        for p in [comp] + styles:
            names.append(f'transfer/{folder}/' + os.path.basename(p))
            flat[f'file{len(names) - 1:02d}'] = np.frombuffer(open(p, 'rb').read(), dtype=np.uint8)
    flat['names'] = np.array(names)
    path = os.path.join(GOLDEN, 'bundled_inputs.npz')
    np.savez_compressed(path, **flat)
    print(path, os.path.getsize(path) // 1024, 'KiB', len(names), 'files')


def flatten_losses(losses):
    row = {}

    def walk(prefix, x):
        if isinstance(x, dict):
            for k, v in x.items():
                walk(prefix + [k], v)
        else:
            row['/'.join(prefix)] = float('nan') if x is None else float(x)
    walk([], losses)
    return row


def run_config0():
    ref = ref_harness.load()
    d, m = ref.data, ref.model
    model = ref_harness.build_model(108)
    optimizer = torch.optim.Adam(model.parameters(), lr=.01)
    scheduler = torch.optim.lr_scheduler.StepLR(optimizer, step_size=200, gamma=.9)
    files = sorted(glob.glob(os.path.join(glob.escape(DRUMS), '*.mid')))
    inputs = d.iter_inputs(files, d.included_instruments, shuffle=False, looped=False)     # one epoch, file order
    rows, names, shapes, fnames = [], None, [], []
    optimizer.zero_grad()
    for iteration, input in enumerate(inputs):
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
        row = flatten_losses(losses)
        names = names or sorted(row)
        rows.append([row.get(k, float('nan')) for k in names])
        shapes.append(list(pitched_channels.shape[1:4]) + [0 if unpitched_channels is None else 1, int(info['bpm'])])
        fnames.append(os.path.basename(filename))
        print('config0 iteration', iteration, os.path.basename(filename), tuple(pitched_channels.shape), row['total'])
        if (iteration + 1) % 2 == 0:
            grads = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
            optimizer.step()
            optimizer.zero_grad()
            scheduler.step()
    sd = model.state_dict()
    pnames = sorted(sd)
    flat = dict(
        files=np.array(fnames), shapes=np.array(shapes, dtype=np.int64), loss_names=np.array(names),
        losses=np.array(rows, dtype=np.float64), param_names=np.array(pnames),
        grad_norms=np.array([float(grads[k].double().norm()) for k in pnames]),
        param_norms=np.array([float(sd[k].double().norm()) for k in pnames]),
        param_sums=np.array([float(sd[k].double().sum()) for k in pnames]),
        param_heads=np.stack([np.pad(sd[k].numpy().ravel()[:8], (0, max(0, 8 - sd[k].numel()))) for k in pnames]))
    path = os.path.join(GOLDEN, 'config0.npz')
    np.savez_compressed(path, **flat)
    print(path, os.path.getsize(path) // 1024, 'KiB')


def chunk_digest(track_msgs, ticks_per_bar):
    """Note messages of a written track -> (rows (abs_time, is_on, channel, note, velocity), per-(channel, bar) count and
    64-bit digest).  The digest of a chunk is the first 8 bytes of SHA-256 over its rows in track order."""
    rows = []
    t = 0
    for m in track_msgs:
        t += m.time
        if m.type in ('note_on', 'note_off'):
            rows.append((t, 1 if m.type == 'note_on' else 0, m.channel, m.note, m.velocity))
    rows = np.array(rows, dtype=np.int32).reshape(-1, 5)
    return rows


def digest_rows(rows, ticks_per_bar):
    bars = rows[:, 0] // ticks_per_bar
    key = rows[:, 2].astype(np.int64) * (1 << 20) + bars
    order = np.argsort(key, kind='stable')
    rows_s, key_s = rows[order], key[order]
    uniq, start, count = np.unique(key_s, return_index=True, return_counts=True)
    dig = np.zeros(len(uniq), dtype=np.uint64)
    for i, (s, c) in enumerate(zip(start, count)):
        dig[i] = np.frombuffer(hashlib.sha256(np.ascontiguousarray(rows_s[s:s + c]).tobytes()).digest()[:8], dtype=np.uint64)[0]
    return uniq, count.astype(np.int32), dig


def run_config1():
    ref = ref_harness.load()
    st = ref.style_transfer
    flat = {}
    folders = transfer_folders()
    flat['folders'] = np.array([f for f, _, _ in folders])
    model = ref_harness.build_model(108)
    # record apply_style's host decisions (instruments / mode / tempo) by wrapping save()
    decisions = []
    orig_save = st.save

    def save_spy(cc, pitched, unpitched, instruments, save_path):
        decisions.append((save_path, [int(i) for i in instruments], cc.info['scale']['mode'] is ref.scales.major_mode,
                          int(cc.info['tempo']), cc.info['scale']['key'], unpitched is not None))
        return orig_save(cc, pitched, unpitched, instruments, save_path)
    st.save = save_spy
    for fi, (folder, comp, styles) in enumerate(folders):
        with tempfile.TemporaryDirectory() as tmp:
            t0 = time.time()
            del decisions[:]
            with torch.no_grad():
                st.transfer_style(model, comp, styles, tmp)
            print(f'config1 {folder}: transfer_style took {time.time() - t0:.0f} s')
            files = sorted(glob.glob(os.path.join(glob.escape(tmp), '**', '*.mid'), recursive=True))
            names = [os.path.relpath(f, tmp) for f in files]
            flat[f'{fi}/names'] = np.array(names)
            dec = {os.path.relpath(p, tmp): rest for p, *rest in decisions}
            for i, (f, n) in enumerate(zip(files, names)):
                data = open(f, 'rb').read()
                pre = f'{fi}/{i:02d}/'
                flat[pre + 'size'] = np.int64(len(data))
                flat[pre + 'sha256'] = np.frombuffer(hashlib.sha256(data).digest(), dtype=np.uint8)
                mid = ref.mido.MidiFile(f)
                tpb = mid.ticks_per_beat
                ts = [m for m in mid.tracks[0] if m.type == 'time_signature'][0]
                tpbar = int(tpb * ts.numerator)
                rows = chunk_digest(mid.tracks[0], tpbar)
                flat[pre + 'n_messages'] = np.int64(len(rows))
                flat[pre + 'ticks_per_bar'] = np.int64(tpbar)
                keys, counts, digs = digest_rows(rows, tpbar)
                flat[pre + 'chunk_keys'], flat[pre + 'chunk_counts'], flat[pre + 'chunk_digests'] = keys, counts, digs
                sample = rows[(rows[:, 0] // tpbar) % SAMPLE_EVERY == 0]
                flat[pre + 'sample_rows'] = sample
                ins, major, tempo, key, has_u = dec[n]
                flat[pre + 'decision'] = np.array([int(major), tempo, int(has_u), ref.scales.key_names.index(key)] + ins, dtype=np.int64)
                if '/original/' in '/' + n or len(data) < (1 << 18):
                    flat[pre + 'bytes'] = np.frombuffer(data, dtype=np.uint8)
                print(f'  {n}: {len(data)} bytes, {len(rows)} note messages, {len(sample)} sampled, instruments {ins}, '
                      f'major {major}, tempo {tempo}')
    st.save = orig_save
    flat['sample_every'] = np.int64(SAMPLE_EVERY)
    path = os.path.join(GOLDEN, 'config1.npz')
    np.savez_compressed(path, **flat)
    print(path, os.path.getsize(path) // 1024, 'KiB')


def run_create_midi():
    """tests/golden/create_midi.npz: the reference's create_midi (style/midi.py:120-168) on seeded random message lists
    (three channels incl. channel 9, np.float32 velocities, negative durations, long gaps) for two max_delta_time values:
    the message arrays and the bytes mid.save() wrote."""
    ref = ref_harness.load()
    NM = ref.midi_conversion.NoteMessage
    rng = np.random.default_rng(3)
    flat = {}
    info = {'ticks_per_beat': 480, 'tempo': 434783, 'time_signature': {'numerator': 3, 'denominator': 4}, 'ticks_per_bar': 1440}
    cases = [(1, {}), (0.25, {'duration': 400000}), (2, {})]
    for ci, (mdt, extra) in enumerate(cases):
        chans = []
        for ch, prog in ((0, 25), (9, -1), (3, 48)):
            msgs, t = [], 0
            for _ in range(150):
                t += int(rng.integers(0, 3000))
                msgs.append(NM('note_on', int(rng.integers(0, 128)), t, np.float32(rng.random())))
                msgs.append(NM('note_off', int(rng.integers(0, 128)), t + int(rng.integers(-50, 900)), 0))
            chans.append({'channel_id': ch, 'instrument_id': prog, 'messages': sorted(msgs, key=lambda m: m.time)})
        mid = ref.midi.create_midi(dict(info, **extra), *chans, max_delta_time=mdt)
        with tempfile.NamedTemporaryFile(suffix='.mid') as f:
            mid.save(f.name)
            flat[f'{ci}/bytes'] = np.frombuffer(open(f.name, 'rb').read(), dtype=np.uint8)
        flat[f'{ci}/max_delta_time'] = np.float64(mdt)
        flat[f'{ci}/duration'] = np.int64(extra.get('duration', -1))
        for k, c in enumerate(chans):
            flat[f'{ci}/{k}/channel'] = np.array([c['channel_id'], c['instrument_id']], dtype=np.int64)
            flat[f'{ci}/{k}/msgs'] = np.array([(m.type == 'note_on', m.note, m.time) for m in c['messages']], dtype=np.int64)
            flat[f'{ci}/{k}/velocity'] = np.array([m.velocity for m in c['messages']], dtype=np.float32)
    flat['n_cases'] = np.int64(len(cases))
    path = os.path.join(GOLDEN, 'create_midi.npz')
    np.savez_compressed(path, **flat)
    print(path, os.path.getsize(path) // 1024, 'KiB')


if __name__ == '__main__':
    if '--create-midi' in sys.argv:
        run_create_midi()
        sys.exit(0)
    todo = sys.argv[1:] or ['--inputs', '--config0', '--config1']
    if '--inputs' in todo:
        write_inputs()
    if '--config0' in todo:
        run_config0()
    if '--config1' in todo:
        run_config1()
