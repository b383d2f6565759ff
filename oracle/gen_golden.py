"""Generate golden vectors by running the UNMODIFIED reference in this container.

    python oracle/gen_golden.py            # writes tests/golden/*.npz

TEST INFRASTRUCTURE.  Needs /root/reference (so it only runs in the build container); the
vectors it writes are committed and are what the `-m "not gpu"` and `-m gpu` tests read.

Codec goldens (tests/golden/codec_bundled.npz, codec_synth.npz), one entry per song:
  input   events (MstEvent records), ticks_per_beat
  stage R read_midi  -> (channel_id, instrument_id, n_messages) per channel entry + info fields
  stage G get_input after the iter_inputs filter (min_n_messages=100 for bundled files, 0 for the
          synthetic ones) -> info['scale'], keys_dist, instrument lists, grid shapes, the f32
          grids as sparse (flat index, value) pairs + SHA-256 of the dense f32 bytes
  decode  decode_midi(cc, ...) of those grids -> the track's messages and the saved SMF bytes
"""
import hashlib
import glob
import io
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))

import ref_harness  # noqa: E402

ref = ref_harness.load()
import torch  # noqa: E402
from mst_native import abi, events as ev_mod  # noqa: E402

mido = ref.mido
MidiFormatError = sys.modules['style.exceptions'].MidiFormatError

KEY_NAMES = ref.scales.key_names


def sparse(a):
    a = np.ascontiguousarray(a, dtype=np.float32).ravel()
    idx = np.flatnonzero(a)
    return idx.astype(np.int64), a[idx]


def run_song(mid, min_n_messages, use_filter=True):
    """Returns a dict of golden arrays for one song."""
    out = {}
    events, tpb = ev_mod.from_mido(mid)
    out['events'] = events
    out['ticks_per_beat'] = np.int32(tpb)
    try:
        channels, info = ref.midi_conversion.read_midi(mid)
    except MidiFormatError as e:
        msg = e.msg
        status = {'Time signature changed': abi.SONG_TIME_SIG_CHANGED,
                  'Key signature changed': abi.SONG_KEY_SIG_CHANGED}.get(msg, abi.SONG_UNKNOWN_MESSAGE)
        out['status'] = np.int32(status)
        return out
    except ValueError:      # min()/max() of an empty sequence: no note_on at all
        out['status'] = np.int32(abi.SONG_NO_NOTES)
        return out
    out['lists'] = np.array(
        [(c['channel_id'], c['instrument_id'], len(c['messages'])) for c in channels], dtype=np.int32
    ).reshape(-1, 3)
    ts = info['time_signature']
    out['info'] = np.array([
        info['ticks_per_beat'], ts['numerator'], int(ts['denominator']).bit_length() - 1,
        info['duration'], info['ticks_per_bar'], int(np.ceil(info['n_bars'])), info['n_beats'],
        info['tempo'], info['bpm']], dtype=np.int64)
    if use_filter:
        channels = [c for c in channels
                    if c['instrument_id'] in [-1, *ref.data.included_instruments]
                    and len(c['messages']) >= min_n_messages]
    if not any(ref.midi.is_pitched(c['instrument_id']) for c in channels):
        out['status'] = np.int32(abi.SONG_NO_PITCHED)
        return out
    info, pitched, feats, instruments, unpitched = ref.data.get_input(channels, info)
    out['status'] = np.int32(abi.SONG_OK)
    scale = info['scale']
    out['scale'] = np.array([KEY_NAMES.index(scale['key']),
                             0 if scale['mode'] is ref.scales.major_mode else 1], dtype=np.int32)
    out['scale_stats'] = np.array([scale['coverage'], scale['loose_coverage'],
                                   scale['cross_entropy'], scale['loss']], dtype=np.float64)
    out['instruments'] = np.array(instruments, dtype=np.int32)
    out['instruments_features'] = feats.astype(np.float32)
    p32 = pitched.astype(np.float32)
    out['pitched_shape'] = np.array(p32.shape, dtype=np.int64)
    out['pitched_idx'], out['pitched_val'] = sparse(p32)
    out['pitched_sha256'] = np.frombuffer(hashlib.sha256(p32.tobytes()).digest(), dtype=np.uint8)
    if unpitched is not None:
        u32 = unpitched.astype(np.float32)
        out['unpitched_shape'] = np.array(u32.shape, dtype=np.int64)
        out['unpitched_idx'], out['unpitched_val'] = sparse(u32)
        out['unpitched_sha256'] = np.frombuffer(hashlib.sha256(u32.tobytes()).digest(), dtype=np.uint8)

    # decode: style_transfer.save (:77-98) -> decode_midi (:145-158) on the f32 grids
    cc = ref.midi_conversion.ChannelConverter(info)
    channels_info = [{'channel_id': i} for i in range(16) if i != 9][:p32.shape[0]]
    if len(channels_info) == p32.shape[0]:
        for instrument_id, channel_info in zip(instruments, channels_info):
            channel_info['instrument_id'] = instrument_id
        pt = torch.tensor(p32).unsqueeze(0)
        ut = torch.tensor(unpitched.astype(np.float32)).unsqueeze(0) if unpitched is not None else None
        midi_out = ref.style_transfer.decode_midi(
            cc, channels_info, pt, {'channel_id': 9, 'instrument_id': -1}, ut)
        track = midi_out.tracks[0]
        rows = []
        for m in track:
            if m.type in ('note_on', 'note_off'):
                rows.append((m.time, 1 if m.type == 'note_on' else 0, m.channel, m.note, m.velocity))
        out['decoded_notes'] = np.array(rows, dtype=np.int32).reshape(-1, 5)
        out['decoded_end_delta'] = np.int32(track[-1].time)
        with tempfile.NamedTemporaryFile(suffix='.mid') as f:
            midi_out.save(f.name)
            out['decoded_smf'] = np.frombuffer(open(f.name, 'rb').read(), dtype=np.uint8)
    return out


# ------------------------------------------------------------------ synthetic quirk songs
def _track(msgs):
    """msgs: list of (abs_tick, message) -> MidiTrack with delta times (stable order kept)."""
    tr = mido.MidiTrack()
    t = 0
    for tick, m in sorted(msgs, key=lambda x: x[0]):
        m = m.copy()
        m.time = tick - t
        t = tick
        tr.append(m)
    return tr


def synth_song(seed, tpb, numerator=4, n_bars=12, n_tracks=5, quirks=()):
    rng = np.random.RandomState(seed)
    mid = mido.MidiFile()
    mid.ticks_per_beat = tpb
    meta = [(0, mido.MetaMessage('time_signature', numerator=numerator, denominator=4)),
            (0, mido.MetaMessage('set_tempo', tempo=int(rng.choice([400000, 500000, 521739, 750000]))))]
    if 'tempo_changes' in quirks:
        for k in range(1, 6):
            meta.append((k * tpb * numerator * 2, mido.MetaMessage(
                'set_tempo', tempo=int(rng.choice([400000, 500000, 600000])))))
    if 'key_sig_before' in quirks:
        meta.append((0, mido.MetaMessage('key_signature', key='D')))
    if 'ts_change' in quirks:
        meta.append((tpb * numerator * 3, mido.MetaMessage('time_signature', numerator=3, denominator=4)))
    if 'key_change' in quirks:
        meta.append((0, mido.MetaMessage('key_signature', key='C')))
        meta.append((tpb * numerator * 3, mido.MetaMessage('key_signature', key='G')))
    if 'unknown_meta' in quirks:
        m = mido.MetaMessage('device_name')
        m.name = 'x'
        meta.append((0, m))
    if 'late_tempo' in quirks:
        meta.append((tpb * numerator * (n_bars + 3), mido.MetaMessage('set_tempo', tempo=300000)))
    mid.tracks.append(_track(meta))
    programs = [0, 25, 33, 48, 27, 30, 29, 1, 61, 73, 52, 4, 5, 24, 35]
    total = tpb * numerator * n_bars
    for t in range(n_tracks):
        ch = 9 if (t == n_tracks - 1 and 'no_drums' not in quirks) else (t if t < 9 else t + 1)
        msgs = []
        if ch != 9:
            prog = programs[t % len(programs)]
            if 'same_instrument' in quirks and t == 1:
                prog = programs[0]
            msgs.append((0, mido.Message('program_change', channel=ch, program=prog)))
            msgs.append((0, mido.Message('control_change', channel=ch, control=7,
                                         value=int(rng.randint(64, 128)))))
        if 'channel_prefix' in quirks and t == 0:
            m = mido.MetaMessage('channel_prefix')
            m.channel = 3
            msgs.append((0, m))
        n_notes = int(rng.randint(60, 90)) * max(1, n_bars // 12)
        for _ in range(n_notes):
            r = rng.rand()
            if r < 0.7:
                onset = int(rng.randint(0, total * 4 // tpb)) * tpb // 4
            elif r < 0.85:
                onset = int(rng.randint(0, total * 3 // tpb)) * tpb // 3
            else:
                onset = int(rng.randint(0, total))
            if rng.rand() < 0.2:
                onset = max(0, onset + int(rng.randint(-3, 4)))
            # a time/key signature that differs from the default is only accepted strictly before
            # the first note_on (is_during_song, midi_conversion.py:128-129,152-158)
            if numerator != 4 or 'key_sig_before' in quirks:
                onset += tpb
            dur = int(rng.geometric(1.0 / tpb))
            pitch = int(rng.randint(27, 88)) if ch == 9 else int(rng.randint(0, 128))
            vel = int(rng.randint(1, 128))
            msgs.append((onset, mido.Message('note_on', channel=ch, note=pitch, velocity=vel)))
            u = rng.rand()
            if u < 0.02:
                pass                                   # unclosed note
            elif u < 0.5:
                msgs.append((onset + dur, mido.Message('note_off', channel=ch, note=pitch, velocity=0)))
            else:
                msgs.append((onset + dur, mido.Message('note_on', channel=ch, note=pitch, velocity=0)))
        if 'program_change_mid' in quirks and ch != 9 and t == 0:
            msgs.append((total // 2, mido.Message('program_change', channel=ch, program=30)))
            msgs.append((total * 3 // 4, mido.Message('program_change', channel=ch, program=programs[0])))
        if 'volume_zero' in quirks and ch != 9 and t == 1:
            msgs.append((total // 3, mido.Message('control_change', channel=ch, control=7, value=0)))
            msgs.append((total // 2, mido.Message('control_change', channel=ch, control=7, value=100)))
        if 'other_messages' in quirks:
            pw = mido.Message('pitchwheel', channel=ch)
            pw.pitch = 100
            msgs.append((total + tpb, pw))
            msgs.append((tpb, mido.Message('aftertouch', channel=ch, value=5)))
        mid.tracks.append(_track(msgs))
    return mid


SYNTH_CASES = [
    dict(seed=1, tpb=480),
    dict(seed=2, tpb=100, quirks=('tempo_changes',)),
    dict(seed=3, tpb=96, numerator=3),
    dict(seed=4, tpb=1000, quirks=('program_change_mid', 'volume_zero')),
    dict(seed=5, tpb=1024, numerator=6, quirks=('same_instrument', 'other_messages')),
    dict(seed=6, tpb=120, quirks=('channel_prefix', 'key_sig_before', 'late_tempo')),
    dict(seed=7, tpb=192, n_tracks=12, n_bars=24),
    dict(seed=8, tpb=384, quirks=('no_drums',)),
    dict(seed=9, tpb=50, numerator=5, n_tracks=3),
    dict(seed=10, tpb=480, quirks=('ts_change',)),
    dict(seed=11, tpb=480, quirks=('key_change',)),
    dict(seed=12, tpb=480, quirks=('unknown_meta',)),
    dict(seed=13, tpb=7, numerator=2, n_tracks=4),
    dict(seed=14, tpb=25, numerator=7, n_tracks=6, quirks=('program_change_mid', 'same_instrument')),
]


def save_npz(path, songs):
    flat = {}
    for i, s in enumerate(songs):
        for k, v in s.items():
            flat[f'{i:03d}/{k}'] = v
    flat['n_songs'] = np.int32(len(songs))
    np.savez_compressed(path, **flat)
    print(path, os.path.getsize(path) // 1024, 'KiB')


def main():
    out_dir = os.path.join(ROOT, 'tests', 'golden')
    os.makedirs(out_dir, exist_ok=True)

    files = sorted(glob.glob(os.path.join(ref_harness.REFERENCE_ROOT, 'examples', '**', '*.mid'),
                             recursive=True))
    songs = []
    for f in files:
        mid = ref.midi.load_midi_from_file(f)
        s = run_song(mid, min_n_messages=100)
        rel = os.path.relpath(f, os.path.join(ref_harness.REFERENCE_ROOT, 'examples'))
        s['name'] = np.frombuffer(rel.encode('utf-8'), dtype=np.uint8)
        print(rel, int(s['status']), s.get('pitched_shape'))
        songs.append(s)
    save_npz(os.path.join(out_dir, 'codec_bundled.npz'), songs)

    songs = []
    for case in SYNTH_CASES:
        mid = synth_song(**case)
        s = run_song(mid, min_n_messages=0)
        s['name'] = np.frombuffer(repr(case).encode('utf-8'), dtype=np.uint8)
        print(case, int(s['status']), s.get('pitched_shape'))
        songs.append(s)
    save_npz(os.path.join(out_dir, 'codec_synth.npz'), songs)


if __name__ == '__main__':
    main()
