"""CPU tests of the drop-in `style` package (music-style-transfer_b200/style): the reference's public names exist with
the reference's signatures, the module constructors reproduce the seed-108 weights bit for bit, and the host-side
helpers (one-hot encoders, info dicts, decode records, SMF loading) agree with the reference goldens.  No compute
call is made here; the GPU legs are in tests/test_gpu_dropin.py."""
import hashlib
import inspect
import os

import numpy as np
import pytest

import golden_util as gu

Z = np.load(os.path.join(gu.GOLDEN_DIR, 'model_small.npz'))

# the reference's public surface on the path (file: names); signatures are checked against these parameter lists
SURFACE = {
    'style.data': ['included_instruments', 'instrument_groups', 'n_instruments', 'instruments_one_hot_encoder',
                   'groups_one_hot_encoder', 'instrument_size', 'percussion_id', 'iter_all_midis', 'iter_inputs',
                   'get_input', 'encode_instruments', 'prepare_input', 'get_used_instruments'],
    'style.model': ['device', 'LSTM', 'Distributed', 'PitchedChannelsEncoder', 'UnpitchedChannelsEncoder',
                    'PitchedRhythmEncoder', 'UnpitchedRhythmEncoder', 'StyleEncoder', 'MelodyEncoder', 'SongInfoModel',
                    'PitchedStyleApplier', 'UnpitchedStyleApplier', 'StyleTransferModel', 'get_total_loss',
                    'hard_output', 'combine'],
    'style.midi_conversion': ['read_midi', 'ChannelConverter', 'get_keys_dist'],
    'style.midi': ['load_midi_from_file', 'create_midi', 'popular_instruments', 'program2instrument', 'program2group',
                   'is_pitched'],
    'style.style_transfer': ['transfer_style', 'get_model_input', 'extract_style', 'save', 'apply_style',
                             'combine_info', 'decode_midi'],
    'style.scales': ['key_names', 'major_mode', 'minor_mode', 'create_mode'],
    'style.exceptions': ['MidiFormatError'],
    'style.utils.misc': ['iter_all_files', 'make_dirs', 'dict_map', 'ProgressBar'],
    'style.utils.data': ['save_to_csv', 'assert_dir'],
    'style.utils.parallel': ['iter_parallel'],
    'style.utils.pytorch': ['LSTM', 'Distributed', 'squash_dims', 'cat_with_broadcast', 'safe_sqrt', 'get_mean'],
}

SIGNATURES = {
    ('style.data', 'iter_inputs'): ['files', 'instruments', 'min_n_messages', 'shuffle', 'looped'],
    ('style.data', 'get_input'): ['channels', 'info'],
    ('style.data', 'prepare_input'): ['input', 'max_n_bars'],
    ('style.data', 'get_used_instruments'): ['instruments_features', 'unpitched_channels'],
    ('style.model', 'get_total_loss'): ['instruments_pred', 'instruments_target', 'mode_pred', 'mode_target', 'bpm_pred',
                                        'bpm_target', 'pitched_pred', 'pitched_target', 'unpitched_pred',
                                        'unpitched_target', 'normalize'],
    ('style.style_transfer', 'transfer_style'): ['model', 'composition_path', 'style_paths', 'output_path'],
    ('style.style_transfer', 'apply_style'): ['model', 'info', 'style', 'melody', 'rhythm', 'n_instruments', 'save_path'],
    ('style.style_transfer', 'decode_midi'): ['channel_converter', 'channels_info', 'pitched_channels',
                                              'unpitched_channel_info', 'unpitched_channels'],
}


def test_public_surface_and_signatures():
    import importlib
    for mod, names in SURFACE.items():
        m = importlib.import_module(mod)
        assert 'music-style-transfer_b200' in m.__file__, m.__file__
        for n in names:
            assert hasattr(m, n), f'{mod}.{n}'
    for (mod, fn), params in SIGNATURES.items():
        got = list(inspect.signature(getattr(importlib.import_module(mod), fn)).parameters)
        assert got[:len(params)] == params, (mod, fn, got)


def build_model(seed=108):
    import torch
    from style import model as m, data as d
    torch.manual_seed(seed)
    pce = m.PitchedChannelsEncoder(64, 128, d.instrument_size)
    uce = m.UnpitchedChannelsEncoder(64, 128)
    pre = m.PitchedRhythmEncoder(32, 64, 128, d.instrument_size)
    ure = m.UnpitchedRhythmEncoder(32, 64, 128)
    se = m.StyleEncoder(256, 128, d.instrument_size)
    me = m.MelodyEncoder(8, 64, 128, d.instrument_size)
    sim = m.SongInfoModel(8, 256, 32, d.n_instruments)
    psa = m.PitchedStyleApplier(256, 8, 32, d.instrument_size)
    usa = m.UnpitchedStyleApplier(256, 32)
    return m.StyleTransferModel(pce, uce, se, me, pre, ure, sim, psa, usa)


def test_constructors_reproduce_the_reference_weights():
    """train-model.py:52-85 order + torch.manual_seed(108) -> the exact tensors of the reference (golden sha256)."""
    model = build_model()
    sd = model.state_dict()
    names = [str(n) for n in Z['param_names']]
    assert sorted(sd) == sorted(names)
    h = hashlib.sha256()
    for n in sorted(sd):
        h.update(n.encode()); h.update(sd[n].detach().cpu().numpy().tobytes())
    assert h.digest() == bytes(Z['weights_sha256'])
    assert sum(p.numel() for p in model.parameters()) == 980325
    # every parameter has a slot of the right size in the library's flat table (what mst_model_* reads)
    from mst_native import model as mm
    table = {n: numel for n, _, numel in mm.param_table()}
    assert {n: p.numel() for n, p in model.named_parameters()} == table


def test_instrument_encoders_match_reference_goldens():
    from style import data as d
    assert (d.instrument_size, d.n_instruments, d.percussion_id) == (51, 41, 40)
    n = 0
    for name in ('codec_bundled', 'codec_synth'):
        for s in gu.load_songs(name):
            if 'instruments_features' not in s:
                continue
            got = d.encode_instruments([int(x) for x in s['instruments']])
            assert np.array_equal(got, s['instruments_features']), s['name']
            n += 1
    assert n > 30
    x = np.zeros((2, 40)); x[0, 3] = 1; x[1, 39] = 1
    inv = d.instruments_one_hot_encoder.inverse_transform(x)
    assert [int(v[0]) for v in inv] == [sorted(d.included_instruments)[3], sorted(d.included_instruments)[39]]


def test_used_instruments_and_mode_bpm_tensors():
    import torch
    from style import data as d
    feats = torch.tensor(d.encode_instruments([0, 25, 25]), dtype=torch.float).unsqueeze(0)
    used = d.get_used_instruments(feats, None)
    assert used.shape == (1, 41) and float(used.sum()) == 2.0 and float(used[0, 40]) == 0.0
    used = d.get_used_instruments(feats, torch.zeros(1))
    assert float(used[0, 40]) == 1.0


def test_load_midi_and_decode_record_match_goldens(tmp_path):
    from mst_native import abi
    from style import midi, midi_conversion as mc, scales
    s = gu.load_songs('codec_synth')[0]
    p = tmp_path / 'a.mid'
    p.write_bytes(bytes(s['decoded_smf']))
    mid = midi.load_midi_from_file(str(p))
    assert mid is not None and mid.ticks_per_beat == int(s['info'][0]) and len(mid.events) > 0
    assert midi.load_midi_from_file(str(tmp_path / 'missing.mid')) is None
    (tmp_path / 'bad.mid').write_bytes(b'not a midi file')
    assert midi.load_midi_from_file(str(tmp_path / 'bad.mid')) is None
    # ChannelConverter(info).decode_song_record == the record the decode goldens were produced with
    grids, want = gu.decode_inputs(s)
    g = s['info']
    info = {'ticks_per_beat': int(g[0]), 'ticks_per_bar': int(g[4]), 'tempo': int(g[7]), 'duration': int(g[3]),
            'time_signature': {'numerator': int(g[1]), 'denominator': 1 << int(g[2])},
            'scale': {'key': scales.key_names[int(s['scale'][0])], 'mode': scales.MODES[int(s['scale'][1])]}}
    C, B = int(s['pitched_shape'][0]), int(s['pitched_shape'][1])
    got = mc.ChannelConverter(info).decode_song_record(C, 'unpitched_shape' in s, B, int(g[6]),
                                                       [i for i in range(16) if i != 9][:C], list(s['instruments']))
    got['unpitched_offset'] = want['unpitched_offset']
    assert got.tobytes() == want.tobytes()
    with pytest.raises(NotImplementedError):
        mc.ChannelConverter(info, beat_divisors=(4,))


def test_combine_info_and_tempo():
    from style import style_transfer as st
    a = {'time_signature': 1, 'scale': 2, 'ticks_per_beat': 3, 'ticks_per_bar': 4, 'tempo': 5}
    b = {'time_signature': 6, 'scale': 7, 'ticks_per_beat': 8, 'ticks_per_bar': 9, 'tempo': 10}
    assert st.combine_info(style_info=a, melody_info=b) == {'time_signature': 6, 'scale': 2, 'ticks_per_beat': 8,
                                                            'ticks_per_bar': 9, 'tempo': 5}
    assert st.bpm2tempo(120) == 500000 and st.bpm2tempo(97) == 618557


def test_model_needs_the_gpu_library():
    """No CPU fallback: without a CUDA device the model's compute entry points raise."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from mst_native.lib import MstError
    model = build_model()
    with pytest.raises((MstError, RuntimeError, AssertionError)):
        model(torch.tensor([[1., 0.]]), torch.tensor([120.]), torch.zeros(1, 1, 1, 4, 10, 56, 5), torch.zeros(1, 1, 51), None)


def test_utils_pytorch_helpers_and_pickle_names():
    """style.utils.pytorch exists because reference torch.save(model) pickles name LSTM / Distributed by that path."""
    import torch
    from style.utils import pytorch as up
    from style import model as m
    assert up.LSTM is m.LSTM and up.Distributed is m.Distributed
    x = torch.arange(2 * 3 * 4 * 5, dtype=torch.float32).view(2, 3, 4, 5)
    assert up.squash_dims(x, 1, 3).shape == (2, 12, 5) and up.squash_dims(x, -2).shape == (2, 3, 20)
    a, b = torch.ones(2, 1, 3), torch.zeros(1, 4, 5)
    assert up.cat_with_broadcast([a, b], -1).shape == (2, 4, 8)
    t = [torch.tensor(3.0), torch.tensor(4.0)]
    assert abs(float(up.get_mean(t, mean_type='quadratic')) - (12.5 ** .5)) < 1e-6
    assert float(up.get_mean([torch.tensor(0.0)], mean_type='quadratic')) == 0.0


def test_create_midi_matches_reference_bytes():
    """style.midi.create_midi on Python message lists (host logic + native SMF writer) against the bytes the reference's
    create_midi + mid.save() produced for the same lists (tests/golden/create_midi.npz, oracle/gen_golden_configs.py)."""
    from style.midi import create_midi
    G = np.load(os.path.join(gu.GOLDEN_DIR, 'create_midi.npz'))

    class Msg:
        def __init__(self, on, note, time, velocity):
            self.type, self.note, self.time, self.velocity = ('note_on' if on else 'note_off'), int(note), int(time), velocity

    info = {'ticks_per_beat': 480, 'tempo': 434783, 'time_signature': {'numerator': 3, 'denominator': 4}, 'ticks_per_bar': 1440}
    for ci in range(int(G['n_cases'])):
        i2 = dict(info)
        if int(G[f'{ci}/duration']) >= 0:
            i2['duration'] = int(G[f'{ci}/duration'])
        chans = []
        for k in range(3):
            ch, prog = (int(x) for x in G[f'{ci}/{k}/channel'])
            msgs = [Msg(on, note, t, np.float32(v) if on else 0) for (on, note, t), v in zip(G[f'{ci}/{k}/msgs'], G[f'{ci}/{k}/velocity'])]
            chans.append({'channel_id': ch, 'instrument_id': prog, 'messages': msgs})
        got = create_midi(i2, *chans, max_delta_time=float(G[f'{ci}/max_delta_time']))
        assert got.data == bytes(G[f'{ci}/bytes']), ci


@pytest.mark.refcheck
def test_train_script_is_the_reference_script_modulo_documented_lines():
    """music-style-transfer_b200/train-model.py claims to be the reference's script (same hyper-parameters, same loop): the
    only differences allowed are the docstring, the flatten_dict import / call (third-party package not needed) and the two
    environment overrides of data_path / n_iterations.  Needs /root/reference (build container only)."""
    import difflib
    ref_path = '/root/reference/train-model.py'
    if not os.path.exists(ref_path):
        pytest.skip('/root/reference is not present on this machine')
    here = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'music-style-transfer_b200', 'train-model.py')
    ref = [l.rstrip() for l in open(ref_path).read().splitlines()]
    ours = [l.rstrip() for l in open(here).read().splitlines()]
    removed, added = [], []
    for l in difflib.ndiff(ref, ours):
        if l.startswith('- '):
            removed.append(l[2:].strip())
        elif l.startswith('+ '):
            added.append(l[2:].strip())
    assert removed == [
        'from flatten_dict import flatten as flatten_dict',
        'from style.utils.data import save_to_csv, assert_dir',
        "data_path = 'data/Lakh MIDI Dataset/clean_midi/'",
        'n_iterations = 5000',
        "losses = flatten_dict(losses, reducer='underscore')",
    ], removed
    code_added = [l for l in added if l and not l.startswith('"""') and 'drop-in package' not in l and 'third-party' not in l]
    assert code_added == [
        'from style.utils.data import save_to_csv, assert_dir, flatten_dict',
        "data_path = os.environ.get('MST_DATA_PATH', 'data/Lakh MIDI Dataset/clean_midi/')",
        "n_iterations = int(os.environ.get('MST_N_ITERATIONS', 5000))",
        'losses = flatten_dict(losses)',
    ], code_added
