"""CPU tests of the C-ABI library: it loads without a GPU, exports every symbol the header declares,
its struct layouts match the numpy mirrors, and the host-side SMF parser follows mido's rules."""
import ctypes
import os
import re
import sys

import numpy as np
import pytest

import golden_util as gu
from mst_native import abi, events as ev_mod, lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def L():
    sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))
    import build
    build.build()
    return lib.load()


def test_exports_every_declared_symbol(L):
    header = open(os.path.join(ROOT, 'include', 'mst_b200.h')).read()
    declared = set(re.findall(r'\b(mst_[a-z0-9_]+)\s*\(', header))
    assert declared, 'no declarations found'
    for name in declared:
        assert hasattr(L, name), f'{name} declared in include/mst_b200.h but not exported'
    assert declared == set(lib.EXPORTS)
    assert L.mst_abi_version() == abi.MST_ABI_VERSION


def test_no_cpu_fallback_without_gpu(L):
    if L.mst_device_count() > 0:
        pytest.skip('GPU present')
    with pytest.raises(lib.MstError):
        lib.require_gpu()


def test_smf_parse_of_reference_written_files(L):
    """The golden 'decoded_smf' blobs are SMF files the reference wrote through mid.save()
    (running status, one track): parse them and compare with the message list that produced them."""
    n_checked = 0
    for name in ('codec_bundled', 'codec_synth'):
        for s in gu.load_songs(name):
            if 'decoded_smf' not in s:
                continue
            res = lib.smf_parse(bytes(s['decoded_smf']))
            assert res is not None
            events, tpb = res
            assert tpb == int(s['info'][0])
            notes = events[(events['kind'] == abi.EV_NOTE_ON) | (events['kind'] == abi.EV_NOTE_OFF)]
            g = s['decoded_notes']
            assert len(notes) == len(g)
            np.testing.assert_array_equal(notes['tick'], np.cumsum(g[:, 0]))
            np.testing.assert_array_equal(notes['kind'], g[:, 1])
            np.testing.assert_array_equal(notes['channel'], g[:, 2])
            np.testing.assert_array_equal(notes['d1'], g[:, 3])
            np.testing.assert_array_equal(notes['d2'], g[:, 4])
            tempo = events[events['kind'] == abi.EV_SET_TEMPO]
            assert len(tempo) == 1 and int(tempo['aux'][0]) == int(s['info'][7])
            n_checked += 1
    assert n_checked >= 40


def _shim_mido():
    p = os.path.join(ROOT, 'oracle', 'refshim')
    if p not in sys.path:
        sys.path.insert(0, p)
    import mido
    return mido


def test_smf_parse_matches_mido_conversion(L, tmp_path):
    """Multi-track file with metas, program changes, pitchwheel, channel_prefix, device_name:
    C++ parser == mido-object conversion (mst_native.events.from_mido) of the same file."""
    mido = _shim_mido()
    mid = mido.MidiFile()
    mid.ticks_per_beat = 96
    t0 = mido.MidiTrack()
    t0.append(mido.MetaMessage('time_signature', numerator=3, denominator=8, time=0))
    t0.append(mido.MetaMessage('set_tempo', tempo=432100, time=0))
    t0.append(mido.MetaMessage('key_signature', key='F#m', time=10))
    t0.append(mido.MetaMessage('end_of_track', time=5))
    t1 = mido.MidiTrack()
    t1.append(mido.Message('program_change', channel=2, program=33, time=0))
    t1.append(mido.Message('control_change', channel=2, control=7, value=101, time=3))
    for k in range(20):
        t1.append(mido.Message('note_on', channel=2, note=40 + k, velocity=64 + k, time=7))
        t1.append(mido.Message('note_on', channel=2, note=40 + k, velocity=0, time=11))
    pw = mido.Message('pitchwheel', channel=2, time=1)
    pw.pitch = -200
    t1.append(pw)
    t1.append(mido.MetaMessage('end_of_track', time=0))
    mid.tracks += [t0, t1]
    path = str(tmp_path / 'a.mid')
    mid.save(path)
    back = mido.MidiFile(path)
    ref_events, ref_tpb = ev_mod.from_mido(back)
    events, tpb = lib.smf_parse(open(path, 'rb').read())
    assert tpb == ref_tpb == 96
    np.testing.assert_array_equal(events, ref_events)


def _smf(tracks, division=480, fmt=1):
    out = b'MThd' + (6).to_bytes(4, 'big') + fmt.to_bytes(2, 'big') + len(tracks).to_bytes(2, 'big') + \
        division.to_bytes(2, 'big')
    for t in tracks:
        out += b'MTrk' + len(t).to_bytes(4, 'big') + t
    return out


def test_smf_parse_edge_cases(L):
    eot = bytes([0, 0xFF, 0x2F, 0])
    # running status + device_name (unknown) + channel_prefix + a realtime byte (unknown message)
    tr = bytes([0, 0x90, 60, 100, 10, 61, 90, 0, 0xFF, 0x09, 1, 65, 0, 0xFF, 0x20, 1, 5, 0, 0xF8,
                0x81, 0x00, 0x80, 60, 0]) + eot
    events, tpb = lib.smf_parse(_smf([tr]))
    assert list(events['kind']) == [abi.EV_NOTE_ON, abi.EV_NOTE_ON, abi.EV_UNKNOWN, abi.EV_CHANNEL_PREFIX,
                                    abi.EV_UNKNOWN, abi.EV_NOTE_OFF]
    assert list(events['tick']) == [0, 10, 10, 10, 10, 138]
    assert int(events['channel'][3]) == 5
    # load failures on which load_midi_from_file returns None (style/midi.py:104-108)
    assert lib.smf_parse(b'RIFF' + bytes(20)) is None                       # no MThd
    assert lib.smf_parse(_smf([tr])[:-3]) is None                           # truncated
    assert lib.smf_parse(_smf([bytes([0, 60, 100]) + eot])) is None         # running status without status
    assert lib.smf_parse(_smf([bytes([0, 0x90, 60, 200]) + eot])) is None   # data byte > 127
    assert lib.smf_parse(_smf([bytes([0, 0xF4]) + eot])) is None            # undefined status byte
    assert lib.smf_parse(_smf([bytes([0, 0xFF, 0x59, 2, 9, 0]) + eot])) is None   # key signature error
    # events after tick 1e7 are dropped (midi_conversion.py:52,59)
    big = bytes([0x85, 0x80, 0x80, 0x00])   # delta 10485760 > 1e7
    events, _ = lib.smf_parse(_smf([bytes([0, 0x90, 60, 100]) + big + bytes([0x80, 60, 0]) + eot]))
    assert len(events) == 1


def test_smf_write_matches_reference_saved_bytes(L):
    """mst_smf_write of the oracle-decoded messages == the bytes the reference's mid.save() produced."""
    import oracle_lib
    n = 0
    for name in ('codec_bundled', 'codec_synth'):
        for s in gu.load_songs(name):
            if 'decoded_smf' not in s:
                continue
            grids, d = gu.decode_inputs(s)
            res, msgs = oracle_lib.decode_song(grids, d)
            assert lib.smf_write(d, msgs, int(res['eot_delta'])) == bytes(s['decoded_smf']), s['name']
            n += 1
    assert n >= 40


def test_smf_parse_hostile_lengths_terminate(L):
    """A meta / sysex / delta varint of any width must end in the EOF load failure mido reports (EOFError ->
    load_midi_from_file returns None), never in a wrapped position that re-reads the same event for ever."""
    head = b'MThd' + (6).to_bytes(4, 'big') + b'\x00\x01\x00\x01\x01\xe0' + b'MTrk' + (0x7FFFFFFF).to_bytes(4, 'big')
    huge = b'\xff' * 9 + b'\x73'                               # 2^64 - 13 as a 10-group varint
    cases = [
        head + b'\x00\xff\x01' + huge + b'abc',                # meta text with a wrapping length
        head + b'\x00\xf0' + huge + b'\x01\x02\xf7',           # sysex with a wrapping length
        head + huge + b'\x90\x40\x40',                         # delta time wider than 64 bits
        head + b'\x00\xff\x51' + b'\xff\xff\xff\xff\x7f' + b'\x07\xa1\x20',   # 5-byte length beyond the file
    ]
    for data in cases:
        assert lib.smf_parse(data) is None


def test_files_read_batch_reads_files_as_they_are(tmp_path):
    """mst_files_read_batch: raw bytes on 16-byte boundaries, unreadable files as 0 bytes + status -1, capacity error with
    the needed size."""
    import numpy as np
    from mst_native import lib
    blobs = [bytes(range(200)) * 3, b'MThd', b'', bytes(17)]
    paths = []
    for i, b in enumerate(blobs):
        p = tmp_path / f'f{i}.bin'
        p.write_bytes(b)
        paths.append(str(p))
    paths.insert(2, str(tmp_path / 'missing.mid'))
    blobs.insert(2, None)
    buf = np.full(4096, 0xEE, dtype=np.uint8)
    total, files, status = lib.files_read_batch(paths, buf, n_threads=3)
    assert list(status) == [0, 0, -1, 0, 0]
    end = 0
    for i, b in enumerate(blobs):
        o, n = int(files[i]['byte_offset']), int(files[i]['n_bytes'])
        assert o % 16 == 0 and o >= end
        if b is None:
            assert n == 0
            continue
        assert n == len(b) and buf[o:o + n].tobytes() == b
        end = o + n
    assert total >= end and total % 16 == 0
    with pytest.raises(lib.MstError) as e:
        lib.files_read_batch(paths, np.zeros(64, dtype=np.uint8))
    assert e.value.needed == total


@pytest.mark.timeout(120)
def test_model_workspace_plan_runs_without_a_gpu():
    """mst_model_workspace_bytes walks the whole forward + backward host code in planning mode (no launches): it must
    terminate and grow with the batch - this is the cheapest guard against a broken schedule in the host code."""
    import ctypes
    from mst_native import model as mm
    L = mm._L()
    sizes = {}
    for n, c, b, u in ((1, 4, 200, 1), (2, 4, 200, 1), (2, 3, 6, 0), (1, 8, 800, 1)):
        for training in (0, 1):
            s = mm.ModelShape(n, c, b, 4, u, 0)
            sizes[(n, c, b, u, training)] = int(L.mst_model_workspace_bytes(ctypes.byref(s), training))
            assert sizes[(n, c, b, u, training)] > 0
    assert sizes[(2, 4, 200, 1, 1)] > sizes[(1, 4, 200, 1, 1)] > sizes[(1, 4, 200, 1, 0)]


def test_raw_bytes_entry_points_fail_loudly_without_gpu(L):
    """The device-side parser path has no CPU fallback either: without a GPU both entry points return MST_E_NO_DEVICE, and
    their workspace queries (pure host arithmetic) grow with the byte count."""
    import ctypes
    import numpy as np
    from mst_native import abi, lib
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    buf, files = lib.concat_files([b'MThd' + bytes(10), b''])
    infos = np.zeros(2, dtype=abi.SONG_INFO_DTYPE)
    filt = abi.make_filter(None, 0, 0).reshape(1)
    dummy = np.zeros(64, dtype=np.uint8)
    rc = L.mst_encode_batch_smf_async(buf.ctypes.data, len(buf), files.ctypes.data, 2, filt.ctypes.data, infos.ctypes.data,
                                      None, 0, dummy.ctypes.data, dummy.size, None)
    assert rc == -4 and b'no CPU fallback' in L.mst_last_error()
    events = np.zeros(16, dtype=abi.EVENT_DTYPE)
    songs = np.zeros(2, dtype=abi.SONG_IN_DTYPE)
    status = np.zeros(2, dtype=np.int32)
    total = ctypes.c_uint64(0)
    rc = L.mst_smf_parse_batch_device(buf.ctypes.data, len(buf), files.ctypes.data, 2, events.ctypes.data, 16, songs.ctypes.data,
                                      status.ctypes.data, ctypes.byref(total), dummy.ctypes.data, dummy.size, None)
    assert rc == -4
    a = L.mst_encode_smf_workspace_bytes(1 << 20, 8, 1 << 20)
    b = L.mst_encode_smf_workspace_bytes(1 << 24, 8, 1 << 20)
    assert 0 < a < b
    assert 0 < L.mst_smf_parse_device_workspace_bytes(1 << 20, 8) < L.mst_smf_parse_device_workspace_bytes(1 << 24, 8)
