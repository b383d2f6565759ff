"""GPU parity tests of the device-side Standard MIDI File parser and the raw-bytes encode path
(csrc/mst_smf_gpu.cu: mst_smf_parse_batch_device, mst_encode_batch_smf_async) against the host parser mst_smf_parse
(csrc/mst_smf.cpp, itself pinned to the reference's loader by tests/test_native_cpu.py and the codec goldens):
bit-exact event records, identical "load failed" verdicts - on the bundled files, on files serialised from the golden
event lists, on hand-made hostile files and on a few thousand random mutations of valid files - and identical song infos /
bit-identical grids through the encoder."""
import hashlib
import os

import numpy as np
import pytest

import golden_util as gu

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def torch():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('GPU tests need a CUDA device (no CPU fallback exists)')
    return torch


def bundled_blobs():
    z = np.load(os.path.join(gu.GOLDEN_DIR, 'bundled_inputs.npz'), allow_pickle=False)
    return [bytes(z[f'file{i:02d}'].tobytes()) for i in range(len(z['names']))]


def golden_event_blobs():
    """Files serialised from the golden event lists (34 bundled + 14 synthetic quirk songs) by the native writer."""
    from mst_native import lib
    blobs = []
    for name in ('codec_bundled.npz', 'codec_synth.npz'):
        z = np.load(os.path.join(gu.GOLDEN_DIR, name), allow_pickle=False)
        n = int(z['n_songs'])
        for i in range(n):
            blobs.append(lib.smf_write_events(z[f'{i:03d}/events'], int(z[f'{i:03d}/ticks_per_beat'])))
    return blobs


def assert_same_parse(host, dev, tag=''):
    assert len(host) == len(dev)
    for i, (h, d) in enumerate(zip(host, dev)):
        assert (h is None) == (d is None), (tag, i, 'host failed' if h is None else 'host ok', 'device failed' if d is None else 'device ok')
        if h is None:
            continue
        assert h[1] == d[1], (tag, i, 'ticks_per_beat', h[1], d[1])
        assert len(h[0]) == len(d[0]), (tag, i, len(h[0]), len(d[0]))
        if not np.array_equal(h[0].view(np.uint8), d[0].view(np.uint8)):
            bad = np.nonzero(h[0] != d[0])[0][:3]
            raise AssertionError((tag, i, bad, h[0][bad], d[0][bad]))


def test_device_parser_matches_host_on_real_files(torch):
    from mst_native import lib
    blobs = bundled_blobs() + golden_event_blobs()
    host = [lib.smf_parse(b) for b in blobs]
    assert sum(h is not None for h in host) >= 60
    assert_same_parse(host, lib.smf_parse_device(blobs), 'real')


def hostile_files(base):
    """Hand-made failures of every LoadFail class of mst_smf.cpp, plus legal oddities."""
    trk = base.index(b'MTrk')
    out = {
        'empty': b'',
        'short': base[:7],
        'no MThd': b'XThd' + base[4:],
        'short header': base[:4] + (4).to_bytes(4, 'big') + base[8:12],
        'long header': base[:4] + (10).to_bytes(4, 'big') + base[8:14] + b'\x00\x00\x00\x00' + base[14:],
        'truncated': base[:len(base) // 2],
        'truncated by one': base[:-1],
        'no MTrk': base[:trk] + b'MTrx' + base[trk + 4:],
        'more tracks than chunks': base[:10] + (40).to_bytes(2, 'big') + base[12:],
        'fewer tracks than chunks': base[:10] + (1).to_bytes(2, 'big') + base[12:],
        'negative track count': base[:10] + (0xFFFF).to_bytes(2, 'big') + base[12:],
        'chunk size + 1': base[:trk + 4] + (int.from_bytes(base[trk + 4:trk + 8], 'big') + 1).to_bytes(4, 'big') + base[trk + 8:],
        'chunk size - 1': base[:trk + 4] + (int.from_bytes(base[trk + 4:trk + 8], 'big') - 1).to_bytes(4, 'big') + base[trk + 8:],
        'huge chunk size': base[:trk + 4] + b'\xff\xff\xff\xff' + base[trk + 8:],
        'trailing garbage': base + b'\x01\x02\x03garbage',
    }

    def one_track(body, division=480):
        return b'MThd' + (6).to_bytes(4, 'big') + (1).to_bytes(2, 'big') + (1).to_bytes(2, 'big') + division.to_bytes(2, 'big') + \
            b'MTrk' + len(body).to_bytes(4, 'big') + body
    eot = b'\x00\xff\x2f\x00'
    out.update({
        'running status without status': one_track(b'\x00\x40\x40' + eot),
        'data byte > 127': one_track(b'\x00\x90\x40\x90' + eot),
        'undefined status f4': one_track(b'\x00\xf4' + eot),
        'undefined status f5 after notes': one_track(b'\x00\x90\x40\x40\x10\xf5' + eot),
        'system common f2': one_track(b'\x00\xf2\x01\x02' + eot),
        'system realtime f8': one_track(b'\x00\xf8\x00\x90\x40\x40' + eot),
        'f1 with bad data': one_track(b'\x00\xf1\x80' + eot),
        'bad key signature': one_track(b'\x00\xff\x59\x02\x08\x00' + eot),
        'good key signature': one_track(b'\x00\xff\x59\x02\xf9\x01' + eot),
        'short tempo': one_track(b'\x00\xff\x51\x02\x07\xa1' + eot),
        'short time signature': one_track(b'\x00\xff\x58\x03\x04\x02\x18' + eot),
        'long time signature': one_track(b'\x00\xff\x58\x05\x03\x03\x18\x08\x00' + eot),
        'empty channel prefix': one_track(b'\x00\xff\x20\x00' + eot),
        'channel prefix 200': one_track(b'\x00\xff\x20\x01\xc8\x00\x90\x40\x40' + eot),
        'device name': one_track(b'\x00\xff\x09\x03abc' + eot),
        'meta payload beyond file': one_track(b'\x00\xff\x01\x7f' + eot),
        'sysex ok': one_track(b'\x00\xf0\x03\x01\x02\xf7\x00\x90\x40\x40' + eot),
        'sysex bad byte': one_track(b'\x00\xf0\x03\x01\x82\xf7' + eot),
        'sysex f7 escape': one_track(b'\x00\xf7\x02\x01\x02' + eot),
        'sysex sets running status': one_track(b'\x00\xf0\x01\xf7\x00\x40\x40' + eot),
        'huge delta': one_track(b'\xff\xff\xff\xff\xff\xff\xff\xff\xff\xff\x7f\x90\x40\x40' + eot),
        'tick beyond the limit': one_track(b'\x00\x90\x40\x40\x85\x80\x80\x00\x80\x40\x00\x00\x90\x41\x40' + eot),
        'varint to eof': one_track(b'\x00\x90\x40\x40\xff\xff'),
        'program change running': one_track(b'\x00\xc0\x05\x10\x06\x10\x07' + eot),
        'aftertouch and pitchwheel': one_track(b'\x00\xd3\x11\x00\xe3\x12\x13\x00\xa3\x14\x15' + eot),
        'smpte division': one_track(b'\x00\x90\x40\x40' + eot, division=0xE728),
        'no events': one_track(b''),
        'zero tracks': b'MThd' + (6).to_bytes(4, 'big') + b'\x00\x01\x00\x00\x01\xe0',
    })
    return out


def test_device_parser_matches_host_on_hostile_files(torch):
    from mst_native import lib
    base = bundled_blobs()[0]
    cases = hostile_files(base)
    names, blobs = list(cases), list(cases.values())
    host = [lib.smf_parse(b) for b in blobs]
    dev = lib.smf_parse_device(blobs)
    failed = [n for n, h in zip(names, host) if h is None]
    assert len(failed) >= 15 and len(failed) < len(names), failed       # both verdicts are exercised
    for i, n in enumerate(names):
        assert_same_parse([host[i]], [dev[i]], n)


def test_device_parser_matches_host_on_random_mutations(torch):
    """Byte flips, deletions and insertions in valid files: whatever the host parser makes of the result, the device parser
    makes the same of it (verdict and records)."""
    from mst_native import lib
    rng = np.random.default_rng(2024)
    sources = [b for b in bundled_blobs() + golden_event_blobs() if len(b) < 60000][:12]
    blobs = []
    for src in sources:
        a = np.frombuffer(src, dtype=np.uint8)
        for k in range(160):
            m = a.copy()
            kind = k % 4
            pos = int(rng.integers(0, len(m)))
            if kind == 0:
                m[pos] = rng.integers(0, 256)
            elif kind == 1:
                for p in rng.integers(14, len(m), size=3):
                    m[p] ^= 1 << int(rng.integers(0, 8))
            elif kind == 2:
                m = np.delete(m, slice(pos, pos + int(rng.integers(1, 4))))
            else:
                m = np.insert(m, pos, rng.integers(0, 256, size=int(rng.integers(1, 4)), dtype=np.uint8))
            blobs.append(m.tobytes())
    host = [lib.smf_parse(b) for b in blobs]
    n_ok = sum(h is not None for h in host)
    assert 50 < n_ok < len(blobs) - 50, n_ok                               # the mutations produce both verdicts
    assert_same_parse(host, lib.smf_parse_device(blobs), 'mutations')


def grid_digest(lib, info, arena):
    p, u = lib.song_grids(info, arena)
    h = hashlib.sha256(p.contiguous().cpu().numpy().tobytes())
    if u is not None:
        h.update(u.contiguous().cpu().numpy().tobytes())
    return h.hexdigest()


def test_raw_bytes_encode_matches_event_encode(torch):
    """mst_encode_batch_smf_async == mst_smf_parse + mst_encode_batch_device: every MstSongInfo field and every grid bit, with a
    rejected file and a file that does not load in the middle of the batch."""
    from mst_native import abi, lib
    blobs = bundled_blobs()[:8] + golden_event_blobs()[34:48]
    blobs.insert(3, blobs[0][:len(blobs[0]) // 3])                      # does not load
    blobs.insert(7, b'')                                                # unreadable / empty
    filt = abi.make_filter(range(128), 0, 0)
    parsed = [lib.smf_parse(b) for b in blobs]
    empty = np.zeros(0, dtype=abi.EVENT_DTYPE)
    infos_ref, arena_ref = lib.encode_songs([empty if p is None else p[0] for p in parsed], [0 if p is None else p[1] for p in parsed], filt)
    infos, arena = lib.encode_smf(blobs, filt)
    assert len(infos) == len(infos_ref) == len(blobs)
    n_ok = 0
    for i in range(len(blobs)):
        if parsed[i] is None:
            assert int(infos[i]['status']) == abi.SONG_LOAD_FAILED, (i, infos[i]['status'])
            continue
        for name in abi.SONG_INFO_DTYPE.names:
            if name in ('pitched_offset', 'unpitched_offset'):
                continue
            assert np.array_equal(infos[i][name], infos_ref[i][name]), (i, name, infos[i][name], infos_ref[i][name])
        if int(infos[i]['status']) == abi.SONG_OK:
            assert grid_digest(lib, infos[i], arena) == grid_digest(lib, infos_ref[i], arena_ref), i
            n_ok += 1
    assert n_ok >= 15


def test_event_pipeline_wire_formats_agree(torch, tmp_path):
    """The product pipeline (mst_native.pipeline.EventPipeline, behind style.data.iter_inputs) yields the same song infos and
    bit-identical grids whether the files are parsed on the device (wire='smf', the default) or on the host (wire='events');
    an unreadable path, an empty file and a truncated file are reported as load failures by both."""
    from mst_native import abi, lib
    from mst_native.pipeline import EventPipeline
    blobs = bundled_blobs()[:6] + golden_event_blobs()[34:44]
    blobs.insert(2, b'')
    blobs.insert(5, blobs[0][:1000])
    paths = []
    for i, b in enumerate(blobs):
        p = tmp_path / f'song{i:02d}.mid'
        p.write_bytes(b)
        paths.append(str(p))
    paths.insert(9, str(tmp_path / 'missing.mid'))
    filt = abi.make_filter(range(128), 0, 0)
    seen = {}
    for wire in ('smf', 'events'):
        pipe = EventPipeline(filt, batch_files=5, slots=3, arena_floats=64 << 20, wire=wire)
        rows = []
        for path, r, arena in pipe.run(paths):
            rec = {'path': path, 'status': int(r['status'])}
            if rec['status'] == abi.SONG_OK:
                rec['info'] = {k: np.array(r[k]).copy() for k in abi.SONG_INFO_DTYPE.names if k not in ('pitched_offset', 'unpitched_offset')}
                rec['digest'] = grid_digest(lib, r, arena)
            rows.append(rec)
        pipe.close()
        seen[wire] = rows
        assert [r['path'] for r in rows] == paths
    failed = [r['path'] for r in seen['smf'] if r['status'] == abi.SONG_LOAD_FAILED]
    assert sorted(os.path.basename(p) for p in failed) == ['missing.mid', 'song02.mid', 'song05.mid']
    n_ok = 0
    for a, b in zip(seen['smf'], seen['events']):
        assert a['status'] == b['status'], (a['path'], a['status'], b['status'])
        if a['status'] == abi.SONG_OK:
            for k in a['info']:
                assert np.array_equal(a['info'][k], b['info'][k]), (a['path'], k)
            assert a['digest'] == b['digest'], a['path']
            n_ok += 1
    assert n_ok >= 10
