"""GPU parity tests of the encode path, through the C ABI (mst_encode_batch_device / _host):
  * golden vectors produced by the unmodified reference (bit-exact grids, identical info/scale),
  * the C oracle on seeded synthetic corpora, including ticks_per_beat not divisible by 24,
  * size-independent properties at BASELINE's corpus shape (idempotence, batch-composition
    independence, arena/crop behaviour)."""
import ctypes
import hashlib

import numpy as np
import pytest

import golden_util as gu
from mst_native import abi, lib, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def torch():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('GPU tests need a CUDA device (no CPU fallback exists)')
    lib.require_gpu()
    return torch


def _check_song_against_golden(s, info, arena):
    assert int(info['status']) == int(s['status']), s['name']
    if 'info' in s:
        got = [info[k] for k in ('ticks_per_beat', 'numerator', 'denominator_log2', 'duration', 'ticks_per_bar',
                                 'n_bars', 'n_beats', 'tempo', 'bpm')]
        assert [int(x) for x in got] == [int(x) for x in s['info']], s['name']
        assert int(info['n_lists']) == len(s['lists'])
    if int(s['status']) != abi.SONG_OK:
        return
    assert (int(info['tonic']), int(info['mode'])) == tuple(int(x) for x in s['scale']), s['name']
    np.testing.assert_allclose(
        [info['scale_coverage'], info['scale_loose_coverage'], info['scale_cross_entropy'], info['scale_loss']],
        s['scale_stats'], rtol=1e-12)
    C = int(info['n_pitched'])
    np.testing.assert_array_equal(info['pitched_instruments'][:C], s['instruments'])
    pitched, unpitched = lib.song_grids(info, arena)
    assert list(pitched.shape) == [int(x) for x in s['pitched_shape']]
    p = pitched.cpu().numpy()
    assert hashlib.sha256(p.tobytes()).digest() == bytes(s['pitched_sha256']), s['name']
    if 'unpitched_shape' in s:
        u = unpitched.cpu().numpy()
        assert list(u.shape) == [int(x) for x in s['unpitched_shape']]
        assert hashlib.sha256(u.tobytes()).digest() == bytes(s['unpitched_sha256']), s['name']
    else:
        assert unpitched is None


@pytest.mark.parametrize('name', ['codec_bundled', 'codec_synth'])
def test_encode_matches_reference_goldens(torch, name):
    songs = gu.load_songs(name)
    filt = abi.make_filter(gu.INCLUDED, gu.min_n_messages(name))
    infos, arena = lib.encode_songs([s['events'] for s in songs], [int(s['ticks_per_beat']) for s in songs], filt)
    for s, info in zip(songs, infos):
        _check_song_against_golden(s, info, arena)


def _oracle_compare(torch, events, songs, filt):
    import oracle_lib
    per_song = synth.split_songs(events, songs)
    infos, arena = lib.encode_songs(per_song, songs['ticks_per_beat'], filt)
    for e, sg, info in zip(per_song, songs, infos):
        oinfo, op, ou = oracle_lib.encode_song(e, sg['ticks_per_beat'], filt)
        for k in ('status', 'duration', 'ticks_per_bar', 'n_bars', 'n_beats', 'tempo', 'bpm', 'tonic', 'mode',
                  'n_pitched', 'n_unpitched', 'n_bars_out', 'n_notes', 'n_lists'):
            assert int(info[k]) == int(oinfo[k]), k
        if int(oinfo['status']) != abi.SONG_OK:
            continue
        np.testing.assert_array_equal(info['pitched_instruments'], oinfo['pitched_instruments'])
        np.testing.assert_array_equal(info['keys_dist'], oinfo['keys_dist'])   # same summation order: bit-equal
        p, u = lib.song_grids(info, arena)
        assert np.array_equal(p.cpu().numpy(), op)
        if ou is not None:
            assert np.array_equal(u.cpu().numpy(), ou)
    return infos, arena


def test_encode_matches_oracle_on_synthetic_corpus(torch):
    events, songs = synth.synth_corpus(6, seed=1234)
    _oracle_compare(torch, events, songs, abi.make_filter(synth.POPULAR, 100))


@pytest.mark.parametrize('tpb', [96, 100, 120, 192, 384, 1000, 1024])
def test_encode_quantisation_quirk_tpb(torch, tpb):
    """tpb not divisible by 24 lands notes one slot early (f64 fmod / trunc / floor-div sequence)."""
    events, songs = synth.synth_corpus(2, seed=tpb, tpb=tpb, n_bars=24, n_tracks=6)
    _oracle_compare(torch, events, songs, abi.make_filter(synth.POPULAR, 0))


def test_encode_tracks_sharing_channels_take_the_full_sort(torch):
    """k_sort's fast path (class partition + sortedness check) only holds when no two tracks share a class: fold the
    tracks of some songs onto three channels so that their class segments interleave in tick order and the full radix
    sort has to run; the other songs of the same batch stay on the fast path."""
    events, songs = synth.synth_corpus(4, seed=77, n_bars=24, n_tracks=8)
    events = events.copy()
    for sg in songs[::2]:
        lo, hi = int(sg['event_offset']), int(sg['event_offset']) + int(sg['n_events'])
        ev = events[lo:hi]
        pitched = (ev['track'] > 0) & (ev['channel'] != 9)
        ev['channel'][pitched] = ev['channel'][pitched] % 3
    infos, _ = _oracle_compare(torch, events, songs, abi.make_filter(synth.POPULAR, 0))
    assert all(int(r['status']) == abi.SONG_OK for r in infos)


def test_encode_without_filter_and_with_crop(torch):
    events, songs = synth.synth_corpus(2, seed=5, n_bars=40, n_tracks=8)
    _oracle_compare(torch, events, songs, None)
    _oracle_compare(torch, events, songs, abi.make_filter(synth.POPULAR, 100, crop_channel_bars=80))


def test_encode_empty_and_degenerate_songs(torch):
    ev = np.zeros(3, dtype=abi.EVENT_DTYPE)
    ev['kind'] = [abi.EV_SET_TEMPO, abi.EV_PROGRAM_CHANGE, abi.EV_NOTE_OFF]
    ev['aux'][0] = 500000
    good, gs = synth.synth_corpus(1, seed=3, n_bars=8, n_tracks=3)
    arrays = [np.zeros(0, dtype=abi.EVENT_DTYPE), ev, good]
    infos, arena = lib.encode_songs(arrays, [480, 480, 480], abi.make_filter(synth.POPULAR, 0))
    assert [int(i['status']) for i in infos] == [abi.SONG_NO_NOTES, abi.SONG_NO_NOTES, abi.SONG_OK]


def test_batch_composition_independence_and_idempotence(torch):
    """A song's grid does not depend on its neighbours in the batch, and a second pass over the same
    device buffers rewrites identical bytes (zero-kept shared-memory tiles are restored correctly)."""
    events, songs = synth.synth_corpus(40, seed=77, n_bars=32, n_tracks=10)
    filt = abi.make_filter(synth.POPULAR, 100)
    per_song = synth.split_songs(events, songs)
    infos, arena = lib.encode_songs(per_song, songs['ticks_per_beat'], filt)
    sums = []
    for info in infos:
        p, u = lib.song_grids(info, arena)
        sums.append((hashlib.sha256(p.cpu().numpy().tobytes()).hexdigest(),
                     hashlib.sha256(u.cpu().numpy().tobytes()).hexdigest() if u is not None else ''))
    order = np.random.default_rng(0).permutation(len(per_song))
    infos2, arena2 = lib.encode_songs([per_song[i] for i in order], songs['ticks_per_beat'][order], filt)
    for j, i in enumerate(order):
        p, u = lib.song_grids(infos2[j], arena2)
        assert hashlib.sha256(p.cpu().numpy().tobytes()).hexdigest() == sums[i][0]
    # idempotence on a poisoned arena
    ev_all, song_tab = lib.make_songs(per_song, songs['ticks_per_beat'])
    batch = lib.DeviceBatch(ev_all, song_tab)
    arena3 = torch.full_like(arena, 7.0)
    batch.encode(filt, arena3)
    batch.encode(filt, arena3)
    torch.cuda.synchronize()
    assert torch.equal(arena3, arena)


def test_arena_full_is_reported_not_written(torch):
    events, songs = synth.synth_corpus(3, seed=9, n_bars=16, n_tracks=4)
    filt = abi.make_filter(synth.POPULAR, 0)
    per_song = synth.split_songs(events, songs)
    infos, arena = lib.encode_songs(per_song, songs['ticks_per_beat'], filt)
    need0 = lib.grid_floats(infos[0])
    ev_all, song_tab = lib.make_songs(per_song, songs['ticks_per_beat'])
    batch = lib.DeviceBatch(ev_all, song_tab)
    small = torch.full((need0 + 10,), -1.0, dtype=torch.float32, device='cuda')
    batch.encode(filt, small)
    got = batch.infos()
    assert [int(i['status']) for i in got] == [abi.SONG_OK, abi.SONG_ARENA_FULL, abi.SONG_ARENA_FULL]
    assert torch.equal(small[:need0], arena[:need0])
    assert bool((small[need0:] == -1.0).all())
    assert int(got[1]['n_pitched']) == int(infos[1]['n_pitched'])   # info stays valid


def test_host_entry_point(torch):
    """mst_encode_batch_host: host buffers in, infos (+ grids) out."""
    L = lib.load()
    events, songs = synth.synth_corpus(3, seed=11, n_bars=16, n_tracks=5)
    filt = abi.make_filter(synth.POPULAR, 0)
    per_song = synth.split_songs(events, songs)
    infos, arena = lib.encode_songs(per_song, songs['ticks_per_beat'], filt)
    n = arena.numel()
    d_arena = torch.empty(n, dtype=torch.float32, device='cuda')
    h_arena = np.empty(n, dtype=np.float32)
    h_info = np.zeros(len(songs), dtype=abi.SONG_INFO_DTYPE)
    wsb = L.mst_encode_host_workspace_bytes(len(events), len(songs), n)
    ws = torch.empty(wsb, dtype=torch.uint8, device='cuda')
    rc = L.mst_encode_batch_host(events.ctypes.data, len(events), songs.ctypes.data, len(songs),
                                 filt.reshape(1).ctypes.data, h_info.ctypes.data, d_arena.data_ptr(), n,
                                 h_arena.ctypes.data, ws.data_ptr(), wsb, None)
    lib.check(rc)
    assert np.array_equal(h_arena, arena.cpu().numpy())
    np.testing.assert_array_equal(h_info['pitched_offset'], infos['pitched_offset'])


def test_host_async_entry_point_two_slots(torch):
    """mst_encode_batch_host_async on two alternating streams / workspaces / arenas (the pipelined host path of
    bench.py's e2e leg): every batch's infos and grids equal the synchronous device path's."""
    L = lib.load()
    filt = abi.make_filter(synth.POPULAR, 0)
    batches = []
    for seed in (21, 22, 23, 24):
        events, songs = synth.synth_corpus(3, seed=seed, n_bars=12, n_tracks=5)
        infos, arena = lib.encode_songs(synth.split_songs(events, songs), songs['ticks_per_beat'], filt)
        batches.append((events, songs, infos, arena))
    n = max(b[3].numel() for b in batches)
    wsb = max(L.mst_encode_host_workspace_bytes(len(b[0]), len(b[1]), n) for b in batches)
    slots = [dict(stream=torch.cuda.Stream(), ws=torch.empty(wsb, dtype=torch.uint8, device='cuda'),
                  arena=torch.empty(n, dtype=torch.float32, device='cuda'),
                  info=torch.zeros(3 * abi.SONG_INFO_DTYPE.itemsize, dtype=torch.uint8).pin_memory(), job=None) for _ in range(2)]
    pinned = [(torch.from_numpy(b[0].view(np.uint8).reshape(-1).copy()).pin_memory(),
               torch.from_numpy(b[1].view(np.uint8).reshape(-1).copy()).pin_memory()) for b in batches]

    def drain(sl):
        sl['stream'].synchronize()
        if sl['job'] is not None:
            events, songs, infos, arena = batches[sl['job']]
            got = sl['info'].numpy().view(abi.SONG_INFO_DTYPE)
            for k in ('status', 'n_pitched', 'n_bars_out', 'pitched_offset', 'unpitched_offset', 'tonic', 'mode'):
                np.testing.assert_array_equal(got[k], infos[k])
            assert torch.equal(sl['arena'][:arena.numel()], arena)
            sl['job'] = None

    for rep in range(2):
        for i, (ev, sg) in enumerate(pinned):
            sl = slots[i % 2]
            drain(sl)
            lib.check(L.mst_encode_batch_host_async(ev.data_ptr(), len(batches[i][0]), sg.data_ptr(), 3,
                                                    filt.reshape(1).ctypes.data, sl['info'].data_ptr(), sl['arena'].data_ptr(), n,
                                                    sl['ws'].data_ptr(), wsb, sl['stream'].cuda_stream))
            sl['job'] = i
    for sl in slots:
        drain(sl)


def test_packed_event_entry_point_equals_the_16_byte_path(torch):
    """mst_events_pack + mst_encode_batch_packed_async (8-byte wire records, widened on the device): infos and grids are
    bit-identical to the MstEvent path, including songs with set_tempo / key_signature (their aux travels in other bytes),
    every bundled golden song and a file written by the pipeline's loader."""
    L = lib.load()
    filt = abi.make_filter(synth.POPULAR, 0)
    events, songs = synth.synth_corpus(4, seed=31, n_bars=10, n_tracks=6)
    per = synth.split_songs(events, songs)
    tpbs = [int(t) for t in songs['ticks_per_beat']]
    for s in gu.load_songs('codec_bundled')[:6] + gu.load_songs('codec_synth'):
        per.append(s['events'])
        tpbs.append(int(s['ticks_per_beat']))
    ev, sg = lib.make_songs(per, tpbs)
    assert (ev['kind'] == abi.EV_SET_TEMPO).any()
    infos, arena = lib.encode_songs(per, tpbs, filt)
    packed = lib.events_pack(ev)
    n = arena.numel()
    wsb = L.mst_encode_packed_workspace_bytes(len(ev), len(sg), n)
    ws = torch.empty(wsb, dtype=torch.uint8, device='cuda')
    d_arena = torch.full((n,), -1.0, dtype=torch.float32, device='cuda')
    h_packed = torch.from_numpy(packed.view(np.uint8).copy()).pin_memory()
    h_songs = torch.from_numpy(sg.view(np.uint8).reshape(-1).copy()).pin_memory()
    h_info = torch.zeros(len(sg) * abi.SONG_INFO_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
    st = torch.cuda.Stream()
    lib.check(L.mst_encode_batch_packed_async(h_packed.data_ptr(), len(ev), h_songs.data_ptr(), len(sg), filt.reshape(1).ctypes.data,
                                              h_info.data_ptr(), d_arena.data_ptr(), n, ws.data_ptr(), wsb, st.cuda_stream))
    st.synchronize()
    got = h_info.numpy().view(abi.SONG_INFO_DTYPE)
    for k in ('status', 'n_pitched', 'n_unpitched', 'n_bars_out', 'pitched_offset', 'unpitched_offset', 'tonic', 'mode', 'tempo', 'bpm',
              'key_signature', 'duration', 'n_notes'):
        np.testing.assert_array_equal(got[k], infos[k], err_msg=k)
    assert torch.equal(d_arena, arena)
