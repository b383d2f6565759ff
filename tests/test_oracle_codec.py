"""CPU tests: the C oracle (oracle/mst_oracle.c) against golden vectors produced by the
unmodified reference (oracle/gen_golden.py).  This is what pins the oracle."""
import hashlib

import numpy as np
import pytest

import golden_util as gu
import oracle_lib
from mst_native import abi

CASES = [(name, i) for name in ('codec_bundled', 'codec_synth') for i in range(len(gu.load_songs(name)))]


@pytest.mark.parametrize('name,i', CASES)
def test_oracle_encode_matches_reference(name, i):
    s = gu.load_songs(name)[i]
    filt = abi.make_filter(gu.INCLUDED, gu.min_n_messages(name))
    info, pitched, unpitched = oracle_lib.encode_song(s['events'], s['ticks_per_beat'], filt)
    assert info['status'] == int(s['status']), s['name']
    if 'lists' in s:
        _, lists = oracle_lib.read_midi_lists(s['events'], s['ticks_per_beat'])
        np.testing.assert_array_equal(lists, s['lists'])
        g = s['info']
        got = [info['ticks_per_beat'], info['numerator'], info['denominator_log2'], info['duration'],
               info['ticks_per_bar'], info['n_bars'], info['n_beats'], info['tempo'], info['bpm']]
        assert [int(x) for x in got] == [int(x) for x in g]
    if int(s['status']) != abi.SONG_OK:
        return
    assert (int(info['tonic']), int(info['mode'])) == tuple(int(x) for x in s['scale'])
    np.testing.assert_allclose(
        [info['scale_coverage'], info['scale_loose_coverage'], info['scale_cross_entropy'], info['scale_loss']],
        s['scale_stats'], rtol=1e-12)
    C = int(info['n_pitched'])
    np.testing.assert_array_equal(info['pitched_instruments'][:C], s['instruments'])
    assert list(pitched.shape) == [int(x) for x in s['pitched_shape']]
    assert hashlib.sha256(pitched.tobytes()).digest() == bytes(s['pitched_sha256'])
    if 'unpitched_shape' in s:
        assert list(unpitched.shape) == [int(x) for x in s['unpitched_shape']]
        assert hashlib.sha256(unpitched.tobytes()).digest() == bytes(s['unpitched_sha256'])
    else:
        assert unpitched is None
