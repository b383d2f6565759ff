"""CPU tests: the C oracle's decode against the messages the unmodified reference produced
(decode_midi on the encoded grids, oracle/gen_golden.py)."""
import numpy as np
import pytest

import golden_util as gu
import oracle_lib
from mst_native import abi

CASES = [(name, i) for name in ('codec_bundled', 'codec_synth') for i, s in enumerate(gu.load_songs(name))
         if 'decoded_notes' in s]


@pytest.mark.parametrize('name,i', CASES)
def test_oracle_decode_matches_reference(name, i):
    s = gu.load_songs(name)[i]
    grids, d = gu.decode_inputs(s)
    res, msgs = oracle_lib.decode_song(grids, d)
    assert int(res['status']) == abi.DECODE_OK
    g = s['decoded_notes']
    assert len(msgs) == len(g)
    np.testing.assert_array_equal(msgs['delta'], g[:, 0])
    np.testing.assert_array_equal(msgs['is_on'], g[:, 1])
    np.testing.assert_array_equal(msgs['channel'], g[:, 2])
    np.testing.assert_array_equal(msgs['note'], g[:, 3])
    np.testing.assert_array_equal(msgs['velocity'], g[:, 4])
    assert int(res['eot_delta']) == int(s['decoded_end_delta'])
