"""Loading of the committed golden vectors (tests/golden/*.npz, written by oracle/gen_golden.py)."""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# popular_instruments of the reference (style/midi.py:23-64), as data for the iter_inputs filter
INCLUDED = [0, 25, 48, 33, 1, 27, 49, 29, 35, 30, 50, 24, 5, 4, 32, 52, 26, 18, 28, 89, 65, 53, 61, 2, 17,
            73, 54, 62, 16, 39, 34, 51, 90, 56, 66, 38, 11, 81, 3, 57]

_cache = {}


def load_songs(name):
    """name: 'codec_bundled' (min_n_messages=100) or 'codec_synth' (min_n_messages=0)."""
    if name not in _cache:
        z = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
        n = int(z['n_songs'])
        songs = [dict() for _ in range(n)]
        for key in z.files:
            if '/' in key:
                i, k = key.split('/', 1)
                songs[int(i)][k] = z[key]
        for s in songs:
            s['name'] = bytes(s['name']).decode('utf-8')
        _cache[name] = songs
    return _cache[name]


def min_n_messages(name):
    return 100 if name == 'codec_bundled' else 0


def dense(shape, idx, val):
    a = np.zeros(int(np.prod(shape)), dtype=np.float32)
    a[idx] = val
    return a.reshape([int(x) for x in shape])


def decode_inputs(s):
    """(grid buffer, MstDecodeSong) for a golden song: what style_transfer.save feeds decode_midi."""
    from mst_native import abi
    p = dense(s['pitched_shape'], s['pitched_idx'], s['pitched_val'])
    C, B = int(s['pitched_shape'][0]), int(s['pitched_shape'][1])
    parts = [p.ravel()]
    has_u = 'unpitched_shape' in s
    if has_u:
        parts.append(dense(s['unpitched_shape'], s['unpitched_idx'], s['unpitched_val']).ravel())
    grids = np.ascontiguousarray(np.concatenate(parts))
    g = s['info']
    info = dict(ticks_per_beat=g[0], numerator=g[1], denominator_log2=g[2], ticks_per_bar=g[4], n_beats=g[6],
                tempo=g[7], tonic=s['scale'][0], mode=s['scale'][1])
    d = abi.make_decode_song(info, C, has_u, B, instrument_ids=list(s['instruments']), duration=int(g[3]),
                             pitched_offset=0, unpitched_offset=p.size)
    return grids, d
