"""ctypes wrapper of oracle/libmst_oracle.so (the CPU restatement in mst_oracle.c).

TEST INFRASTRUCTURE ONLY — importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never from the product package.
"""
import ctypes
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(_HERE), 'music-style-transfer_b200'))
from mst_native import abi  # noqa: E402  (struct layouts only)

_lib = None


def build():
    subprocess.check_call(['make', '-s', '-C', _HERE])


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, 'libmst_oracle.so')
        src = os.path.join(_HERE, 'mst_oracle.c')
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
            build()
        _lib = ctypes.CDLL(path)
        _lib.msto_sizeof_song_info.restype = ctypes.c_size_t
        _lib.msto_sizeof_event.restype = ctypes.c_size_t
        assert _lib.msto_sizeof_song_info() == abi.SONG_INFO_DTYPE.itemsize
        assert _lib.msto_sizeof_event() == abi.EVENT_DTYPE.itemsize
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def encode_song(events, tpb, filt=None):
    """events: EVENT_DTYPE array.  Returns (info record, pitched f32 ndarray or None, unpitched or None)."""
    L = lib()
    events = np.ascontiguousarray(events, dtype=abi.EVENT_DTYPE)
    info = np.zeros(1, dtype=abi.SONG_INFO_DTYPE)
    fptr = _p(filt.reshape(1)) if filt is not None else None
    L.msto_encode_song(_p(events), ctypes.c_uint32(len(events)), ctypes.c_int32(int(tpb)), fptr, _p(info),
                       None, ctypes.c_uint64(0), None, ctypes.c_uint64(0))
    rec = info[0]
    if rec['status'] != abi.SONG_OK:
        return rec.copy(), None, None
    C, U, B, nb = int(rec['n_pitched']), int(rec['n_unpitched']), int(rec['n_bars_out']), int(rec['n_beats'])
    pitched = np.empty((C, B, nb, 10, 56, 5), dtype=np.float32)
    unpitched = np.empty((U, B, nb, 10, 47, 2), dtype=np.float32) if U else None
    info2 = np.zeros(1, dtype=abi.SONG_INFO_DTYPE)
    L.msto_encode_song(_p(events), ctypes.c_uint32(len(events)), ctypes.c_int32(int(tpb)), fptr, _p(info2),
                       _p(pitched), ctypes.c_uint64(pitched.size),
                       _p(unpitched), ctypes.c_uint64(unpitched.size if U else 0))
    return info2[0].copy(), pitched, unpitched


def read_midi_lists(events, tpb):
    L = lib()
    events = np.ascontiguousarray(events, dtype=abi.EVENT_DTYPE)
    info = np.zeros(1, dtype=abi.SONG_INFO_DTYPE)
    cap = 16 * abi.MAX_LISTS_PER_CHANNEL
    ch = np.zeros(cap, np.int32); ins = np.zeros(cap, np.int32); cnt = np.zeros(cap, np.int32)
    n = L.msto_read_midi_lists(_p(events), ctypes.c_uint32(len(events)), ctypes.c_int32(int(tpb)), _p(info),
                               _p(ch), _p(ins), _p(cnt), ctypes.c_int(cap))
    return info[0].copy(), np.stack([ch[:n], ins[:n], cnt[:n]], 1)


def encode_batch(events, songs, filt=None, n_threads=1):
    """Timed CPU baseline: encode every song of the batch on n_threads host threads (OpenMP).
    Returns (infos, checksum)."""
    L = lib()
    L.msto_encode_batch.restype = ctypes.c_double
    events = np.ascontiguousarray(events, dtype=abi.EVENT_DTYPE)
    songs = np.ascontiguousarray(songs, dtype=abi.SONG_IN_DTYPE)
    infos = np.zeros(len(songs), dtype=abi.SONG_INFO_DTYPE)
    fptr = _p(filt.reshape(1)) if filt is not None else None
    chk = L.msto_encode_batch(_p(events), _p(songs), ctypes.c_uint32(len(songs)), fptr, _p(infos),
                              ctypes.c_int(int(n_threads)))
    return infos, float(chk)


def decode_song(grids, dsong):
    """grids: contiguous float32 array the song's offsets index into (mutated when write_back).
    Returns (result record, messages[OUT_MSG_DTYPE])."""
    L = lib()
    for f, dt in (('msto_sizeof_decode_song', abi.DECODE_SONG_DTYPE), ('msto_sizeof_out_msg', abi.OUT_MSG_DTYPE),
                  ('msto_sizeof_decode_result', abi.DECODE_RESULT_DTYPE)):
        getattr(L, f).restype = ctypes.c_size_t
        assert getattr(L, f)() == dt.itemsize, f
    assert grids.dtype == np.float32 and grids.flags['C_CONTIGUOUS']
    res = np.zeros(1, dtype=abi.DECODE_RESULT_DTYPE)
    d = np.ascontiguousarray(dsong.reshape(1))
    cap = 1 << 16
    while True:
        out = np.zeros(cap, dtype=abi.OUT_MSG_DTYPE)
        L.msto_decode_song(_p(grids), _p(d), _p(out), ctypes.c_uint32(cap), _p(res))
        if res[0]['status'] == abi.DECODE_CAPACITY:
            cap = int(res[0]['n_messages'])
            continue
        break
    n = int(res[0]['n_messages']) if res[0]['status'] == abi.DECODE_OK else 0
    return res[0].copy(), out[:n].copy()
