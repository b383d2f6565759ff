"""ctypes binding of libmst_b200.so (C ABI: include/mst_b200.h).

torch is used only for device memory and streams.  There is no CPU fallback: every compute entry
point raises when the library or a CUDA device is missing.
"""
import ctypes
import os

import numpy as np

from . import abi

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(_PKG, 'libmst_b200.so')

_lib = None

# every symbol include/mst_b200.h declares (tests check that the library exports them all)
EXPORTS = [
    'mst_abi_version', 'mst_last_error', 'mst_device_count',
    'mst_profile_enable', 'mst_profile_encode_ms',
    'mst_smf_parse', 'mst_smf_load_batch', 'mst_smf_write_events',
    'mst_encode_workspace_bytes', 'mst_encode_batch_device',
    'mst_encode_host_workspace_bytes', 'mst_encode_batch_host', 'mst_encode_batch_host_async',
    'mst_events_pack', 'mst_encode_packed_workspace_bytes', 'mst_encode_batch_packed_async',
    'mst_files_read_batch', 'mst_smf_parse_device_workspace_bytes', 'mst_smf_parse_batch_device',
    'mst_encode_smf_workspace_bytes', 'mst_encode_batch_smf_async',
    'mst_decode_workspace_bytes', 'mst_decode_batch_device', 'mst_smf_write', 'mst_transfer_select',
    'mst_model_param_count', 'mst_model_param_name', 'mst_model_param_numel', 'mst_model_param_offset',
    'mst_model_workspace_bytes', 'mst_model_forward', 'mst_model_train_step', 'mst_model_predict_apply', 'mst_model_song_info', 'mst_model_backward', 'mst_model_loss_workspace_bytes',
    'mst_model_loss', 'mst_model_loss_noalloc', 'mst_adam_step', 'mst_model_gemm',
]


class MstError(RuntimeError):
    pass


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MstError(f'{LIB_PATH} is missing: build it with `python music-style-transfer_b200/build.py` '
                       '(there is no CPU fallback)')
    L = ctypes.CDLL(LIB_PATH)
    L.mst_abi_version.restype = ctypes.c_int
    L.mst_last_error.restype = ctypes.c_char_p
    L.mst_device_count.restype = ctypes.c_int
    L.mst_profile_enable.restype = ctypes.c_int
    L.mst_profile_enable.argtypes = [ctypes.c_int]
    L.mst_profile_encode_ms.restype = ctypes.c_float
    L.mst_profile_encode_ms.argtypes = [ctypes.c_int]
    L.mst_smf_parse.restype = ctypes.c_int64
    L.mst_smf_parse.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t,
                                ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32),
                                ctypes.POINTER(ctypes.c_size_t)]
    L.mst_smf_load_batch.restype = ctypes.c_int
    L.mst_smf_load_batch.argtypes = [ctypes.POINTER(ctypes.c_char_p), ctypes.c_uint32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_uint64,
                                     ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64)]
    L.mst_smf_write_events.restype = ctypes.c_int64
    L.mst_smf_write_events.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int32, ctypes.c_void_p, ctypes.c_size_t,
                                       ctypes.POINTER(ctypes.c_size_t)]
    for name in ('mst_encode_workspace_bytes', 'mst_encode_host_workspace_bytes'):
        f = getattr(L, name)
        f.restype = ctypes.c_size_t
        f.argtypes = [ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint64]
    L.mst_encode_batch_device.restype = ctypes.c_int
    L.mst_encode_batch_device.argtypes = [
        ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p,
        ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    L.mst_encode_batch_host.restype = ctypes.c_int
    L.mst_encode_batch_host.argtypes = [
        ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p,
        ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    L.mst_encode_batch_host_async.restype = ctypes.c_int
    L.mst_encode_batch_host_async.argtypes = [
        ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p,
        ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    L.mst_events_pack.restype = None
    L.mst_events_pack.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
    L.mst_encode_packed_workspace_bytes.restype = ctypes.c_size_t
    L.mst_encode_packed_workspace_bytes.argtypes = [ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint64]
    L.mst_encode_batch_packed_async.restype = ctypes.c_int
    L.mst_encode_batch_packed_async.argtypes = L.mst_encode_batch_host_async.argtypes
    L.mst_files_read_batch.restype = ctypes.c_int
    L.mst_files_read_batch.argtypes = [ctypes.POINTER(ctypes.c_char_p), ctypes.c_uint32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_uint64,
                                       ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64)]
    L.mst_smf_parse_device_workspace_bytes.restype = ctypes.c_size_t
    L.mst_smf_parse_device_workspace_bytes.argtypes = [ctypes.c_uint64, ctypes.c_uint32]
    L.mst_smf_parse_batch_device.restype = ctypes.c_int
    L.mst_smf_parse_batch_device.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p,
                                             ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64),
                                             ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    L.mst_encode_smf_workspace_bytes.restype = ctypes.c_size_t
    L.mst_encode_smf_workspace_bytes.argtypes = [ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint64]
    L.mst_encode_batch_smf_async.restype = ctypes.c_int
    L.mst_encode_batch_smf_async.argtypes = [
        ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p,
        ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    L.mst_decode_workspace_bytes.restype = ctypes.c_size_t
    L.mst_decode_workspace_bytes.argtypes = [ctypes.c_uint32, ctypes.c_uint64, ctypes.c_uint64]
    L.mst_decode_batch_device.restype = ctypes.c_int
    L.mst_decode_batch_device.argtypes = [
        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64,
        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    L.mst_smf_write.restype = ctypes.c_int64
    L.mst_smf_write.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int32, ctypes.c_void_p,
                                ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)]
    L.mst_transfer_select.restype = ctypes.c_int
    L.mst_transfer_select.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                      ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p,
                                      ctypes.c_void_p, ctypes.c_void_p]
    if L.mst_abi_version() != abi.MST_ABI_VERSION:
        raise MstError('libmst_b200.so ABI version mismatch: rebuild')
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise MstError(f'libmst_b200 error {rc}: {load().mst_last_error().decode()}')


def require_gpu():
    L = load()
    if L.mst_device_count() == 0:
        raise MstError('no CUDA device visible: the B200 path has no CPU fallback')
    return L


def _np_ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


# ---------------------------------------------------------------------------------- SMF parsing
def smf_parse(data):
    """bytes -> (events ndarray[EVENT_DTYPE], ticks_per_beat) or None when the load fails the way
    mido.MidiFile would (style/midi.py:104-108)."""
    L = load()
    buf = np.frombuffer(data, dtype=np.uint8)
    tpb, ntr, need = ctypes.c_int32(0), ctypes.c_int32(0), ctypes.c_size_t(0)
    # an event needs at least 2 bytes (delta + running-status data byte): size once, no second pass
    cap = max(1, len(buf) // 2)
    out = np.empty(cap, dtype=abi.EVENT_DTYPE)
    n = L.mst_smf_parse(_np_ptr(buf), len(buf), _np_ptr(out), cap, ctypes.byref(tpb), ctypes.byref(ntr),
                        ctypes.byref(need))
    if n < 0:
        return None
    return out[:n].copy(), int(tpb.value)


def smf_load_batch(paths, events_out, songs_out=None, status_out=None, n_threads=0):
    """Parse `paths` on a C++ thread pool into `events_out` (ndarray[EVENT_DTYPE], ideally a view of page-locked memory).
    Returns (n_events_total, songs ndarray[SONG_IN_DTYPE], status int32 ndarray) or raises MstError(capacity) carrying the
    needed record count in `.needed`.  The GIL is released for the duration of the call."""
    L = load()
    n = len(paths)
    songs = np.zeros(n, dtype=abi.SONG_IN_DTYPE) if songs_out is None else songs_out[:n]
    status = np.zeros(n, dtype=np.int32) if status_out is None else status_out[:n]
    arr = (ctypes.c_char_p * max(n, 1))(*[os.fsencode(p) for p in paths])
    total = ctypes.c_uint64(0)
    rc = L.mst_smf_load_batch(arr, n, int(n_threads), _np_ptr(events_out), len(events_out), _np_ptr(songs), _np_ptr(status),
                              ctypes.byref(total))
    if rc == -3:
        e = MstError(f'event buffer too small: {total.value} records needed')
        e.needed = int(total.value)
        raise e
    check(rc)
    return int(total.value), songs, status


def files_read_batch(paths, bytes_out, files_out=None, status_out=None, n_threads=0):
    """Read `paths` as they are into `bytes_out` (uint8 ndarray, ideally a view of page-locked memory) on a C++ thread pool.
    Returns (n_bytes_total, files ndarray[FILE_IN_DTYPE], status int32 ndarray) or raises MstError carrying `.needed`."""
    L = load()
    n = len(paths)
    files = np.zeros(n, dtype=abi.FILE_IN_DTYPE) if files_out is None else files_out[:n]
    status = np.zeros(n, dtype=np.int32) if status_out is None else status_out[:n]
    arr = (ctypes.c_char_p * max(n, 1))(*[os.fsencode(p) for p in paths])
    total = ctypes.c_uint64(0)
    rc = L.mst_files_read_batch(arr, n, int(n_threads), _np_ptr(bytes_out), len(bytes_out), _np_ptr(files), _np_ptr(status),
                                ctypes.byref(total))
    if rc == -3:
        e = MstError(f'byte buffer too small: {total.value} bytes needed')
        e.needed = int(total.value)
        raise e
    check(rc)
    return int(total.value), files, status


def concat_files(blobs):
    """[bytes, ...] -> (uint8 ndarray of the files on 16-byte boundaries, files ndarray[FILE_IN_DTYPE]): the batch buffer of
    mst_encode_batch_smf_async for files already in memory."""
    files = np.zeros(len(blobs), dtype=abi.FILE_IN_DTYPE)
    off = 0
    for i, b in enumerate(blobs):
        files[i]['byte_offset'], files[i]['n_bytes'] = off, len(b)
        off += (len(b) + 15) & ~15
    buf = np.zeros(max(off, 1), dtype=np.uint8)
    for i, b in enumerate(blobs):
        o = int(files[i]['byte_offset'])
        buf[o:o + len(b)] = np.frombuffer(bytes(b), dtype=np.uint8)
    return buf, files


def smf_parse_device(blobs):
    """Parse files given as raw bytes ON THE DEVICE (mst_smf_parse_batch_device).  Returns [(events, ticks_per_beat) or None]
    per file, exactly what smf_parse returns for each."""
    import torch
    L = require_gpu()
    buf, files = concat_files(blobs)
    n = len(blobs)
    ws = torch.empty(L.mst_smf_parse_device_workspace_bytes(len(buf), n), dtype=torch.uint8, device='cuda')
    events = np.empty(len(buf) // 2 + 1, dtype=abi.EVENT_DTYPE)
    songs = np.zeros(n, dtype=abi.SONG_IN_DTYPE)
    status = np.zeros(n, dtype=np.int32)
    total = ctypes.c_uint64(0)
    check(L.mst_smf_parse_batch_device(_np_ptr(buf), len(buf), _np_ptr(files), n, _np_ptr(events), len(events), _np_ptr(songs),
                                       _np_ptr(status), ctypes.byref(total), ws.data_ptr(), ws.numel(),
                                       torch.cuda.current_stream().cuda_stream))
    out = []
    for i in range(n):
        if status[i] != 0:
            out.append(None)
        else:
            o, c = int(songs[i]['event_offset']), int(songs[i]['n_events'])
            out.append((events[o:o + c].copy(), int(songs[i]['ticks_per_beat'])))
    return out


def encode_smf(blobs, filt=None, arena_floats=None):
    """Encode files given as raw SMF bytes through mst_encode_batch_smf_async (device-side parse).  Synchronous convenience
    wrapper for tests and small jobs: returns (infos ndarray[SONG_INFO_DTYPE], arena CUDA tensor)."""
    import torch
    L = require_gpu()
    buf, files = concat_files(blobs)
    n = len(blobs)
    if filt is None:
        filt = abi.make_filter(None, 0, 0)
    filt = np.ascontiguousarray(filt).reshape(1)
    infos = np.zeros(n, dtype=abi.SONG_INFO_DTYPE)
    if arena_floats is None:
        # planning pass (no arena) for the sizes, then the real pass
        ws0 = torch.empty(L.mst_encode_smf_workspace_bytes(len(buf), n, 0), dtype=torch.uint8, device='cuda')
        check(L.mst_encode_batch_smf_async(_np_ptr(buf), len(buf), _np_ptr(files), n, _np_ptr(filt), _np_ptr(infos), None, 0,
                                           ws0.data_ptr(), ws0.numel(), torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        arena_floats = max(1, sum(grid_floats(r) for r in infos if r['status'] in (abi.SONG_OK, abi.SONG_ARENA_FULL)))
        del ws0
    arena = torch.empty(int(arena_floats), dtype=torch.float32, device='cuda')
    ws = torch.empty(L.mst_encode_smf_workspace_bytes(len(buf), n, arena.numel()), dtype=torch.uint8, device='cuda')
    check(L.mst_encode_batch_smf_async(_np_ptr(buf), len(buf), _np_ptr(files), n, _np_ptr(filt), _np_ptr(infos), arena.data_ptr(),
                                       arena.numel(), ws.data_ptr(), ws.numel(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return infos, arena


def events_pack(events, out=None):
    """MstEvent records -> the 8-byte wire format (uint64 ndarray; `out` may be a view of page-locked memory)."""
    L = load()
    events = np.ascontiguousarray(events, dtype=abi.EVENT_DTYPE)
    if out is None:
        out = np.empty(len(events), dtype=np.uint64)
    L.mst_events_pack(_np_ptr(events), len(events), _np_ptr(out))
    return out


def smf_write_events(events, ticks_per_beat):
    """MstEvent records (track-major, absolute ticks) -> Standard MIDI File bytes (mst_smf_write_events)."""
    L = load()
    events = np.ascontiguousarray(events, dtype=abi.EVENT_DTYPE)
    cap = 64 + 16 * (int(events['track'].max()) + 2 if len(events) else 2) + 8 * len(events)
    out = np.empty(cap, dtype=np.uint8)
    need = ctypes.c_size_t(0)
    n = L.mst_smf_write_events(_np_ptr(events), len(events), int(ticks_per_beat), _np_ptr(out), cap, ctypes.byref(need))
    if n < 0:
        raise MstError(f'mst_smf_write_events failed ({n})')
    return out[:n].tobytes()


# ---------------------------------------------------------------------------------- batches
def make_songs(event_arrays, tpbs):
    """Concatenate per-song event arrays into (events, songs[SONG_IN_DTYPE])."""
    songs = np.zeros(len(event_arrays), dtype=abi.SONG_IN_DTYPE)
    off = 0
    for i, (e, t) in enumerate(zip(event_arrays, tpbs)):
        songs[i]['event_offset'] = off
        songs[i]['n_events'] = len(e)
        songs[i]['ticks_per_beat'] = int(t)
        off += len(e)
    events = np.concatenate(event_arrays) if event_arrays else np.zeros(0, abi.EVENT_DTYPE)
    return np.ascontiguousarray(events, dtype=abi.EVENT_DTYPE), songs


def grid_floats(info):
    """Floats one encoded song occupies in the arena (pitched then unpitched)."""
    per_bar = int(info['n_beats']) * (int(info['n_pitched']) * abi.PITCHED_BEAT_FLOATS
                                       + int(info['n_unpitched']) * abi.UNPITCHED_BEAT_FLOATS)
    return int(info['n_bars_out']) * per_bar


class DeviceBatch:
    """A batch of songs resident in HBM: events + song table (+ reusable workspace / info)."""

    def __init__(self, events, songs, device=None):
        import torch
        require_gpu()
        self.torch = torch
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        self.n_events = int(len(events))
        self.n_songs = int(len(songs))
        ev_bytes = torch.from_numpy(events.view(np.uint8).reshape(-1)) if self.n_events else torch.zeros(16, dtype=torch.uint8)
        self.d_events = ev_bytes.to(self.device)
        self.d_songs = torch.from_numpy(songs.view(np.uint8).reshape(-1)).to(self.device)
        self.d_info = torch.empty(max(1, self.n_songs) * abi.SONG_INFO_DTYPE.itemsize, dtype=torch.uint8, device=self.device)
        self._ws = None
        self._ws_arena = -1

    def workspace(self, arena_floats):
        L = load()
        need = L.mst_encode_workspace_bytes(self.n_events, self.n_songs, arena_floats)
        if self._ws is None or self._ws.numel() < need:
            self._ws = self.torch.empty(need, dtype=self.torch.uint8, device=self.device)
        return self._ws

    def encode(self, filt, arena, stream=None):
        """Launch the encode pipeline (asynchronous).  arena: float32 CUDA tensor or None (plan only)."""
        L = load()
        torch = self.torch
        arena_floats = 0 if arena is None else arena.numel()
        ws = self.workspace(arena_floats)
        s = torch.cuda.current_stream(self.device) if stream is None else stream
        check(L.mst_encode_batch_device(
            self.d_events.data_ptr(), self.n_events, self.d_songs.data_ptr(), self.n_songs,
            _np_ptr(filt.reshape(1)) if filt is not None else None, self.d_info.data_ptr(),
            arena.data_ptr() if arena is not None else None, arena_floats,
            ws.data_ptr(), ws.numel(), s.cuda_stream))

    def infos(self):
        """Synchronising D2H read of the MstSongInfo records."""
        raw = self.d_info.cpu().numpy()
        return raw[:self.n_songs * abi.SONG_INFO_DTYPE.itemsize].view(abi.SONG_INFO_DTYPE).copy()


def encode_songs(event_arrays, tpbs, filt=None, device=None):
    """Plan + fill: returns (infos, arena float32 CUDA tensor).  Grids of song i live at
    infos[i]['pitched_offset'] / ['unpitched_offset'] (float offsets into the arena)."""
    import torch
    events, songs = make_songs(event_arrays, tpbs)
    batch = DeviceBatch(events, songs, device)
    batch.encode(filt, None)
    plan = batch.infos()
    total = sum(grid_floats(r) for r in plan if r['status'] in (abi.SONG_OK, abi.SONG_ARENA_FULL))
    arena = torch.empty(max(total, 1), dtype=torch.float32, device=batch.device)
    batch.encode(filt, arena)
    return batch.infos(), arena


def song_grids(info, arena):
    """Views (no copy) of one song's grids inside the arena: (pitched, unpitched or None)."""
    C, U, B, nb = (int(info[k]) for k in ('n_pitched', 'n_unpitched', 'n_bars_out', 'n_beats'))
    po, uo = int(info['pitched_offset']), int(info['unpitched_offset'])
    pitched = arena[po:po + C * B * nb * abi.PITCHED_BEAT_FLOATS].view(C, B, nb, 10, 56, 5)
    unpitched = None
    if U:
        unpitched = arena[uo:uo + U * B * nb * abi.UNPITCHED_BEAT_FLOATS].view(U, B, nb, 10, 47, 2)
    return pitched, unpitched


# ---------------------------------------------------------------------------------- decode
def decode_chunks(dsongs):
    """Number of (channel, bar) chunks of a MstDecodeSong table."""
    return int(((dsongs['n_pitched'].astype(np.int64) + (dsongs['has_unpitched'] != 0)) * dsongs['n_bars']).sum())


def decode_batch(grids, dsongs, msg_capacity=None, stream=None):
    """grids: float32 CUDA tensor the songs' offsets index into; dsongs: ndarray[DECODE_SONG_DTYPE].
    Returns (results ndarray[DECODE_RESULT_DTYPE], messages ndarray[OUT_MSG_DTYPE]); song i's messages are
    messages[msg_offset : msg_offset + n_messages].  Grows the message buffer once if it was too small."""
    import torch
    L = require_gpu()
    dsongs = np.ascontiguousarray(dsongs, dtype=abi.DECODE_SONG_DTYPE).reshape(-1)
    S = len(dsongs)
    n_chunks = decode_chunks(dsongs)
    dev = grids.device
    d_songs = torch.from_numpy(dsongs.view(np.uint8).reshape(-1)).to(dev)
    d_res = torch.empty(S * abi.DECODE_RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    cap = int(msg_capacity) if msg_capacity else max(1024, n_chunks * 16)
    s = torch.cuda.current_stream(dev) if stream is None else stream
    while True:
        d_msgs = torch.empty(cap * abi.OUT_MSG_DTYPE.itemsize, dtype=torch.uint8, device=dev)
        wsb = L.mst_decode_workspace_bytes(S, n_chunks, cap)
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        check(L.mst_decode_batch_device(grids.data_ptr(), d_songs.data_ptr(), S, n_chunks, d_msgs.data_ptr(), cap,
                                        d_res.data_ptr(), ws.data_ptr(), wsb, s.cuda_stream))
        res = d_res.cpu().numpy().view(abi.DECODE_RESULT_DTYPE).copy()
        if (res['status'] == abi.DECODE_CAPACITY).any():
            cap = int(res['n_messages'].astype(np.int64).sum())
            continue
        break
    msgs = d_msgs.cpu().numpy().view(abi.OUT_MSG_DTYPE)
    return res, msgs


def smf_write(dsong, msgs, eot_delta):
    """Serialise one decoded song to Standard MIDI File bytes (mid.save, style_transfer.py:98)."""
    L = load()
    d = np.ascontiguousarray(dsong.reshape(1))
    msgs = np.ascontiguousarray(msgs, dtype=abi.OUT_MSG_DTYPE)
    cap = 64 + 8 * abi.MAX_DECODE_CHANNELS + 8 * len(msgs)
    out = np.empty(cap, dtype=np.uint8)
    need = ctypes.c_size_t(0)
    n = L.mst_smf_write(_np_ptr(d), _np_ptr(msgs), len(msgs), int(eot_delta), _np_ptr(out), cap, ctypes.byref(need))
    if n < 0:
        raise MstError(f'mst_smf_write failed ({n})')
    return out[:n].tobytes()


# ---------------------------------------------------------------------------------- style transfer glue
def transfer_select(instr_logits, mode_logits, bpm, n_instruments, programs, groups, max_channels, d_songs, stream=None):
    """apply_style's host decisions on the device (mst_transfer_select).  instr_logits (S,41), mode_logits (S,2), bpm (S):
    float32 CUDA tensors; n_instruments: list of ints; programs / groups: the 40 sorted included programs and the group
    one-hot index of each; d_songs: uint8 CUDA tensor holding S MstDecodeSong records (updated in place).
    Returns (features (S, max_channels, 51) float32 CUDA, choices uint8 CUDA tensor of S MstTransferChoice records)."""
    import torch
    L = require_gpu()
    S = int(instr_logits.shape[0])
    dev = instr_logits.device
    n_dev = torch.tensor([int(x) for x in n_instruments], dtype=torch.int32, device=dev)
    feats = torch.empty(S, int(max_channels), 51, dtype=torch.float32, device=dev)
    choice = torch.empty(S * abi.TRANSFER_CHOICE_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    prog = np.ascontiguousarray(programs, dtype=np.int16)
    grp = np.ascontiguousarray(groups, dtype=np.int16)
    assert len(prog) == 40 and len(grp) == 40
    s = torch.cuda.current_stream(dev) if stream is None else stream
    check(L.mst_transfer_select(instr_logits.contiguous().data_ptr(), mode_logits.contiguous().data_ptr(), bpm.contiguous().data_ptr(),
                                n_dev.data_ptr(), S, _np_ptr(prog), _np_ptr(grp), int(max_channels), feats.data_ptr(),
                                d_songs.data_ptr(), choice.data_ptr(), s.cuda_stream))
    return feats, choice
