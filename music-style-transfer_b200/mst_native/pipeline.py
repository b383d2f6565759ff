"""File -> grid pipeline of the encoder: what `load_midi_from_file` + the prefetch thread + `get_input`
do one file at a time in the reference (style/midi.py:104-108, style/utils/parallel.py:34-51, style/data.py:51-100), as a
batched, overlapped stream:

    paths --(mst_files_read_batch: C++ thread pool, GIL released)--> the files' raw bytes in a page-locked buffer
          --(mst_encode_batch_smf_async on a rotating slot = stream + workspace + arena + info buffer:
             H2D of the bytes, parse ON THE DEVICE (thread per track), encode)--> grids in HBM

(`wire='events'` keeps the round-2a path: mst_smf_load_batch parses on the host's cores into page-locked MstEvent records
and mst_encode_batch_host_async uploads those - 4-5x the bytes, and the host parse bounds it.)

While the GPU encodes batch i, a helper thread parses batch i + 1 and the H2D copy of batch i + 1 is queued behind the
kernels of batch i on another slot's stream.  Nothing is allocated per batch (buffers grow geometrically and are kept), and
the only host synchronisation per batch is the wait for its own slot before its song infos are read.
"""
import concurrent.futures
import ctypes

import numpy as np

from . import abi, lib as _lib


class _HostSlot:
    """Page-locked staging of one batch of files."""

    def __init__(self, torch, n_files, n_events):
        self.torch = torch
        self.n_files = n_files
        self.songs_t = torch.empty(n_files * abi.SONG_IN_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
        self.files_t = torch.empty(n_files * abi.FILE_IN_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
        self.info_t = torch.empty(n_files * abi.SONG_INFO_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
        self.status = np.zeros(n_files, dtype=np.int32)
        self.events_t = None
        self.bytes_t = None
        self.n_bytes = 0
        self.reserve(n_events)

    def reserve_bytes(self, n_bytes):
        if self.bytes_t is None or self.bytes_t.numel() < n_bytes:
            self.bytes_t = self.torch.empty(int(n_bytes * 1.25) + 4096, dtype=self.torch.uint8).pin_memory()

    @property
    def files(self):
        return self.files_t.numpy().view(abi.FILE_IN_DTYPE)

    def reserve(self, n_events):
        if self.events_t is None or self.events_t.numel() < n_events * abi.EVENT_DTYPE.itemsize:
            self.events_t = self.torch.empty(int(n_events * 1.25) * abi.EVENT_DTYPE.itemsize, dtype=self.torch.uint8).pin_memory()

    @property
    def events(self):
        return self.events_t.numpy().view(abi.EVENT_DTYPE)

    @property
    def songs(self):
        return self.songs_t.numpy().view(abi.SONG_IN_DTYPE)

    @property
    def infos(self):
        return self.info_t.numpy().view(abi.SONG_INFO_DTYPE)


class EventPipeline:
    """Encodes MIDI files in batches; `run(paths)` yields (path, MstSongInfo record, arena tensor) per file in file order
    (record['status'] tells whether the file loaded / encoded; the grids of an OK song are views of `arena`, valid until
    `slots - 1` more batches have been produced: consume or clone them before that)."""

    def __init__(self, filt, batch_files=16, slots=3, arena_floats=256 << 20, n_threads=0, device=None, wire='smf'):
        import torch
        assert wire in ('smf', 'events')
        self.wire = wire
        self.torch = torch
        self.L = _lib.require_gpu()
        self.filt = np.ascontiguousarray(filt).reshape(1)
        self.batch_files, self.n_threads = int(batch_files), int(n_threads)
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        self.arena_floats = int(arena_floats)
        self.host = [_HostSlot(torch, self.batch_files, 1 << 16) for _ in range(slots)]
        self.dev = [{'stream': torch.cuda.Stream(self.device), 'ws': None,
                     'arena': torch.empty(self.arena_floats, dtype=torch.float32, device=self.device)} for _ in range(slots)]
        self.pool = concurrent.futures.ThreadPoolExecutor(max_workers=1)
        self.files_loaded = self.files_failed = 0

    # ---- stage 1: files -> page-locked events (helper thread; the C call releases the GIL)
    def _load(self, paths, h):
        if self.wire == 'smf':
            h.reserve_bytes(1 << 20)
            while True:
                try:
                    h.n_bytes, _, _ = _lib.files_read_batch(paths, h.bytes_t.numpy(), h.files, h.status, self.n_threads)
                    return h.n_bytes
                except _lib.MstError as e:
                    if not hasattr(e, 'needed'):
                        raise
                    h.reserve_bytes(e.needed)
        while True:
            try:
                total, _, _ = _lib.smf_load_batch(paths, h.events, h.songs, h.status, self.n_threads)
                return total
            except _lib.MstError as e:
                if not hasattr(e, 'needed'):
                    raise
                h.reserve(e.needed)

    # ---- stage 2: events -> grids (asynchronous on the slot's stream)
    def _encode(self, n_files, n_events, h, d):
        if self.wire == 'smf':
            n_bytes = n_events                         # _load returned the byte count
            need = self.L.mst_encode_smf_workspace_bytes(n_bytes, n_files, d['arena'].numel())
            if d['ws'] is None or d['ws'].numel() < need:
                d['ws'] = None
                d['ws'] = self.torch.empty(int(need * 1.25), dtype=self.torch.uint8, device=self.device)
            _lib.check(self.L.mst_encode_batch_smf_async(
                h.bytes_t.data_ptr(), n_bytes, h.files_t.data_ptr(), n_files, self.filt.ctypes.data, h.info_t.data_ptr(),
                d['arena'].data_ptr(), d['arena'].numel(), d['ws'].data_ptr(), d['ws'].numel(), d['stream'].cuda_stream))
            return
        need = self.L.mst_encode_host_workspace_bytes(n_events, n_files, d['arena'].numel())
        if d['ws'] is None or d['ws'].numel() < need:
            d['ws'] = self.torch.empty(int(need * 1.25), dtype=self.torch.uint8, device=self.device)
        _lib.check(self.L.mst_encode_batch_host_async(
            h.events_t.data_ptr(), n_events, h.songs_t.data_ptr(), n_files, self.filt.ctypes.data, h.info_t.data_ptr(),
            d['arena'].data_ptr(), d['arena'].numel(), d['ws'].data_ptr(), d['ws'].numel(), d['stream'].cuda_stream))

    def _drain(self, job):
        paths, h, d = job
        d['stream'].synchronize()
        infos = h.infos[:len(paths)].copy()
        for i, p in enumerate(paths):
            r = infos[i]
            if h.status[i] != 0:
                r['status'] = abi.SONG_LOAD_FAILED
            if int(r['status']) == abi.SONG_LOAD_FAILED:
                self.files_failed += 1
            self.files_loaded += 1
            if int(r['status']) == abi.SONG_ARENA_FULL:
                # rare: this batch's grids outgrew the slot's arena; encode the song on its own into a fitting arena
                if self.wire == 'smf':
                    o, k = int(h.files[i]['byte_offset']), int(h.files[i]['n_bytes'])
                    one, arena = _lib.encode_smf([h.bytes_t.numpy()[o:o + k].tobytes()], self.filt[0])
                else:
                    ev = h.events[int(h.songs[i]['event_offset']):int(h.songs[i]['event_offset']) + int(h.songs[i]['n_events'])].copy()
                    one, arena = _lib.encode_songs([ev], [int(h.songs[i]['ticks_per_beat'])], self.filt[0])
                yield p, one[0], arena
            else:
                yield p, r, d['arena']

    def run(self, paths):
        paths = list(paths)
        batches = [paths[i:i + self.batch_files] for i in range(0, len(paths), self.batch_files)]
        if not batches:
            return
        S = len(self.host)
        fut = self.pool.submit(self._load, batches[0], self.host[0])
        pending = []
        for i, batch in enumerate(batches):
            n_events = fut.result()
            h, d = self.host[i % S], self.dev[i % S]
            if i + 1 < len(batches):
                # the next batch's host slot was drained at least one iteration ago (S >= 2 slots rotate)
                fut = self.pool.submit(self._load, batches[i + 1], self.host[(i + 1) % S])
            self._encode(len(batch), n_events, h, d)
            pending.append((batch, h, d))
            if len(pending) >= S - 1:
                yield from self._drain(pending.pop(0))
        while pending:
            yield from self._drain(pending.pop(0))

    def close(self):
        self.pool.shutdown(wait=True)
