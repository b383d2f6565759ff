"""Single-slot latency of the raw-SMF encode path against the packed-event path on one synthetic batch (development probe)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))
import numpy as np
import torch
from mst_native import abi, lib, synth

n_songs = int(sys.argv[1]) if len(sys.argv) > 1 else 512
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
L = lib.require_gpu()
events, songs = synth.synth_corpus(n_songs, seed=1234)
filt = abi.make_filter(synth.POPULAR, 100).reshape(1)
batch = lib.DeviceBatch(events, songs)
batch.encode(filt, None)
floats = sum(lib.grid_floats(r) for r in batch.infos() if r['status'] in (0, 8))
arena = torch.empty(floats, dtype=torch.float32, device='cuda')
blobs = []
for srec in songs:
    o, k = int(srec['event_offset']), int(srec['n_events'])
    blobs.append(lib.smf_write_events(events[o:o + k], int(srec['ticks_per_beat'])))
buf, files = lib.concat_files(blobs)
h_bytes = torch.from_numpy(buf).pin_memory()
h_files = torch.from_numpy(files.view(np.uint8).reshape(-1)).pin_memory()
h_events = torch.from_numpy(lib.events_pack(events).view(np.uint8)).pin_memory()
h_songs = torch.from_numpy(songs.view(np.uint8).reshape(-1)).pin_memory()
info = torch.empty(n_songs * abi.SONG_INFO_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
wsb = max(L.mst_encode_packed_workspace_bytes(len(events), n_songs, arena.numel()), L.mst_encode_smf_workspace_bytes(len(buf), n_songs, arena.numel()))
ws = torch.empty(wsb, dtype=torch.uint8, device='cuda')
st = torch.cuda.Stream()
print(f'{n_songs} songs, {len(events)} events, {len(buf)} file bytes ({len(buf) / len(events):.2f} B/event), arena {floats * 4 / 1e9:.2f} GB', flush=True)


def smf():
    lib.check(L.mst_encode_batch_smf_async(h_bytes.data_ptr(), len(buf), h_files.data_ptr(), n_songs, filt.ctypes.data, info.data_ptr(),
                                           arena.data_ptr(), arena.numel(), ws.data_ptr(), wsb, st.cuda_stream))


def packed():
    lib.check(L.mst_encode_batch_packed_async(h_events.data_ptr(), len(events), h_songs.data_ptr(), n_songs, filt.ctypes.data,
                                              info.data_ptr(), arena.data_ptr(), arena.numel(), ws.data_ptr(), wsb, st.cuda_stream))


for name, fn in (('smf', smf), ('packed', packed)):
    fn(); st.synchronize()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); st.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    ok = int((info.numpy().view(abi.SONG_INFO_DTYPE)['status'] == 0).sum())
    print(f'{name}: single-slot latency {min(ts):.3f} ms (median {sorted(ts)[len(ts) // 2]:.3f}), {ok} songs ok', flush=True)
