"""Small end-to-end pass over every kernel family for compute-sanitizer (memcheck / racecheck / synccheck / initcheck):
encode (both sort paths, crop, host + packed entry points), decode, the model forward / training step in the fp32 mode
and in the bf16 mode (tcgen05 GEMMs, the TMA-fed GEMM, the opt-in tcgen05 tails), Adam, the transfer-select kernel.

    compute-sanitizer --tool memcheck python tools/sanitize_target.py
Shapes are tiny: the sanitizers serialise and instrument every access."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))
import numpy as np
import torch
from mst_native import abi, lib, model as mm, synth, trainer as tr


def main():
    L = lib.require_gpu()
    filt = abi.make_filter(synth.POPULAR, 0)
    # ---- encode: one-channel-per-track songs (fast path) and tracks sharing channels (full radix sort)
    events, songs = synth.synth_corpus(3, seed=3, n_bars=6, n_tracks=5)
    per = synth.split_songs(events, songs)
    shared = per[0].copy()
    shared['channel'][shared['kind'] <= abi.EV_PITCHWHEEL] %= 2
    per.append(shared)
    tpbs = [int(t) for t in songs['ticks_per_beat']] + [int(songs['ticks_per_beat'][0])]
    infos, arena = lib.encode_songs(per, tpbs, filt)
    assert (infos['status'] == 0).any()
    ev, sg = lib.make_songs(per, tpbs)
    n = arena.numel()
    wsb = L.mst_encode_packed_workspace_bytes(len(ev), len(sg), n)
    ws = torch.empty(wsb, dtype=torch.uint8, device='cuda')
    d_arena = torch.empty(n, dtype=torch.float32, device='cuda')
    h_info = np.zeros(len(sg), dtype=abi.SONG_INFO_DTYPE)
    packed = lib.events_pack(ev)
    lib.check(L.mst_encode_batch_packed_async(packed.ctypes.data, len(ev), sg.ctypes.data, len(sg), filt.reshape(1).ctypes.data,
                                              h_info.ctypes.data, d_arena.data_ptr(), n, ws.data_ptr(), wsb, None))
    torch.cuda.synchronize()
    assert torch.equal(d_arena, arena)
    # ---- decode of the first encoded song
    r = infos[0]
    C = min(int(r['n_pitched']), 15)
    d = abi.make_decode_song(r, C, int(r['n_unpitched']) > 0, int(r['n_bars_out']), instrument_ids=list(r['pitched_instruments'][:C]),
                             duration=int(r['duration']), write_back=True, pitched_offset=int(r['pitched_offset']),
                             unpitched_offset=int(r['unpitched_offset']))
    res, msgs = lib.decode_batch(arena, np.array([d], dtype=abi.DECODE_SONG_DTYPE))
    assert int(res[0]['status']) == 0 and int(res[0]['n_messages']) > 0
    # ---- model: training steps in both modes, the opt-in tcgen05 tails included
    sd = synth.make_state_dict(108)
    s = synth.synthetic_sample(seed=3, C=2, B=3, nb=4, drums=True, density=0.02)
    cu = lambda t: t.cuda().contiguous()
    for prec, env in (('fp32', {}), ('bf16', {}), ('bf16', {'MST_PSA_FWD_TC': '1', 'MST_PSA_BWD_TC': '1'})):
        os.environ.update(env)
        t = tr.Trainer(sd, precision=prec)
        losses = t.accumulate(cu(s['pitched']), cu(s['unpitched']), cu(s['feats']), cu(s['mode']), cu(s['bpm']), cu(s['used']),
                              torch.tensor([120.], device='cuda'))
        t.optimizer_step()
        torch.cuda.synchronize()
        assert torch.isfinite(losses).all(), (prec, env)
        for k in env:
            os.environ.pop(k)
    # ---- device-side apply_style decisions
    from style import style_transfer as st
    rec = abi.make_decode_song(dict(ticks_per_beat=480, ticks_per_bar=1920, tempo=500000, tonic=0, mode=0, numerator=4, denominator_log2=2,
                                    n_beats=4), 4, True, 2, write_back=False)
    d_songs = torch.from_numpy(np.array([rec, rec], dtype=abi.DECODE_SONG_DTYPE).view(np.uint8).reshape(-1)).cuda()
    lib.transfer_select(torch.randn(2, 41, device='cuda'), torch.randn(2, 2, device='cuda'), torch.full((2,), 120.5, device='cuda'),
                        [3, 1], st._PROGRAMS, st._GROUPS, 4, d_songs)
    torch.cuda.synchronize()
    print('sanitize target ok')


if __name__ == '__main__':
    main()
