"""style/style_transfer.py of the reference (style_transfer.py:22-158): encode two songs, swap the style vector,
predict instruments / mode / bpm, decode and save.  Everything between the MIDI bytes and the finished files runs on
the GPU: the encoder, the network, apply_style's own decisions (which instruments, which mode, which tempo:
mst_transfer_select) and the decoder; `transfer_style` renders the composition in ALL requested styles as one batched
device pass with a single read-back of the finished messages.  decode_midi hands a device tensor
straight to the decode kernels (no .cpu().numpy() of the grid) and returns the file the native SMF writer produced."""
import os

import numpy as np
import torch

from mst_native import abi, lib as _lib
from style.data import (
    included_instruments,
    get_input,
    prepare_input,
    percussion_id,
    instruments_one_hot_encoder,
    groups_one_hot_encoder,
    encode_instruments,
)
from style.midi import load_midi_from_file, program2group, DecodedMidi
from style.midi_conversion import ChannelConverter, read_midi
from style.model import device
from style.scales import major_mode, minor_mode, MODES
from style.utils.misc import make_dirs


def bpm2tempo(bpm):
    """mido.bpm2tempo: microseconds per beat, rounded."""
    return int(round(60 * 1e6 / bpm))


def transfer_style(model, composition_path, style_paths, output_path):
    """style_transfer.py:22-54.  Writes original/<song>.mid for the composition and every style, `<composition>
    (reconstructed).mid` and `<composition> (<style> style).mid` per style: the same files in the same places; the
    renderings are produced together by apply_styles()."""
    composition_name = os.path.splitext(os.path.basename(composition_path))[0]
    composition_input = get_model_input(composition_path)
    (_, (composition_info, composition_pitched_channels, _, composition_instruments,
         composition_unpitched_channels)) = composition_input
    composition_cc = ChannelConverter(composition_info)
    style_, melody, rhythm = extract_style(model, composition_input)

    output_path = os.path.join(output_path, composition_name)
    # style_transfer.py:31-34 tries shutil.rmtree(output_path) but never imports shutil and swallows the NameError:
    # the directory is in effect never removed.  Same here (files are overwritten one by one).

    save(
        composition_cc, composition_pitched_channels, composition_unpitched_channels,
        composition_instruments, os.path.join(output_path, f'original/{composition_name}.mid'))
    jobs = [(composition_info, style_, len(composition_instruments),
             os.path.join(output_path, f'{composition_name} (reconstructed).mid'))]
    for style_path in style_paths:
        style_name = os.path.splitext(os.path.basename(style_path))[0]
        style_input = get_model_input(style_path)
        _, (style_info, style_pitched_channels, _, style_instruments, style_unpitched_channels) =\
            style_input
        style_cc = ChannelConverter(style_info)
        style, _, _ = extract_style(model, style_input)

        save(style_cc, style_pitched_channels, style_unpitched_channels, style_instruments,
             os.path.join(output_path, f'original/{style_name}.mid'))
        info = combine_info(style_info=style_info, melody_info=composition_info)
        jobs.append((info, style, len(style_instruments),
                     os.path.join(output_path, f'{composition_name} ({style_name} style).mid')))
    apply_styles(model, jobs, melody, rhythm)


def get_model_input(path):
    mid = load_midi_from_file(path)
    if mid is None:
        return None
    channels, info = read_midi(mid)
    channels = channels.filtered(included_instruments, min_n_messages=0)     # style_transfer.py:62: no length filter
    input = get_input(channels, info)
    return path, input


def extract_style(model, input):
    pitched_channels = input[1][1]
    max_n_bars = 1000 // pitched_channels.shape[0]
    mode, bpm, pitched_channels, instruments_features, unpitched_channels = prepare_input(
        input, max_n_bars)
    with torch.no_grad():
        style, melody, rhythm = model.extract_style(
            mode, bpm, pitched_channels, instruments_features, unpitched_channels)
    return style.detach(), melody.detach(), rhythm.detach()


def save(cc, pitched_channels, unpitched_channels, instruments, save_path):
    channels_info = [{
        'channel_id': i,
    } for i in range(16) if i != 9][:pitched_channels.shape[1]]     # sic (:79): n_bars for a 6-d grid; decode_midi zips
    for instrument_id, channel_info in zip(instruments, channels_info):
        channel_info['instrument_id'] = instrument_id
    unpitched_channel_info = {
        'channel_id': 9,
        'instrument_id': -1,
    }

    make_dirs(os.path.dirname(save_path))

    if len(pitched_channels.shape) == 6:
        pitched_channels = torch.as_tensor(pitched_channels, dtype=torch.float).unsqueeze(0).to(device)
        if unpitched_channels is not None:
            unpitched_channels = torch.as_tensor(
                unpitched_channels, dtype=torch.float).unsqueeze(0).to(device)

    mid = decode_midi(
        cc, channels_info, pitched_channels, unpitched_channel_info, unpitched_channels)
    mid.save(save_path)


# the one-hot encoders' category tables as the select kernel takes them: sorted included programs, and for each the
# index of its group among the sorted group names (style/data.py:23-27)
_PROGRAMS = [int(p) for p in instruments_one_hot_encoder.categories_[0]]
_GROUPS = [list(groups_one_hot_encoder.categories_[0]).index(program2group[p]) for p in _PROGRAMS]


def apply_styles(model, jobs, melody, rhythm):
    """Device-side apply_style (style_transfer.py:101-131) for a batch: jobs = [(info, style (1,256), n_instruments,
    save_path)], all rendered on the composition's melody / rhythm.

      1. predict_song_info for all S styles at once (mst_model_song_info, N = S);
      2. mst_transfer_select: top-k instruments with the percussion special case, mode arg-max, bpm2tempo(round(bpm)),
         the instrument one-hot features and the decode-table entries - on the device, nothing is read back;
      3. model.apply_style per style into one grid buffer (mst_model_predict_apply; the number of channels a style
         renders is known up front: n_instruments, of which the kernel's choice may drop the percussion slot);
      4. one mst_decode_batch_device over all S renderings, ONE device->host copy of results + messages + choices;
      5. the native SMF writer per file.
    Like the reference, every job's info dict gets info['tempo'] and info['scale']['mode'] of the prediction."""
    from mst_native import model as _mm
    S = len(jobs)
    if S == 0:
        return
    model._flatten()
    f = lambda t: t.detach().to('cuda', torch.float32).contiguous()
    melody, rhythm = f(melody), f(rhythm)
    _, B, nb = rhythm.shape[:3]
    styles = torch.cat([f(j[1]).reshape(1, 256) for j in jobs])
    n_inst = [int(j[2]) for j in jobs]
    with torch.no_grad():
        # 1. the three heads, all styles at once (the composition's rhythm is shared)
        ws = _mm.Workspace(S, 1, B, nb, False, training=False)
        heads = ws.song_info(model._flat, styles, rhythm.expand(S, *rhythm.shape[1:]).contiguous())
        # 2. the decisions, on the device.  A style renders at most n (+1 for the lone-percussion case) channels; the
        # decode table's grid offsets are fixed by that bound, so no decision has to come back to the host first
        max_c = [max(1, n) for n in n_inst]
        pitched_floats = [c * B * nb * abi.PITCHED_BEAT_FLOATS for c in max_c]
        unpitched_floats = B * nb * abi.UNPITCHED_BEAT_FLOATS
        offsets, total = [], 0
        for pf in pitched_floats:
            offsets.append(total)
            total += pf + unpitched_floats
        recs = []
        for (info, _, _, _), c, off, pf in zip(jobs, max_c, offsets, pitched_floats):
            cc = ChannelConverter(info)
            rec = cc.decode_song_record(c, True, B, nb, [i for i in range(16) if i != 9][:c], [0] * min(c, 15), write_back=False)
            rec['pitched_offset'], rec['unpitched_offset'] = off, off + pf
            recs.append(rec)
        dsongs = np.array(recs, dtype=abi.DECODE_SONG_DTYPE)
        d_songs = torch.from_numpy(dsongs.view(np.uint8).reshape(-1)).cuda()
        feats, d_choice = _lib.transfer_select(heads['instruments'], heads['mode'], heads['bpm'], n_inst, _PROGRAMS, _GROUPS,
                                                max(max_c), d_songs)
        # 3. the renderings, one per style, into one grid buffer
        grids = torch.empty(total, dtype=torch.float32, device='cuda')
        for s, (c, off, pf) in enumerate(zip(max_c, offsets, pitched_floats)):
            wa = model._workspace(c, B, nb, True, False)
            out = wa.predict_apply(model._flat, styles[s:s + 1], melody, rhythm, feats[s:s + 1, :c].contiguous())
            grids[off:off + pf].copy_(out['pitched'].reshape(-1))
            grids[off + pf:off + pf + unpitched_floats].copy_(out['unpitched'].reshape(-1))
        # 4. decode everything; the chunk count is an upper bound (the kernel's choices may drop channels)
        n_chunks = sum((min(c, 15) + 1) * B for c in max_c)
        cap = 2 * sum((min(c, 15) * 56 + 47) * B * nb * 10 for c in max_c) + 1024
        L = _lib.require_gpu()
        d_res = torch.empty(S * abi.DECODE_RESULT_DTYPE.itemsize, dtype=torch.uint8, device='cuda')
        d_msgs = torch.empty(cap * abi.OUT_MSG_DTYPE.itemsize, dtype=torch.uint8, device='cuda')
        wsb = L.mst_decode_workspace_bytes(S, n_chunks, cap)
        dws = torch.empty(wsb, dtype=torch.uint8, device='cuda')
        _lib.check(L.mst_decode_batch_device(grids.data_ptr(), d_songs.data_ptr(), S, n_chunks, d_msgs.data_ptr(), cap, d_res.data_ptr(),
                                             dws.data_ptr(), wsb, torch.cuda.current_stream().cuda_stream))
        # the single host round trip of the whole transfer
        res = d_res.cpu().numpy().view(abi.DECODE_RESULT_DTYPE)
        songs_out = d_songs.cpu().numpy().view(abi.DECODE_SONG_DTYPE)
        choices = d_choice.cpu().numpy().view(abi.TRANSFER_CHOICE_DTYPE)
        n_used = int(res['msg_offset'][-1]) + int(res['n_messages'][-1]) if int(res['status'][-1]) == abi.DECODE_OK else cap
        msgs = d_msgs[:min(cap, n_used) * abi.OUT_MSG_DTYPE.itemsize].cpu().numpy().view(abi.OUT_MSG_DTYPE)
    # 5. files + the reference's side effects on the info dicts
    for s, (info, _, _, save_path) in enumerate(jobs):
        info['tempo'] = int(choices[s]['tempo'])
        info['scale']['mode'] = MODES[int(choices[s]['mode'])]
        r = res[s]
        _raise_for_decode_status(int(r['status']))
        m = msgs[int(r['msg_offset']):int(r['msg_offset']) + int(r['n_messages'])]
        data = _lib.smf_write(songs_out[s], m, int(r['eot_delta']))
        make_dirs(os.path.dirname(save_path))
        DecodedMidi(m.copy(), data, int(songs_out[s]['ticks_per_beat'])).save(save_path)
    return choices


def apply_style(model, info, style, melody, rhythm, n_instruments, save_path):
    """style_transfer.py:101-131 for one style: the batch of one of apply_styles()."""
    apply_styles(model, [(info, style, n_instruments, save_path)], melody, rhythm)


def combine_info(style_info, melody_info):
    info = {
        'time_signature': melody_info['time_signature'],
        'scale': style_info['scale'],
        'ticks_per_beat': melody_info['ticks_per_beat'],
        'ticks_per_bar': melody_info['ticks_per_bar'],
        'tempo': style_info['tempo'],
    }
    return info


def _raise_for_decode_status(status):
    if status == abi.DECODE_NO_MESSAGES:
        raise IndexError('list index out of range')               # messages[-1] in create_midi (midi.py:158)
    if status == abi.DECODE_VELOCITY:
        raise AssertionError('velocity out of range')             # the assert at midi.py:148
    if status != abi.DECODE_OK:
        raise ValueError(f'cannot decode grid (status {status})')


def decode_midi(channel_converter, channels_info, pitched_channels, unpitched_channel_info=None,
                unpitched_channels=None):
    """hard_output + vchannel2channel per channel + create_midi(max_delta_time=1) (style_transfer.py:145-158) in one
    batched GPU call.  Like the reference, the velocity gate is written back into `pitched_channels` in place."""
    pitched = pitched_channels[0] if pitched_channels.dim() == 7 else pitched_channels
    n_pitched = min(len(channels_info), pitched.shape[0])          # zip(channels_info, vchannels)
    _, n_bars, n_beats = pitched.shape[:3]
    has_u = unpitched_channels is not None
    p = pitched.detach().to('cuda', torch.float32)
    parts = [p.reshape(-1)]
    if has_u:
        u = unpitched_channels.detach().to('cuda', torch.float32)
        parts.append(u.reshape(-1))
    contiguous_in_place = (not has_u) and p.is_contiguous() and p.data_ptr() == pitched.data_ptr()
    grids = parts[0] if contiguous_in_place else torch.cat(parts)
    rec = channel_converter.decode_song_record(
        n_pitched, has_u, n_bars, n_beats, [c['channel_id'] for c in channels_info[:n_pitched]],
        [c.get('instrument_id', 0) for c in channels_info[:n_pitched]])
    rec['pitched_offset'] = 0
    rec['unpitched_offset'] = p.numel()
    if has_u and unpitched_channel_info is not None:
        rec['unpitched_channel_id'] = unpitched_channel_info['channel_id']
    dsongs = np.array([rec], dtype=abi.DECODE_SONG_DTYPE)
    res, msgs = _lib.decode_batch(grids, dsongs)
    r = res[0]
    _raise_for_decode_status(int(r['status']))
    if not contiguous_in_place:                                    # hard_output's in-place gate on the caller's tensors
        pitched.copy_(grids[:p.numel()].view_as(pitched))
        if has_u:
            unpitched_channels.copy_(grids[p.numel():].view_as(unpitched_channels))
    m = msgs[int(r['msg_offset']):int(r['msg_offset']) + int(r['n_messages'])]
    data = _lib.smf_write(dsongs[0], m, int(r['eot_delta']))
    return DecodedMidi(m.copy(), data, int(rec['ticks_per_beat']))
