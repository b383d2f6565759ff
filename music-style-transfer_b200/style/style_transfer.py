"""style/style_transfer.py of the reference (style_transfer.py:22-158): encode two songs, swap the style vector,
predict instruments / mode / bpm, decode and save.  Encode, the model and decode run on the GPU; decode_midi hands
the model's device tensor straight to the decode kernels (no .cpu().numpy() of the grid) and returns the finished
file produced by the native SMF writer."""
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
    encode_instruments,
)
from style.midi import load_midi_from_file, DecodedMidi
from style.midi_conversion import ChannelConverter, read_midi
from style.model import device
from style.scales import major_mode, minor_mode
from style.utils.misc import make_dirs


def bpm2tempo(bpm):
    """mido.bpm2tempo: microseconds per beat, rounded."""
    return int(round(60 * 1e6 / bpm))


def transfer_style(model, composition_path, style_paths, output_path):
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
    apply_style(model, composition_info, style_, melody, rhythm, len(composition_instruments),
                os.path.join(output_path, f'{composition_name} (reconstructed).mid'))
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

        apply_style(model, info, style, melody, rhythm, len(style_instruments),
                    os.path.join(output_path, f'{composition_name} ({style_name} style).mid'))


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


def apply_style(model, info, style, melody, rhythm, n_instruments, save_path):
    with torch.no_grad():
        instruments_pred, mode, bpm = model.predict_song_info(style, rhythm)
    info['tempo'] = bpm2tempo(round(float(bpm)))

    instruments_pred = instruments_pred.detach().cpu().numpy()
    instruments = np.argsort(-instruments_pred[0])[:n_instruments]
    if len(instruments) == 1 and instruments[0] == [percussion_id]:
        instruments = np.argsort(-instruments_pred[0])[:n_instruments+1]
    unpitched = percussion_id in instruments
    instruments = [instrument for instrument in instruments if instrument != percussion_id]

    instruments_encoded = np.zeros([len(instruments), len(included_instruments)])
    for i, instrument in enumerate(instruments):
        instruments_encoded[i, instrument] = 1
    instruments = instruments_one_hot_encoder.inverse_transform(instruments_encoded)
    instruments = [int(x[0]) for x in instruments]

    if mode[0].argmax() == 0:
        mode = major_mode
    else:
        mode = minor_mode
    info['scale']['mode'] = mode

    cc = ChannelConverter(info)
    instruments_features = encode_instruments(instruments)
    instruments_features = torch.tensor(
        instruments_features, dtype=torch.float).to(device).unsqueeze(0)
    with torch.no_grad():
        pitched_pred, unpitched_pred = model.apply_style(
            style, melody, rhythm, instruments_features, unpitched)

    save(cc, pitched_pred, unpitched_pred, instruments, save_path)


def combine_info(style_info, melody_info):
    info = {
        'time_signature': melody_info['time_signature'],
        'scale': style_info['scale'],
        'ticks_per_beat': melody_info['ticks_per_beat'],
        'ticks_per_bar': melody_info['ticks_per_bar'],
        'tempo': style_info['tempo'],
    }
    return info


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
    status = int(r['status'])
    if status == abi.DECODE_NO_MESSAGES:
        raise IndexError('list index out of range')               # messages[-1] in create_midi (midi.py:158)
    if status == abi.DECODE_VELOCITY:
        raise AssertionError('velocity out of range')             # the assert at midi.py:148
    if status != abi.DECODE_OK:
        raise ValueError(f'cannot decode grid (status {status})')
    if not contiguous_in_place:                                    # hard_output's in-place gate on the caller's tensors
        pitched.copy_(grids[:p.numel()].view_as(pitched))
        if has_u:
            unpitched_channels.copy_(grids[p.numel():].view_as(unpitched_channels))
    m = msgs[int(r['msg_offset']):int(r['msg_offset']) + int(r['n_messages'])]
    data = _lib.smf_write(dsongs[0], m, int(r['eot_delta']))
    return DecodedMidi(m.copy(), data, int(rec['ticks_per_beat']))
