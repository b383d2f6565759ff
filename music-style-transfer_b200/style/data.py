"""style/data.py of the reference on top of the CUDA encoder: same names, signatures and return structures
(data.py:19-169).  A song's grids are float32 CUDA tensors (views into the encoder's arena) instead of float64
numpy arrays; everything train-model.py:97-111 does with them (`.shape[0]`, `.sum()`, `prepare_input`) works unchanged.

iter_inputs encodes `encode_batch` files per GPU call (the reference encodes one file at a time in Python); the yield
order and the per-file skip rules (unreadable file, MidiFormatError, no pitched channel) are the reference's."""
import numpy as np
import torch

from mst_native import abi, lib as _lib
from style.exceptions import MidiFormatError
from style.midi import load_midi_from_file, program2group, popular_instruments
from style.midi_conversion import Channels, read_midi, info_from_record, scale_from_record, raise_for_status
from style.scales import major_mode
from style.model import device

included_instruments = popular_instruments
instrument_groups = [program2group[p] for p in included_instruments]
n_instruments = len(included_instruments) + 1  # also percussion

# OneHotEncoder(categories='auto') orders categories by sorted value (data.py:23-27)
_sorted_instruments = sorted(included_instruments)
_sorted_groups = sorted(set(instrument_groups))
instrument_size = len(_sorted_groups) + len(_sorted_instruments)
percussion_id = len(included_instruments)


class _OneHot:
    """The two sklearn OneHotEncoder objects of the reference, reduced to what callers use (transform /
    inverse_transform / categories_)."""

    def __init__(self, categories):
        self.categories_ = [np.array(categories)]
        self._index = {c: i for i, c in enumerate(categories)}

    def transform(self, x):
        x = np.asarray(x).reshape(-1)
        out = np.zeros((len(x), len(self._index)))
        for i, v in enumerate(x):
            out[i, self._index[v.item() if hasattr(v, 'item') else v]] = 1.0
        return out

    def inverse_transform(self, onehot):
        return np.array([[self.categories_[0][int(np.argmax(row))]] for row in np.asarray(onehot)])


instruments_one_hot_encoder = _OneHot(_sorted_instruments)
groups_one_hot_encoder = _OneHot(_sorted_groups)


def iter_all_midis(files, shuffle=False, looped=False):
    if shuffle:
        files = files[:]
        np.random.shuffle(files)
    if looped:
        while True:
            yield from iter_all_midis(files, looped=False)
    for file in files:
        mid = load_midi_from_file(file)
        if mid is not None:
            try:
                channels, info = read_midi(mid)
                yield file, channels, info
            except MidiFormatError:
                continue


def _encode_many(items, included, min_n_messages):
    """items: [(events, ticks_per_beat)] -> (infos, arena) through one mst_encode_batch_device plan + fill."""
    filt = abi.make_filter(included, min_n_messages)
    return _lib.encode_songs([e for e, _ in items], [t for _, t in items], filt)


def _input_from_record(r, arena):
    """One MstSongInfo record -> the tuple get_input returns (data.py:100)."""
    info = info_from_record(r)
    info['scale'] = scale_from_record(r)
    pitched, unpitched = _lib.song_grids(r, arena)
    instruments = [int(x) for x in r['pitched_instruments'][:int(r['n_pitched'])]]
    return info, pitched, encode_instruments(instruments), instruments, unpitched


def iter_inputs(files, instruments, min_n_messages=100, shuffle=False, looped=False, encode_batch=16):
    """Yields (filename, (info, pitched, instruments_features, instruments, unpitched)) like data.py:51-63.

    The files go through mst_native.pipeline.EventPipeline: a C++ thread pool parses `encode_batch` files at a time into
    page-locked buffers while the GPU encodes the previous batch (the job of the reference's prefetch thread,
    utils/parallel.py:34-51, and of its per-file Python encoder).  Order and skip rules are the reference's: unreadable
    files, MidiFormatError rejects and songs without a pitched channel are skipped (data.py:42-47,57)."""
    from mst_native.pipeline import EventPipeline
    order = list(files)
    if shuffle:
        np.random.shuffle(order)       # once: the reference shuffles a copy, then loops that permutation (data.py:35-41)
    pipe = EventPipeline(abi.make_filter(instruments, min_n_messages), batch_files=encode_batch)
    try:
        while True:
            for f, r, arena in pipe.run(order):
                if int(r['status']) != abi.SONG_OK:
                    continue
                info, pitched, feats, ins, unpitched = _input_from_record(r, arena)
                # the arena slot is rewritten two batches later: the caller gets its own copy of the grids
                yield f, (info, pitched.clone(), feats, ins, None if unpitched is None else unpitched.clone())
            if not looped:
                return
    finally:
        pipe.close()


def get_input(channels, info):
    """get_input(channels, info) (data.py:66-100): encodes ONE song; mutates info['scale'] like the reference (:86)."""
    if not isinstance(channels, Channels):
        raise TypeError('get_input needs the channel collection read_midi returned')
    infos, arena = _encode_many([(channels.events, channels.ticks_per_beat)], channels.included, channels.min_n_messages)
    r = infos[0]
    if int(r['status']) == abi.SONG_NO_PITCHED:
        raise ValueError('need at least one array to stack')      # np.stack([]) in the reference (:91)
    raise_for_status(r['status'])
    out = _input_from_record(r, arena)
    info.update({k: v for k, v in out[0].items() if k not in info or k == 'scale'})
    return (info,) + out[1:]


def encode_instruments(instruments):
    groups = [program2group[p] for p in instruments]
    x = np.concatenate([instruments_one_hot_encoder.transform(instruments), groups_one_hot_encoder.transform(groups)], 1) \
        if len(instruments) else np.zeros((0, instrument_size))
    return x


def prepare_input(input, max_n_bars=None):
    _, (info, pitched_channels, instruments_features, _, unpitched_channels) = input

    if max_n_bars is None:
        max_n_bars = pitched_channels.shape[1]

    pitched_channels = torch.as_tensor(pitched_channels[:, :max_n_bars], dtype=torch.float).to(device).unsqueeze(0).contiguous()
    instruments_features = torch.as_tensor(instruments_features, dtype=torch.float).to(device).unsqueeze(0)

    if unpitched_channels is not None:
        unpitched_channels = torch.as_tensor(
            unpitched_channels[:, :max_n_bars], dtype=torch.float).to(device).unsqueeze(0).contiguous()

    if info['scale']['mode'] == major_mode:
        mode = torch.tensor([1., 0.])
    else:
        mode = torch.tensor([0., 1.])
    mode = mode.unsqueeze(0).to(device)

    bpm = torch.tensor(info['bpm'], dtype=torch.float).unsqueeze(0).to(device)

    return mode, bpm, pitched_channels, instruments_features, unpitched_channels


def get_used_instruments(instruments_features, unpitched_channels):
    used_instruments = instruments_features[:, :, :len(included_instruments)]
    used_instruments = used_instruments.sum(1)
    used_instruments = (used_instruments > 0).float()

    percussion_used = torch.tensor(float(unpitched_channels is not None), dtype=torch.float, device=used_instruments.device)
    percussion_used = percussion_used.view(1, 1)

    return torch.cat([used_instruments, percussion_used], 1)

