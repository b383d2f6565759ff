"""style/midi_conversion.py of the reference, as a thin front end of the CUDA codec.

read_midi(mid) (midi_conversion.py:216-232) runs the encoder's planning pass (sort, state walk, song info) and returns
`(channels, info)`: `info` is the reference's dict; `channels` is an opaque collection carrying the events (the
reference's per-message NoteMessage lists are never built).  ChannelConverter (:349-609) keeps its constructor and is
the handle decode_midi needs; its per-note Python methods are replaced by the batched kernels behind
style.data.get_input (encode) and style.style_transfer.decode_midi (decode)."""

from mst_native import abi, lib as _lib
from style.exceptions import MidiFormatError
from style.midi import program2instrument
from style.scales import MODES, key_names


class Channels:
    """The (channel, instrument) message lists of one song, resident as an event buffer; filter settings travel with it
    (iter_inputs data.py:53-56 / get_model_input style_transfer.py:62 set them)."""

    def __init__(self, events, ticks_per_beat, n_lists, included=None, min_n_messages=0):
        self.events, self.ticks_per_beat, self.n_lists = events, ticks_per_beat, n_lists
        self.included, self.min_n_messages = included, min_n_messages

    def __len__(self):
        return self.n_lists

    def filtered(self, included, min_n_messages=0):
        return Channels(self.events, self.ticks_per_beat, self.n_lists, list(included), min_n_messages)

    def make_filter(self, crop_channel_bars=0):
        return abi.make_filter(self.included, self.min_n_messages, crop_channel_bars)


def info_from_record(r):
    """MstSongInfo record -> the reference's info dict (get_midi_info :117-179; 'scale' once get_input chose it)."""
    num, den = int(r['numerator']), 1 << int(r['denominator_log2'])
    info = {
        'ticks_per_beat': int(r['ticks_per_beat']),
        'time_signature': {'numerator': num, 'denominator': den, 'value': num / den},
        'key': None if int(r['key_signature']) < 0 else int(r['key_signature']),
        'duration': int(r['duration']),
        'ticks_per_bar': int(r['ticks_per_bar']),
        'n_bars': int(r['duration']) / int(r['ticks_per_bar']) if int(r['ticks_per_bar']) else 0.0,
        'n_beats': int(r['n_beats']),
        'tempo': int(r['tempo']),
        'bpm': int(r['bpm']),
    }
    return info


def scale_from_record(r):
    return {'key': key_names[int(r['tonic'])], 'mode': MODES[int(r['mode'])], 'coverage': float(r['scale_coverage']),
            'loose_coverage': float(r['scale_loose_coverage']), 'cross_entropy': float(r['scale_cross_entropy']),
            'ndcg': 0.0, 'loss': float(r['scale_loss'])}


def raise_for_status(status):
    status = int(status)
    if status in abi.STATUS_MESSAGE:
        raise MidiFormatError(abi.STATUS_MESSAGE[status])
    if status in (abi.SONG_NO_NOTES, abi.SONG_BAD_TIMING, abi.SONG_TOO_MANY_INSTRUMENTS):
        raise MidiFormatError(f'unusable file (status {status})')


def read_midi(mid):
    """mid: style.midi.MidiEvents (native loader) or any mido-like object with .tracks / .ticks_per_beat."""
    if hasattr(mid, 'events'):
        events, tpb = mid.events, mid.ticks_per_beat
    else:
        from mst_native import events as _events
        events, tpb = _events.from_mido(mid)
    ev, songs = _lib.make_songs([events], [tpb])
    batch = _lib.DeviceBatch(ev, songs)
    batch.encode(None, None)                       # planning pass: no grids written
    r = batch.infos()[0]
    raise_for_status(r['status'])
    return Channels(events, tpb, int(r['n_lists'])), info_from_record(r)


class ChannelConverter:
    """ChannelConverter(info, beat_divisors=(8, 3), n_octaves=8, min_percussion=35, max_percussion=81) (:349-369)."""

    def __init__(self, info, beat_divisors=(8, 3), n_octaves=8, min_percussion=35, max_percussion=81):
        if tuple(beat_divisors) != (8, 3) or n_octaves != 8 or (min_percussion, max_percussion) != (35, 81):
            raise NotImplementedError('the CUDA codec implements the reference defaults (8, 3) / 8 octaves / 35..81')
        self.info = info
        self.n_pitched_notes, self.n_unpitched_notes = 56, 47
        self.n_pitched_note_features, self.n_unpitched_note_features = 5, 2
        self.n_beat_fractions = 10

    def decode_song_record(self, n_pitched, has_unpitched, n_bars, n_beats, channel_ids, instrument_ids, write_back=True):
        """MstDecodeSong for this converter's info (tick arithmetic of vchannel2qchannel / qchannel2channel :458-562)."""
        info = self.info
        ts = info['time_signature']
        rec = dict(ticks_per_beat=info['ticks_per_beat'], ticks_per_bar=info['ticks_per_bar'], tempo=info['tempo'],
                   tonic=key_names.index(info['scale']['key']), mode=MODES.index(info['scale']['mode']),
                   numerator=ts['numerator'], denominator_log2=int(ts['denominator']).bit_length() - 1, n_beats=n_beats)
        return abi.make_decode_song(rec, n_pitched, has_unpitched, n_bars, channel_ids=channel_ids,
                                    instrument_ids=instrument_ids, duration=info.get('duration'), write_back=write_back)

    def _unsupported(self, *a, **k):
        raise NotImplementedError('per-note conversion is fused into the batched kernels: use style.data.get_input / '
                                  'style.style_transfer.decode_midi')

    channel2nchannel = nchannel2kchannel = kchannel2qchannel = qchannel2vchannel = nchannel2vchannel = _unsupported
    vchannel2channel = vchannel2qchannel = qchannel2channel = _unsupported


def get_keys_dist(info, nchannel):
    raise NotImplementedError('the key histogram is computed inside the encoder (k_scale); see MstSongInfo.keys_dist')


def get_notes_dist(info, nchannel):
    raise NotImplementedError('per-note statistics need the reference\'s Note lists, which the CUDA codec never builds')
