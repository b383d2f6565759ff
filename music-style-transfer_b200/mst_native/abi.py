"""numpy mirrors of the structs in include/mst_b200.h (kept in lock-step; tests check sizeof)."""
import numpy as np

MST_ABI_VERSION = 2
PRECISION_FP32, PRECISION_BF16 = 0, 1
MST_MAX_TICK = 10_000_000

# event kinds
EV_NOTE_OFF, EV_NOTE_ON, EV_POLYTOUCH, EV_CONTROL_CHANGE = 0, 1, 2, 3
EV_PROGRAM_CHANGE, EV_AFTERTOUCH, EV_PITCHWHEEL = 4, 5, 6
EV_SET_TEMPO, EV_TIME_SIGNATURE, EV_KEY_SIGNATURE = 16, 17, 18
EV_CHANNEL_PREFIX, EV_META_IGNORED, EV_UNKNOWN = 19, 20, 21

# per-song status
SONG_OK = 0
SONG_TIME_SIG_CHANGED = 1
SONG_KEY_SIG_CHANGED = 2
SONG_UNKNOWN_MESSAGE = 3
SONG_NO_NOTES = 4
SONG_NO_PITCHED = 5
SONG_BAD_TIMING = 6
SONG_TOO_MANY_INSTRUMENTS = 7
SONG_ARENA_FULL = 8
SONG_LOAD_FAILED = 9

# messages of the reference's MidiFormatError for each reject status
# (style/midi_conversion.py:153,157,165)
STATUS_MESSAGE = {
    SONG_TIME_SIG_CHANGED: 'Time signature changed',
    SONG_KEY_SIG_CHANGED: 'Key signature changed',
    SONG_UNKNOWN_MESSAGE: 'Unknown message type',
}

MAX_PITCHED = 120
MAX_UNPITCHED = 12
MAX_LISTS_PER_CHANNEL = 16
N_FRACTIONS = 10
N_PITCHED_ROWS, N_PITCHED_FEATS = 56, 5
N_UNPITCHED_ROWS, N_UNPITCHED_FEATS = 47, 2
PITCHED_BEAT_FLOATS = N_FRACTIONS * N_PITCHED_ROWS * N_PITCHED_FEATS
UNPITCHED_BEAT_FLOATS = N_FRACTIONS * N_UNPITCHED_ROWS * N_UNPITCHED_FEATS

EVENT_DTYPE = np.dtype([
    ('tick', '<u4'), ('aux', '<u4'), ('kind', 'u1'), ('channel', 'u1'), ('d1', 'u1'), ('d2', 'u1'),
    ('track', '<u2'), ('reserved', '<u2'),
])
assert EVENT_DTYPE.itemsize == 16

SONG_IN_DTYPE = np.dtype([('event_offset', '<u8'), ('n_events', '<u4'), ('ticks_per_beat', '<i4')])
FILE_IN_DTYPE = np.dtype([('byte_offset', '<u8'), ('n_bytes', '<u4'), ('reserved', '<u4')])      # MstFileIn
assert SONG_IN_DTYPE.itemsize == 16

FILTER_DTYPE = np.dtype([
    ('apply', '<i4'), ('min_n_messages', '<i4'), ('included', 'u1', (128,)),
    ('crop_channel_bars', '<i4'), ('reserved', '<i4'),
])
assert FILTER_DTYPE.itemsize == 144

SONG_INFO_DTYPE = np.dtype([
    ('status', '<i4'), ('ticks_per_beat', '<i4'), ('numerator', '<i4'), ('denominator_log2', '<i4'),
    ('key_signature', '<i4'), ('duration', '<i4'), ('ticks_per_bar', '<i4'), ('n_bars', '<i4'),
    ('n_beats', '<i4'), ('tempo', '<i4'), ('bpm', '<i4'), ('tonic', '<i4'), ('mode', '<i4'),
    ('n_pitched', '<i4'), ('n_unpitched', '<i4'), ('n_bars_out', '<i4'),
    ('scale_coverage', '<f8'), ('scale_loose_coverage', '<f8'), ('scale_cross_entropy', '<f8'),
    ('scale_loss', '<f8'), ('scale_margin', '<f8'), ('keys_dist', '<f8', (12,)),
    ('pitched_offset', '<u8'), ('unpitched_offset', '<u8'),
    ('n_notes', '<i4'), ('n_lists', '<i4'),
    ('pitched_instruments', '<i2', (MAX_PITCHED,)), ('unpitched_instruments', '<i2', (MAX_UNPITCHED,)),
])
assert SONG_INFO_DTYPE.itemsize == 488


MAX_DECODE_CHANNELS = 32
DECODE_OK, DECODE_NO_MESSAGES, DECODE_BAD_VALUE, DECODE_VELOCITY, DECODE_CAPACITY = 0, 1, 2, 3, 4

DECODE_SONG_DTYPE = np.dtype([
    ('pitched_offset', '<u8'), ('unpitched_offset', '<u8'),
    ('n_pitched', '<i4'), ('has_unpitched', '<i4'), ('n_bars', '<i4'), ('n_beats', '<i4'),
    ('ticks_per_beat', '<i4'), ('ticks_per_bar', '<i4'), ('tempo', '<i4'), ('tonic', '<i4'), ('mode', '<i4'),
    ('duration', '<i4'), ('numerator', '<i4'), ('denominator_log2', '<i4'),
    ('channel_id', 'u1', (MAX_DECODE_CHANNELS,)), ('instrument_id', '<i2', (MAX_DECODE_CHANNELS,)),
    ('unpitched_channel_id', '<i4'), ('write_back', '<i4'),
])
assert DECODE_SONG_DTYPE.itemsize == 168

OUT_MSG_DTYPE = np.dtype([('time', '<i4'), ('delta', '<i4'), ('is_on', 'u1'), ('channel', 'u1'),
                          ('note', 'u1'), ('velocity', 'u1')])
assert OUT_MSG_DTYPE.itemsize == 12

DECODE_RESULT_DTYPE = np.dtype([('status', '<i4'), ('n_messages', '<u4'), ('msg_offset', '<u8'),
                                ('eot_delta', '<i4'), ('max_delta', '<i4')])
assert DECODE_RESULT_DTYPE.itemsize == 24


def make_decode_song(info, n_pitched, has_unpitched, n_bars, channel_ids=None, instrument_ids=None,
                     duration=None, write_back=True, pitched_offset=0, unpitched_offset=0):
    """MstDecodeSong from an info record/dict with ticks_per_beat, ticks_per_bar, tempo, tonic, mode,
    numerator, denominator_log2 (+ n_beats).  channel ids default to 0..15 without 9
    (style/style_transfer.py:78-80)."""
    d = np.zeros((), dtype=DECODE_SONG_DTYPE)
    d['pitched_offset'], d['unpitched_offset'] = pitched_offset, unpitched_offset
    d['n_pitched'], d['has_unpitched'], d['n_bars'] = n_pitched, int(bool(has_unpitched)), n_bars
    for k in ('n_beats', 'ticks_per_beat', 'ticks_per_bar', 'tempo', 'tonic', 'mode', 'numerator',
              'denominator_log2'):
        d[k] = int(info[k])
    d['duration'] = -1 if duration is None else int(duration)
    if channel_ids is None:
        channel_ids = [i for i in range(16) if i != 9][:n_pitched]
    ch = np.zeros(MAX_DECODE_CHANNELS, np.uint8)
    ch[:len(channel_ids)] = channel_ids
    d['channel_id'] = ch
    ins = np.zeros(MAX_DECODE_CHANNELS, np.int16)
    if instrument_ids is not None:
        ins[:len(instrument_ids)] = instrument_ids
    d['instrument_id'] = ins
    d['unpitched_channel_id'] = 9
    d['write_back'] = int(bool(write_back))
    return d


def make_filter(included=None, min_n_messages=100, crop_channel_bars=0):
    """MstFilter for iter_inputs (style/data.py:53-56); included=None disables the filter."""
    f = np.zeros((), dtype=FILTER_DTYPE)
    if included is not None:
        f['apply'] = 1
        f['min_n_messages'] = min_n_messages
        inc = np.zeros(128, dtype=np.uint8)
        inc[list(included)] = 1
        f['included'] = inc
    f['crop_channel_bars'] = crop_channel_bars
    return f


TRANSFER_CHOICE_DTYPE = np.dtype([('n_pitched', '<i4'), ('has_unpitched', '<i4'), ('mode', '<i4'), ('tempo', '<i4'),
                                  ('instruments', '<i2', (MAX_DECODE_CHANNELS,))])
assert TRANSFER_CHOICE_DTYPE.itemsize == 80
