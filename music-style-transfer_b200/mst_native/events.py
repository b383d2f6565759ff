"""mido-like message objects -> MstEvent records (include/mst_b200.h).

`read_midi(mid)` (style/midi_conversion.py:216) receives a mido.MidiFile.  The drop-in accepts
any object with `.tracks` (iterables of messages with mido's attribute names) and
`.ticks_per_beat`; this module flattens it into the SoA event buffer the kernels consume.
Files loaded through the native SMF parser never take this (slow, per-message Python) route.
"""
import numpy as np

from . import abi

_CHANNEL_KINDS = {
    'note_off': abi.EV_NOTE_OFF, 'note_on': abi.EV_NOTE_ON, 'polytouch': abi.EV_POLYTOUCH,
    'control_change': abi.EV_CONTROL_CHANGE, 'program_change': abi.EV_PROGRAM_CHANGE,
    'aftertouch': abi.EV_AFTERTOUCH, 'pitchwheel': abi.EV_PITCHWHEEL,
}
# messages_to_ignore without a channel attribute (style/midi_conversion.py:80-98)
_IGNORED = {
    'smpte_offset', 'midi_port', 'sysex', 'end_of_track', 'track_name', 'copyright', 'lyrics',
    'marker', 'sequencer_specific', 'text', 'instrument_name', 'cue_marker', 'unknown_meta',
    'sequence_number',
}
_KEY_IDS = {}
for _i, _name in enumerate(['Cb', 'Gb', 'Db', 'Ab', 'Eb', 'Bb', 'F', 'C', 'G', 'D', 'A', 'E', 'B',
                            'F#', 'C#']):
    _KEY_IDS[_name] = _i * 2
for _i, _name in enumerate(['Abm', 'Ebm', 'Bbm', 'Fm', 'Cm', 'Gm', 'Dm', 'Am', 'Em', 'Bm', 'F#m',
                            'C#m', 'G#m', 'D#m', 'A#m']):
    _KEY_IDS[_name] = _i * 2 + 1


def from_mido(mid):
    """Return (events, ticks_per_beat): events in track-major order with absolute ticks."""
    records = []
    for t, track in enumerate(mid.tracks):
        time = 0
        for msg in track:
            time += msg.time
            if time > abi.MST_MAX_TICK:
                # split_channels drops these (style/midi_conversion.py:52,59); ticks are
                # non-decreasing inside a track, so the rest of the track is dropped too
                break
            typ = msg.type
            kind = _CHANNEL_KINDS.get(typ)
            aux = ch = d1 = d2 = 0
            if kind is not None:
                ch = msg.channel
                if kind in (abi.EV_NOTE_ON, abi.EV_NOTE_OFF):
                    d1, d2 = msg.note, msg.velocity
                elif kind == abi.EV_CONTROL_CHANGE:
                    d1, d2 = msg.control, msg.value
                elif kind == abi.EV_PROGRAM_CHANGE:
                    d1 = msg.program
            elif typ == 'set_tempo':
                kind, aux = abi.EV_SET_TEMPO, msg.tempo
            elif typ == 'time_signature':
                kind, d1 = abi.EV_TIME_SIGNATURE, msg.numerator
                d2 = int(msg.denominator).bit_length() - 1
            elif typ == 'key_signature':
                kind, aux = abi.EV_KEY_SIGNATURE, _KEY_IDS[msg.key]
            elif typ == 'channel_prefix':
                kind, ch = abi.EV_CHANNEL_PREFIX, msg.channel
            elif typ in _IGNORED:
                continue          # carries no information for the encoder
            else:
                kind = abi.EV_UNKNOWN
            records.append((time, aux, kind, ch, d1, d2, min(t, 65535), 0))
    events = np.array(records, dtype=abi.EVENT_DTYPE) if records else np.zeros(0, abi.EVENT_DTYPE)
    return events, int(mid.ticks_per_beat)
