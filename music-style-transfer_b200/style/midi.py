"""style/midi.py of the reference: General MIDI tables and predicates (midi.py:17-101), file loading (:104-108) and
the MIDI-file object decode produces (create_midi :120-168 + mid.save).  Parsing and serialisation are the native
SMF reader / writer of libmst_b200.so (mst_smf_parse / mst_smf_write); nothing here walks messages in Python."""
import math

from mst_native import lib as _lib

default_tempo = 500000
default_volume = 96
max_volume = 127
max_velocity = 127

# the 40 most used programs of the Lakh data set (midi.py:23-64)
popular_instruments = [0, 25, 48, 33, 1, 27, 49, 29, 35, 30, 50, 24, 5, 4, 32, 52, 26, 18, 28, 89, 65, 53, 61, 2, 17,
                       73, 54, 62, 16, 39, 34, 51, 90, 56, 66, 38, 11, 81, 3, 57]

_GM_FAMILIES = ['Piano', 'Chromatic Percussion', 'Organ', 'Guitar', 'Bass', 'Strings', 'Ensemble', 'Brass', 'Reed', 'Pipe',
                'Synth Lead', 'Synth Pad', 'Synth Effects', 'Ethnic', 'Percussive', 'Sound Effects']
_GM_NAMES = [
    'Acoustic Grand Piano', 'Bright Acoustic Piano', 'Electric Grand Piano', 'Honky-tonk Piano', 'Electric Piano 1',
    'Electric Piano 2', 'Harpsichord', 'Clavinet', 'Celesta', 'Glockenspiel', 'Music Box', 'Vibraphone', 'Marimba',
    'Xylophone', 'Tubular Bells', 'Dulcimer', 'Drawbar Organ', 'Percussive Organ', 'Rock Organ', 'Church Organ',
    'Reed Organ', 'Accordion', 'Harmonica', 'Tango Accordion', 'Acoustic Guitar (nylon)', 'Acoustic Guitar (steel)',
    'Electric Guitar (jazz)', 'Electric Guitar (clean)', 'Electric Guitar (muted)', 'Overdriven Guitar',
    'Distortion Guitar', 'Guitar Harmonics', 'Acoustic Bass', 'Electric Bass (finger)', 'Electric Bass (pick)',
    'Fretless Bass', 'Slap Bass 1', 'Slap Bass 2', 'Synth Bass 1', 'Synth Bass 2', 'Violin', 'Viola', 'Cello',
    'Contrabass', 'Tremolo Strings', 'Pizzicato Strings', 'Orchestral Harp', 'Timpani', 'String Ensemble 1',
    'String Ensemble 2', 'Synth Strings 1', 'Synth Strings 2', 'Choir Aahs', 'Voice Oohs', 'Synth Choir', 'Orchestra Hit',
    'Trumpet', 'Trombone', 'Tuba', 'Muted Trumpet', 'French Horn', 'Brass Section', 'Synth Brass 1', 'Synth Brass 2',
    'Soprano Sax', 'Alto Sax', 'Tenor Sax', 'Baritone Sax', 'Oboe', 'English Horn', 'Bassoon', 'Clarinet', 'Piccolo',
    'Flute', 'Recorder', 'Pan Flute', 'Blown Bottle', 'Shakuhachi', 'Whistle', 'Ocarina', 'Lead 1 (square)',
    'Lead 2 (sawtooth)', 'Lead 3 (calliope)', 'Lead 4 (chiff)', 'Lead 5 (charang)', 'Lead 6 (voice)', 'Lead 7 (fifths)',
    'Lead 8 (bass + lead)', 'Pad 1 (new age)', 'Pad 2 (warm)', 'Pad 3 (polysynth)', 'Pad 4 (choir)', 'Pad 5 (bowed)',
    'Pad 6 (metallic)', 'Pad 7 (halo)', 'Pad 8 (sweep)', 'FX 1 (rain)', 'FX 2 (soundtrack)', 'FX 3 (crystal)',
    'FX 4 (atmosphere)', 'FX 5 (brightness)', 'FX 6 (goblins)', 'FX 7 (echoes)', 'FX 8 (sci-fi)', 'Sitar', 'Banjo',
    'Shamisen', 'Koto', 'Kalimba', 'Bagpipe', 'Fiddle', 'Shanai', 'Tinkle Bell', 'Agogo', 'Steel Drums', 'Woodblock',
    'Taiko Drum', 'Melodic Tom', 'Synth Drum', 'Reverse Cymbal', 'Guitar Fret Noise', 'Breath Noise', 'Seashore',
    'Bird Tweet', 'Telephone Ring', 'Helicopter', 'Applause', 'Gunshot']
assert len(_GM_NAMES) == 128

program2instrument = dict(enumerate(_GM_NAMES))
program2group = {p: _GM_FAMILIES[p // 8] for p in range(128)}
program2instrument[-1] = 'Percussion'


def get_instrument_id(program, channel=0):
    return -1 if channel == 9 else program


def is_sound_effect(instrument_id):
    return instrument_id > 119


def is_pitched(instrument_id):
    return instrument_id >= 0 and not is_sound_effect(instrument_id)


class MidiEvents:
    """What load_midi_from_file returns instead of a mido.MidiFile: the file's events as MstEvent records
    (include/mst_b200.h) plus ticks_per_beat — the SoA buffer the encode kernels consume."""

    def __init__(self, events, ticks_per_beat, path=None):
        self.events, self.ticks_per_beat, self.path = events, int(ticks_per_beat), path


def load_midi_from_file(path):
    """mido.MidiFile(path) or None on the load failures the reference swallows (midi.py:104-108)."""
    try:
        with open(path, 'rb') as f:
            data = f.read()
    except OSError:
        return None
    parsed = _lib.smf_parse(data)
    if parsed is None:
        return None
    return MidiEvents(parsed[0], parsed[1], path)


class DecodedMidi:
    """Result of decode_midi / create_midi: the ordered note messages (MstOutMsg records) and the file bytes the
    native writer produced from them; `.save(path)` like mido.MidiFile."""

    def __init__(self, messages, data, ticks_per_beat):
        self.messages, self.data, self.ticks_per_beat = messages, data, ticks_per_beat

    def save(self, path):
        with open(path, 'wb') as f:
            f.write(self.data)


def create_midi(info, *instruments, max_delta_time=math.inf):
    """create_midi (midi.py:120-168) on Python message lists: `instruments` are channel dicts {'channel_id', 'instrument_id',
    'messages': [objects with .type ('note_on' | 'note_off'), .note, .time (absolute ticks), .velocity]}.  Host logic of the
    reference restated on numpy (int velocities, stable sort by time, end_of_track, delta times with the max_delta_time clip);
    the bytes come from the native SMF writer.  Returns a DecodedMidi (`.save(path)` like mido.MidiFile).
    The fused GPU path (style_transfer.decode_midi) does the same on the device for whole grids."""
    import numpy as np
    from mst_native import abi
    tpb, tempo = int(info['ticks_per_beat']), int(info['tempo'])
    clip = None
    if math.isfinite(max_delta_time):
        clip = int(round(max_delta_time / (tempo * 1e-6 / tpb)))          # mido.second2tick, then int() (:121-123)
    if len(instruments) > abi.MAX_DECODE_CHANNELS:
        raise ValueError(f'at most {abi.MAX_DECODE_CHANNELS} channels per file')
    rows = []
    for idx, instrument in enumerate(instruments):
        ch = int(instrument['channel_id'])
        if not 0 <= ch <= 15:
            raise ValueError('channel must be in range 0..15')
        for msg in instrument['messages']:
            velocity = int(msg.velocity * max_velocity)                    # np.float32 * int stays f32 under numpy >= 2 (:147)
            assert velocity <= 127, (velocity, msg.velocity)
            if not (0 <= int(msg.note) <= 127 and velocity >= 0):
                raise ValueError('data byte must be in range 0..127')     # what mido.Message raises
            rows.append((int(msg.time), 1 if msg.type == 'note_on' else 0, ch, int(msg.note), velocity))
    rows.sort(key=lambda r: r[0])                                          # sorted(messages, key=time): stable (:157)
    duration = info.get('duration', (rows[-1][0] + info['ticks_per_bar']) if rows else None)
    if duration is None:
        raise IndexError('list index out of range')                       # messages[-1] on an empty list (:158)
    out = np.zeros(len(rows), dtype=abi.OUT_MSG_DTYPE)
    time = 0
    for i, (t, on, ch, note, vel) in enumerate(rows):
        delta = t - time if clip is None else min(t - time, clip)
        time = t
        out[i] = (t, max(0, delta), on, ch, note, vel if on else vel)
    eot = int(duration) - time if clip is None else min(int(duration) - time, clip)
    ts = info['time_signature']
    rec = dict(ticks_per_beat=tpb, ticks_per_bar=int(info.get('ticks_per_bar', tpb * ts['numerator'])), tempo=tempo, tonic=0, mode=0,
               numerator=int(ts['numerator']), denominator_log2=int(ts['denominator']).bit_length() - 1, n_beats=int(ts['numerator']))
    song = abi.make_decode_song(rec, len(instruments), False, 1, channel_ids=[int(i['channel_id']) for i in instruments],
                                instrument_ids=[max(0, int(i.get('instrument_id', 0))) for i in instruments], write_back=False)
    data = _lib.smf_write(song, out, max(0, eot))
    return DecodedMidi(out, data, tpb)
