"""Minimal stand-in for the third-party `mido` package (absent from this image, un-pinned in
the reference's requirements.txt:2).  TEST INFRASTRUCTURE ONLY: it exists so that the
*unmodified* reference under /root/reference can be imported in this container to generate
golden vectors (oracle/gen_golden.py).  Nothing in the product path imports it.

It restates the published behaviour of mido >= 1.3 for exactly the call sites the reference
uses (style/midi.py:6-7,106,121,125-167; style/midi_conversion.py:8,177,333,343,394;
style/style_transfer.py:4,98,103):

  MidiFile(path) / MidiFile() / .tracks / .ticks_per_beat / .save(path)
  MidiTrack, Message, MetaMessage, KeySignatureError
  tick2second, second2tick (int(round()) as in mido >= 1.2.10), tempo2bpm, bpm2tempo

Parity note ("parity unpinned"): the reference's own tests hold no vectors for these calls,
so the SMF parse/serialise semantics are pinned to this restatement and recorded in DESIGN.md.
"""
import struct
from numbers import Integral

DEFAULT_TEMPO = 500000
DEFAULT_TICKS_PER_BEAT = 480


class KeySignatureError(Exception):
    pass


# (status high nibble) -> (type name, data field names)
_CHANNEL_SPECS = {
    0x80: ('note_off', ('note', 'velocity')),
    0x90: ('note_on', ('note', 'velocity')),
    0xA0: ('polytouch', ('note', 'value')),
    0xB0: ('control_change', ('control', 'value')),
    0xC0: ('program_change', ('program',)),
    0xD0: ('aftertouch', ('value',)),
    0xE0: ('pitchwheel', ('pitch',)),
}
_CHANNEL_STATUS = {v[0]: k for k, v in _CHANNEL_SPECS.items()}

# system common / realtime messages that mido can decode (no channel attribute)
_SYSTEM_SPECS = {
    0xF1: ('quarter_frame', 1),
    0xF2: ('songpos', 2),
    0xF3: ('song_select', 1),
    0xF6: ('tune_request', 0),
    0xF8: ('clock', 0),
    0xFA: ('start', 0),
    0xFB: ('continue', 0),
    0xFC: ('stop', 0),
    0xFE: ('active_sensing', 0),
}

_META_NAMES = {
    0x00: 'sequence_number',
    0x01: 'text',
    0x02: 'copyright',
    0x03: 'track_name',
    0x04: 'instrument_name',
    0x05: 'lyrics',
    0x06: 'marker',
    0x07: 'cue_marker',
    0x09: 'device_name',
    0x20: 'channel_prefix',
    0x21: 'midi_port',
    0x2F: 'end_of_track',
    0x51: 'set_tempo',
    0x54: 'smpte_offset',
    0x58: 'time_signature',
    0x59: 'key_signature',
    0x7F: 'sequencer_specific',
}

_KEYS = {
    (-7, 0): 'Cb', (-6, 0): 'Gb', (-5, 0): 'Db', (-4, 0): 'Ab', (-3, 0): 'Eb', (-2, 0): 'Bb',
    (-1, 0): 'F', (0, 0): 'C', (1, 0): 'G', (2, 0): 'D', (3, 0): 'A', (4, 0): 'E', (5, 0): 'B',
    (6, 0): 'F#', (7, 0): 'C#',
    (-7, 1): 'Abm', (-6, 1): 'Ebm', (-5, 1): 'Bbm', (-4, 1): 'Fm', (-3, 1): 'Cm', (-2, 1): 'Gm',
    (-1, 1): 'Dm', (0, 1): 'Am', (1, 1): 'Em', (2, 1): 'Bm', (3, 1): 'F#m', (4, 1): 'C#m',
    (5, 1): 'G#m', (6, 1): 'D#m', (7, 1): 'A#m',
}
_KEYS_INV = {v: k for k, v in _KEYS.items()}


class _Base:
    is_meta = False

    def copy(self, **overrides):
        new = self.__class__.__new__(self.__class__)
        new.__dict__.update(self.__dict__)
        new.__dict__.update(overrides)
        return new

    def __copy__(self):
        return self.copy()

    def __repr__(self):
        kv = ' '.join(f'{k}={v!r}' for k, v in self.__dict__.items() if k != 'type')
        return f'<{self.__class__.__name__} {self.type} {kv}>'

    def __eq__(self, other):
        return type(self) is type(other) and self.__dict__ == other.__dict__


class Message(_Base):
    def __init__(self, type, **kw):
        self.type = type
        if type in _CHANNEL_STATUS:
            self.channel = 0
            for name in _CHANNEL_SPECS[_CHANNEL_STATUS[type]][1]:
                setattr(self, name, 64 if name == 'velocity' else 0)
        elif type == 'sysex':
            self.data = ()
        self.time = 0
        for k, v in kw.items():
            if k in ('note', 'velocity', 'program', 'control', 'value') and type != 'pitchwheel':
                if not (isinstance(v, Integral) and 0 <= v <= 127):
                    raise ValueError(f'{k} must be int in range 0..127')
            if k == 'channel' and not (isinstance(v, Integral) and 0 <= v <= 15):
                raise ValueError('channel must be int in range 0..15')
            setattr(self, k, v)

    def bytes(self):
        if self.type in _CHANNEL_STATUS:
            status = _CHANNEL_STATUS[self.type] | self.channel
            names = _CHANNEL_SPECS[_CHANNEL_STATUS[self.type]][1]
            if self.type == 'pitchwheel':
                v = self.pitch + 8192
                return [status, v & 0x7F, v >> 7]
            return [status] + [getattr(self, n) for n in names]
        raise ValueError(f'cannot serialise {self.type}')


class MetaMessage(_Base):
    is_meta = True

    def __init__(self, type, **kw):
        self.type = type
        if type == 'time_signature':
            self.numerator, self.denominator = 4, 4
            self.clocks_per_click, self.notated_32nd_notes_per_beat = 24, 8
        elif type == 'set_tempo':
            self.tempo = DEFAULT_TEMPO
        elif type == 'key_signature':
            self.key = 'C'
        elif type == 'channel_prefix':
            self.channel = 0
        self.time = 0
        for k, v in kw.items():
            setattr(self, k, v)

    def bytes(self):
        if self.type == 'time_signature':
            dd = self.denominator.bit_length() - 1
            data = [self.numerator, dd, self.clocks_per_click, self.notated_32nd_notes_per_beat]
            t = 0x58
        elif self.type == 'set_tempo':
            data = [self.tempo >> 16, (self.tempo >> 8) & 0xFF, self.tempo & 0xFF]
            t = 0x51
        elif self.type == 'end_of_track':
            data, t = [], 0x2F
        elif self.type == 'key_signature':
            sf, mi = _KEYS_INV[self.key]
            data, t = [sf & 0xFF, mi], 0x59
        else:
            raise ValueError(f'cannot serialise meta {self.type}')
        return [0xFF, t] + _encode_varint(len(data)) + data


class UnknownMetaMessage(MetaMessage):
    def __init__(self, type_byte, data=None, time=0):
        self.type = 'unknown_meta'
        self.type_byte = type_byte
        self.data = tuple(data or ())
        self.time = time


class MidiTrack(list):
    pass


def _encode_varint(value):
    out = [value & 0x7F]
    value >>= 7
    while value:
        out.append((value & 0x7F) | 0x80)
        value >>= 7
    return out[::-1]


class _Reader:
    def __init__(self, data):
        self.data = data
        self.pos = 0

    def byte(self):
        if self.pos >= len(self.data):
            raise EOFError
        b = self.data[self.pos]
        self.pos += 1
        return b

    def read(self, n):
        if self.pos + n > len(self.data):
            raise EOFError
        b = self.data[self.pos:self.pos + n]
        self.pos += n
        return b

    def varint(self):
        delta = 0
        while True:
            b = self.byte()
            delta = (delta << 7) | (b & 0x7F)
            if b < 0x80:
                return delta


def _decode_meta(type_byte, data, delta):
    name = _META_NAMES.get(type_byte)
    if name is None:
        return UnknownMetaMessage(type_byte, data, time=delta)
    msg = MetaMessage(name)
    msg.time = delta
    if name == 'set_tempo':
        msg.tempo = (data[0] << 16) | (data[1] << 8) | data[2]
    elif name == 'time_signature':
        msg.numerator = data[0]
        msg.denominator = 2 ** data[1]
        msg.clocks_per_click = data[2]
        msg.notated_32nd_notes_per_beat = data[3]
    elif name == 'key_signature':
        sf = data[0] - 256 if data[0] > 127 else data[0]
        try:
            msg.key = _KEYS[(sf, data[1])]
        except KeyError:
            raise KeySignatureError(
                f'Could not decode key with {sf} flats/sharps and mode {data[1]}')
    elif name == 'channel_prefix':
        msg.channel = data[0]
    elif name == 'midi_port':
        msg.port = data[0] if data else 0
    elif name == 'sequence_number':
        msg.number = ((data[0] << 8) | data[1]) if len(data) >= 2 else 0
    elif name in ('text', 'copyright', 'track_name', 'instrument_name', 'lyrics', 'marker',
                  'cue_marker', 'device_name'):
        setattr(msg, 'name' if name.endswith('_name') else 'text', bytes(data).decode('latin1'))
    elif name in ('sequencer_specific', 'smpte_offset'):
        msg.data = tuple(data)
    return msg


def _read_track(r):
    name = r.read(4)
    size = struct.unpack('>L', r.read(4))[0]
    if name != b'MTrk':
        raise IOError('no MTrk header at start of track')
    track = MidiTrack()
    start = r.pos
    last_status = None
    while True:
        if r.pos - start == size:
            break
        delta = r.varint()
        status = r.byte()
        if status < 0x80:
            if last_status is None:
                raise IOError('running status without last_status')
            peek = [status]
            status = last_status
        else:
            if status != 0xFF:
                last_status = status
            peek = []
        if status == 0xFF:
            type_byte = r.byte()
            length = r.varint()
            data = list(r.read(length))
            track.append(_decode_meta(type_byte, data, delta))
        elif status in (0xF0, 0xF7):
            length = r.varint()
            data = r.read(length)
            msg = Message('sysex')
            msg.data = tuple(data)
            msg.time = delta
            track.append(msg)
        elif status < 0xF0:
            tname, fields = _CHANNEL_SPECS[status & 0xF0]
            need = len(fields) if tname != 'pitchwheel' else 2
            data = peek + list(r.read(need - len(peek)))
            for b in data:
                if b > 127:
                    raise IOError('data byte must be in range 0..127')
            msg = Message.__new__(Message)
            msg.type = tname
            msg.channel = status & 0x0F
            if tname == 'pitchwheel':
                msg.pitch = (data[0] | (data[1] << 7)) - 8192
            else:
                for f, b in zip(fields, data):
                    setattr(msg, f, b)
            msg.time = delta
            track.append(msg)
        else:
            spec = _SYSTEM_SPECS.get(status)
            if spec is None:
                raise IOError(f'undefined status byte 0x{status:02x}')
            data = peek + list(r.read(spec[1] - len(peek)))
            msg = Message.__new__(Message)
            msg.type = spec[0]
            msg.data = tuple(data)
            msg.time = delta
            track.append(msg)
    return track


class MidiFile:
    def __init__(self, filename=None, type=1, ticks_per_beat=DEFAULT_TICKS_PER_BEAT):
        self.filename = filename
        self.type = type
        self.ticks_per_beat = ticks_per_beat
        self.tracks = []
        if filename is not None:
            with open(filename, 'rb') as f:
                self._load(f.read())

    def _load(self, data):
        r = _Reader(data)
        try:
            header = r.read(8)
        except EOFError:
            raise EOFError
        if header[:4] != b'MThd':
            raise IOError('MThd not found. Probably not a MIDI file')
        size = struct.unpack('>L', header[4:])[0]
        hdr = r.data[r.pos:r.pos + size]
        r.pos += size
        if len(hdr) < 6:
            raise EOFError
        self.type, num_tracks, self.ticks_per_beat = struct.unpack('>hhh', hdr[:6])
        for _ in range(num_tracks):
            self.tracks.append(_read_track(r))

    def save(self, filename):
        out = bytearray()
        out += b'MThd' + struct.pack('>LHHH', 6, self.type, len(self.tracks), self.ticks_per_beat)
        for track in self.tracks:
            data = bytearray()
            running = None
            for msg in track:
                if not isinstance(msg.time, Integral):
                    raise ValueError('message time must be int in MIDI file')
                if msg.time < 0:
                    raise ValueError('message time must be non-negative in MIDI file')
                data += bytes(_encode_varint(msg.time))
                if msg.is_meta:
                    data += bytes(msg.bytes())
                    running = None
                else:
                    b = msg.bytes()
                    if b[0] == running:
                        data += bytes(b[1:])
                    else:
                        data += bytes(b)
                    running = b[0] if b[0] < 0xF0 else None
            out += b'MTrk' + struct.pack('>L', len(data)) + data
        with open(filename, 'wb') as f:
            f.write(out)


def tick2second(tick, ticks_per_beat, tempo):
    scale = tempo * 1e-6 / ticks_per_beat
    return tick * scale


def second2tick(second, ticks_per_beat, tempo):
    scale = tempo * 1e-6 / ticks_per_beat
    return int(round(second / scale))


def tempo2bpm(tempo, time_signature=(4, 4)):
    return 60 * 1e6 / tempo * time_signature[1] / 4.


def bpm2tempo(bpm, time_signature=(4, 4)):
    return int(round(60 * 1e6 / bpm * 4. / time_signature[1]))


def open_output(*a, **k):
    raise RuntimeError('no MIDI ports in the test shim')
