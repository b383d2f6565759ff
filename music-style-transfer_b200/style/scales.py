"""Scale vocabulary of the reference (style/scales.py): key names and the two Mode objects the pipeline compares
against (`info['scale']['mode'] == major_mode`, data.py:145; style_transfer.py:118-122).  Key finding itself
(get_scales / get_key_scores, scales.py:160-221) runs on the GPU inside the encoder (k_scale)."""

key_names = ['C', 'C#', 'D', 'D#', 'E', 'F', 'F#', 'G', 'G#', 'A', 'A#', 'B']
interval2key = dict(enumerate(key_names))
key2interval = {key: interval for interval, key in interval2key.items()}


class Mode:
    """Diatonic mode: interval pattern + rotation (scales.py:27-92, the part the hot path uses)."""
    names = ['Ionian', 'Dorian', 'Phrygian', 'Lydian', 'Mixolydian', 'Aeolian', 'Locrian']

    def __init__(self, intervals, shift=0):
        self.intervals = list(intervals)
        self.shift = shift % 7
        self.absolute_intervals = [0]
        for interval in self.intervals[:-1]:
            self.absolute_intervals.append(self.absolute_intervals[-1] + interval)

    @property
    def name(self):
        return self.names[self.shift]

    def __eq__(self, other):
        return isinstance(other, Mode) and self.intervals == other.intervals

    def __hash__(self):
        return hash(tuple(self.intervals))

    def __repr__(self):
        return f'Mode({self.name})'


def create_mode(mode, shift):
    k = shift % 7
    return Mode(mode.intervals[k:] + mode.intervals[:k], mode.shift + shift)


major_mode = Mode([2, 2, 1, 2, 2, 2, 1])
minor_mode = create_mode(major_mode, shift=-2)      # Aeolian [2, 1, 2, 2, 1, 2, 2]
MODES = (major_mode, minor_mode)                    # index = MstSongInfo.mode
