"""style/exceptions.py of the reference (exceptions.py:1-4)."""


class MidiFormatError(Exception):
    def __init__(self, msg=''):
        super().__init__(msg)
        self.msg = msg
