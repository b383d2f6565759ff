"""Stub for the un-listed `py_utils` dependency (style/scales.py:203).  Test infrastructure only."""
