"""Run the UNMODIFIED reference (/root/reference) in this container.  TEST INFRASTRUCTURE ONLY.

Only oracle/gen_golden.py and the `refcheck`-marked tests (which skip when /root/reference is
absent, i.e. on the GPU box) import this.  No reference source is copied: the modules are
imported from where they lie, behind import shims for the packages this image lacks
(oracle/refshim: mido, py_utils, flatten_dict) and two API-drift patches applied from outside:

  * sklearn >= 1.4 dropped OneHotEncoder(sparse=...) (style/data.py:23,26) -> the keyword is
    mapped to sparse_output before style.data is imported;
  * pandas >= 3 hands out a read-only view from np.asarray(series), so the in-place `/=` at
    style/data.py:81-83 raises -> style.data's module-level `np` is replaced by a proxy whose
    asarray() copies.  No arithmetic changes.

Environment pin: numpy 2.x scalar promotion (np.float32 * int stays
f32), torch 2.11 CPU, pandas 3, sklearn 1.9, mido semantics >= 1.3 as restated in the shim.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get('MST_REFERENCE_ROOT', '/root/reference')
_HERE = os.path.dirname(os.path.abspath(__file__))


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, 'style'))


_loaded = None


def load():
    """Import the reference's `style` package; returns a namespace of its modules."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError('reference tree not present')
    for name in list(sys.modules):
        if name == 'style' or name.startswith('style.'):
            raise RuntimeError('a different `style` package is already imported in this process; '
                               'run the reference harness in its own process')
    sys.path.insert(0, os.path.join(_HERE, 'refshim'))
    sys.path.insert(0, REFERENCE_ROOT)

    import numpy as np
    import sklearn.preprocessing as skp

    _orig = skp.OneHotEncoder

    def _OneHotEncoder(*a, sparse=None, **kw):
        if sparse is not None:
            kw['sparse_output'] = sparse
        return _orig(*a, **kw)

    skp.OneHotEncoder = _OneHotEncoder
    try:
        import style.data as data
    finally:
        skp.OneHotEncoder = _orig

    class _NpProxy(types.ModuleType):
        def __getattr__(self, item):
            return getattr(np, item)

        @staticmethod
        def asarray(x, *a, **kw):
            return np.array(x, *a, **kw)

    data.np = _NpProxy('numpy_proxy_for_style_data')

    import style.midi as midi
    import style.midi_conversion as midi_conversion
    import style.model as model
    import style.scales as scales
    import style.style_transfer as style_transfer
    import mido

    _loaded = types.SimpleNamespace(
        data=data, midi=midi, midi_conversion=midi_conversion, model=model, scales=scales,
        style_transfer=style_transfer, mido=mido)
    return _loaded


def build_model(seed=108):
    """train-model.py:52-85 — the reference's own constructors with its hard-coded sizes."""
    import torch
    ref = load()
    m, d = ref.model, ref.data
    torch.manual_seed(seed)
    beat_size, bar_size, n_rhythm_features = 64, 128, 8
    style_size, melody_size, rhythm_size = 256, 8, 32
    # construction order of train-model.py:62-76 (it fixes the RNG stream, hence the weights)
    pce = m.PitchedChannelsEncoder(beat_size, bar_size, d.instrument_size)
    uce = m.UnpitchedChannelsEncoder(beat_size, bar_size)
    pre = m.PitchedRhythmEncoder(rhythm_size, beat_size, bar_size, d.instrument_size)
    ure = m.UnpitchedRhythmEncoder(rhythm_size, beat_size, bar_size)
    se = m.StyleEncoder(style_size, bar_size, d.instrument_size)
    me = m.MelodyEncoder(melody_size, beat_size, bar_size, d.instrument_size)
    sim = m.SongInfoModel(n_rhythm_features, style_size, rhythm_size, d.n_instruments)
    psa = m.PitchedStyleApplier(style_size, melody_size, rhythm_size, d.instrument_size)
    usa = m.UnpitchedStyleApplier(style_size, rhythm_size)
    return m.StyleTransferModel(pce, uce, se, me, pre, ure, sim, psa, usa)
