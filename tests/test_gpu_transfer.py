"""mst_transfer_select (apply_style's host decisions on the device, style_transfer.py:101-122) against the reference's own
Python statements restated here on numpy: np.argsort(-logits)[:n], the lone-percussion special case, percussion removal,
the one-hot inverse transform, mode arg-max, bpm2tempo(round(bpm)), encode_instruments."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def torch():
    import torch
    if not torch.cuda.is_available():
        pytest.fail('GPU tests need a CUDA device (no CPU fallback exists)')
    return torch


def reference_choice(logits, mode_logits, bpm, n_instruments):
    """style_transfer.py:103-122 on host arrays (the statements of the reference, numpy only)."""
    from style import data as d
    tempo = int(round(60 * 1e6 / round(float(bpm))))
    instruments = np.argsort(-logits)[:n_instruments]
    if len(instruments) == 1 and instruments[0] == [d.percussion_id]:
        instruments = np.argsort(-logits)[:n_instruments + 1]
    unpitched = d.percussion_id in instruments
    instruments = [i for i in instruments if i != d.percussion_id]
    enc = np.zeros([len(instruments), len(d.included_instruments)])
    for i, ins in enumerate(instruments):
        enc[i, ins] = 1
    programs = [int(x[0]) for x in d.instruments_one_hot_encoder.inverse_transform(enc)] if len(instruments) else []
    mode = 0 if int(np.argmax(mode_logits)) == 0 else 1
    return programs, unpitched, mode, tempo, d.encode_instruments(programs)


def test_select_matches_reference_statements(torch):
    from mst_native import abi, lib
    from style import style_transfer as st
    rng = np.random.default_rng(11)
    S, max_c = 96, 16
    logits = rng.normal(size=(S, 41)).astype(np.float32)
    n_inst = rng.integers(1, 13, size=S)
    n_inst[:8] = 1
    logits[:4, 40] = 10.0                                    # lone pick = percussion: one more instrument is taken
    logits[8:16, 40] = 10.0                                  # percussion among several picks
    mode_logits = rng.normal(size=(S, 2)).astype(np.float32)
    mode_logits[20] = [0.25, 0.25]                           # tie: argmax takes the first (major)
    bpm = rng.uniform(50, 200, size=S).astype(np.float32)
    bpm[:6] = [60.5, 61.5, 120.5, 121.5, 99.5, 100.5]        # halves: Python's round() is half-to-even
    info = dict(ticks_per_beat=480, ticks_per_bar=1920, tempo=500000, tonic=0, mode=0, numerator=4, denominator_log2=2, n_beats=4)
    recs = [abi.make_decode_song(info, max_c, True, 8, write_back=False, pitched_offset=1000 * s, unpitched_offset=1000 * s + 7)
            for s in range(S)]
    d_songs = torch.from_numpy(np.array(recs, dtype=abi.DECODE_SONG_DTYPE).view(np.uint8).reshape(-1)).cuda()
    feats, choice = lib.transfer_select(torch.from_numpy(logits).cuda(), torch.from_numpy(mode_logits).cuda(),
                                        torch.from_numpy(bpm).cuda(), list(n_inst), st._PROGRAMS, st._GROUPS, max_c, d_songs)
    torch.cuda.synchronize()
    ch = choice.cpu().numpy().view(abi.TRANSFER_CHOICE_DTYPE)
    songs = d_songs.cpu().numpy().view(abi.DECODE_SONG_DTYPE)
    feats = feats.cpu().numpy()
    for s in range(S):
        programs, unpitched, mode, tempo, enc = reference_choice(logits[s], mode_logits[s], bpm[s], int(n_inst[s]))
        assert int(ch[s]['n_pitched']) == len(programs), s
        assert [int(x) for x in ch[s]['instruments'][:len(programs)]] == programs, s
        assert bool(ch[s]['has_unpitched']) == bool(unpitched) and int(ch[s]['mode']) == mode and int(ch[s]['tempo']) == tempo, s
        np.testing.assert_array_equal(feats[s, :len(programs)], enc.astype(np.float32))
        assert not feats[s, len(programs):].any()
        d = songs[s]
        assert (int(d['n_pitched']), int(d['has_unpitched']), int(d['mode']), int(d['tempo'])) == (min(len(programs), 15), int(unpitched), mode, tempo)
        assert [int(x) for x in d['instrument_id'][:len(programs)]] == programs
        assert [int(x) for x in d['channel_id'][:15]] == [i for i in range(16) if i != 9]
        assert (int(d['pitched_offset']), int(d['unpitched_offset']), int(d['n_bars'])) == (1000 * s, 1000 * s + 7, 8)   # untouched
