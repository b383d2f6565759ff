"""Seeded synthetic MIDI corpora as MstEvent batches (BASELINE.json configs[2]).

Per song: format-1 layout = one conductor track (time_signature 4/4 + set_tempo, 60-180 bpm) and
`n_tracks` instrument tracks: n_tracks-1 pitched channels (programs drawn from the reference's
popular_instruments, one program_change + one CC7 in [64,127] at tick 0) and channel 9 drums.
Pitches uniform 0..127 (drums 27..87) so out-of-range rows are exercised; onsets on the 1/16-note
grid with 10 % triplet onsets and +-3 tick jitter on 20 %; geometric durations (mean one beat);
5 % retriggers before the note-off; 1 % unclosed notes; velocities 1..127.
"""
import numpy as np

from . import abi

POPULAR = [0, 25, 48, 33, 1, 27, 49, 29, 35, 30, 50, 24, 5, 4, 32, 52, 26, 18, 28, 89, 65, 53, 61, 2, 17,
           73, 54, 62, 16, 39, 34, 51, 90, 56, 66, 38, 11, 81, 3, 57]


def synth_corpus(n_songs, seed=1234, tpb=480, n_bars=128, n_tracks=16, notes_per_bar=8):
    """Returns (events[EVENT_DTYPE], songs[SONG_IN_DTYPE]) for n_songs songs."""
    rng = np.random.default_rng(seed)
    S, T, N = n_songs, n_tracks, n_bars * notes_per_bar
    tpbs = np.broadcast_to(np.asarray(tpb, dtype=np.int64), (S,)).reshape(S, 1, 1)

    g16 = rng.integers(0, n_bars * 16, size=(S, T, N)) * tpbs // 4
    g3 = rng.integers(0, n_bars * 12, size=(S, T, N)) * tpbs // 3
    onset = np.where(rng.random((S, T, N)) < 0.1, g3, g16)
    onset = onset + np.where(rng.random((S, T, N)) < 0.2, rng.integers(-3, 4, size=(S, T, N)), 0)
    onset = np.clip(onset, 0, None)
    dur = rng.geometric(1.0 / np.broadcast_to(tpbs, (S, T, N)))
    pitch = rng.integers(0, 128, size=(S, T, N))
    pitch[:, T - 1, :] = rng.integers(27, 88, size=(S, N))
    vel = rng.integers(1, 128, size=(S, T, N))
    unclosed = rng.random((S, T, N)) < 0.01
    retrig = rng.random((S, T, N)) < 0.05
    big = np.int64(1) << 40

    # per track: [program_change, cc7, N note_on, N note_off, N retrigger note_on], sorted stably by tick
    M = 2 + 3 * N
    tick = np.empty((S, T, M), dtype=np.int64)
    kind = np.empty((S, T, M), dtype=np.uint8)
    d1 = np.empty((S, T, M), dtype=np.uint8)
    d2 = np.empty((S, T, M), dtype=np.uint8)
    tick[:, :, 0] = 0
    kind[:, :, 0] = abi.EV_PROGRAM_CHANGE
    progs = np.array(POPULAR)[rng.integers(0, len(POPULAR), size=(S, T))]
    d1[:, :, 0] = progs
    d2[:, :, 0] = 0
    tick[:, :, 1] = 0
    kind[:, :, 1] = abi.EV_CONTROL_CHANGE
    d1[:, :, 1] = 7
    d2[:, :, 1] = rng.integers(64, 128, size=(S, T))
    tick[:, T - 1, 0:2] = big                      # the drums track carries no program / volume
    a, b, c = 2, 2 + N, 2 + 2 * N
    tick[:, :, a:b] = onset
    kind[:, :, a:b] = abi.EV_NOTE_ON
    d1[:, :, a:b] = pitch
    d2[:, :, a:b] = vel
    tick[:, :, b:c] = np.where(unclosed, big, onset + dur)
    kind[:, :, b:c] = abi.EV_NOTE_OFF
    d1[:, :, b:c] = pitch
    d2[:, :, b:c] = 0
    tick[:, :, c:] = np.where(retrig, onset + dur // 2, big)
    kind[:, :, c:] = abi.EV_NOTE_ON
    d1[:, :, c:] = pitch
    d2[:, :, c:] = rng.integers(1, 128, size=(S, T, N))

    order = np.argsort(tick, axis=2, kind='stable')
    tick = np.take_along_axis(tick, order, 2)
    kind = np.take_along_axis(kind, order, 2)
    d1 = np.take_along_axis(d1, order, 2)
    d2 = np.take_along_axis(d2, order, 2)
    valid = tick < big

    chan = np.broadcast_to(np.concatenate([np.arange(T - 1) + (np.arange(T - 1) >= 9), [9]]).astype(np.uint8)
                           .reshape(1, T, 1), (S, T, M))
    track = np.broadcast_to((np.arange(T) + 1).astype(np.uint16).reshape(1, T, 1), (S, T, M))

    per_song = valid.reshape(S, -1).sum(1) + 2            # + conductor track
    songs = np.zeros(S, dtype=abi.SONG_IN_DTYPE)
    songs['n_events'] = per_song
    songs['event_offset'] = np.concatenate([[0], np.cumsum(per_song)[:-1]])
    songs['ticks_per_beat'] = np.broadcast_to(np.asarray(tpb), (S,))
    events = np.zeros(int(per_song.sum()), dtype=abi.EVENT_DTYPE)

    # conductor track first (track 0), then the instrument tracks in order
    pos0 = songs['event_offset'].astype(np.int64)
    events['kind'][pos0] = abi.EV_TIME_SIGNATURE
    events['d1'][pos0] = 4
    events['d2'][pos0] = 2
    events['kind'][pos0 + 1] = abi.EV_SET_TEMPO
    bpm = rng.integers(60, 181, size=S)
    events['aux'][pos0 + 1] = (60_000_000 // bpm).astype(np.uint32)
    body = np.ones(len(events), dtype=bool)
    body[pos0] = False
    body[pos0 + 1] = False
    events['tick'][body] = tick[valid]
    events['kind'][body] = kind[valid]
    events['channel'][body] = chan[valid]
    events['d1'][body] = d1[valid]
    events['d2'][body] = d2[valid]
    events['track'][body] = track[valid]
    return events, songs


def split_songs(events, songs):
    """-> list of per-song event arrays (views)."""
    return [events[int(s['event_offset']):int(s['event_offset']) + int(s['n_events'])] for s in songs]


# ---------------------------------------------------------------------------------- training workload
def build_model(seed=108):
    """The network train-model.py:52-85 creates: torch.manual_seed(seed), then the nine sub-modules in the script's
    construction order (it fixes the RNG stream, hence the weights), through the drop-in `style.model` classes."""
    import torch
    from style import model as m, data as d
    torch.manual_seed(seed)
    pce = m.PitchedChannelsEncoder(64, 128, d.instrument_size)
    uce = m.UnpitchedChannelsEncoder(64, 128)
    pre = m.PitchedRhythmEncoder(32, 64, 128, d.instrument_size)
    ure = m.UnpitchedRhythmEncoder(32, 64, 128)
    se = m.StyleEncoder(256, 128, d.instrument_size)
    me = m.MelodyEncoder(8, 64, 128, d.instrument_size)
    sim = m.SongInfoModel(8, 256, 32, d.n_instruments)
    psa = m.PitchedStyleApplier(256, 8, 32, d.instrument_size)
    usa = m.UnpitchedStyleApplier(256, 32)
    return m.StyleTransferModel(pce, uce, se, me, pre, ure, sim, psa, usa)


def make_state_dict(seed=108):
    """state_dict (CPU tensors, reference names) of build_model(seed)."""
    return {k: v.detach().clone() for k, v in build_model(seed).state_dict().items()}


def synthetic_sample(seed, C=4, B=200, nb=4, drums=True, density=0.003):
    """One synthetic training sample (CPU tensors with a leading batch dim of 1): Bernoulli(density) occupancy, velocity
    U(0,1], duration U(0,4], one-hot accidentals; instruments = the first C sorted popular programs; mode major; bpm 120."""
    import torch
    from style import data as d
    g = torch.Generator().manual_seed(seed)
    p = torch.zeros(1, C, B, nb, 10, 56, 5)
    occ = torch.rand(p.shape[:-1], generator=g) < density
    n = int(occ.sum())
    vals = torch.zeros(n, 5)
    vals[:, 0] = torch.rand(n, generator=g) * 4 + 1e-3
    vals[:, 1] = torch.rand(n, generator=g) * 0.999 + 1e-3
    vals[torch.arange(n), 2 + torch.randint(0, 3, (n,), generator=g)] = 1.0
    p[occ] = vals
    u = None
    if drums:
        u = torch.zeros(1, 1, B, nb, 10, 47, 2)
        occ = torch.rand(u.shape[:-1], generator=g) < density
        n = int(occ.sum())
        uv = torch.zeros(n, 2)
        uv[:, 0] = torch.rand(n, generator=g) * 4 + 1e-3
        uv[:, 1] = torch.rand(n, generator=g) * 0.999 + 1e-3
        u[occ] = uv
    feats = torch.as_tensor(d.encode_instruments(sorted(POPULAR)[:C]), dtype=torch.float32).unsqueeze(0)
    used = d.get_used_instruments(feats, u)
    return dict(mode=torch.tensor([[1., 0.]]), bpm=torch.tensor([120.]), pitched=p, feats=feats, unpitched=u, bpm_target=120,
                used=used)
