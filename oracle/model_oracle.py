"""Plain-PyTorch fp32 restatement of the reference's style-transfer network and losses.

TEST INFRASTRUCTURE ONLY (the checker for the CUDA training step); imported from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never from the product.

Functional style: every function takes the reference's state_dict (same parameter names and shapes as
style/model.py's nn.Modules, so `reference_model.state_dict()` plugs in directly) and tensors with the
reference's layouts:
    pitched   (1, C, B, nb, 10, 56, 5)     unpitched (1, 1, B, nb, 10, 47, 2) or None
    instruments_features (1, C, 51)        mode (1, 2)        bpm (1,)
Pinned against the unmodified reference (style/model.py imported from /root/reference, same weights and
inputs): tests/test_oracle_model.py (refcheck) and the committed goldens tests/golden/model_*.npz.
"""
import torch
import torch.nn.functional as F

EPSILON = 1e-7            # style/model.py:11
BPM_MIN, BPM_RANGE = 50.0, 150.0   # :22-25


def leaky(x):
    return F.leaky_relu(x)   # slope 0.01 everywhere in the reference


def linear(sd, name, x):
    return F.linear(x, sd[name + '.weight'], sd[name + '.bias'])


def lstm(sd, prefix, x, reverse=False, suffix=''):
    """One direction of nn.LSTM (batch_first, 1 layer): x (N, T, I) -> all hidden states (N, T, H).
    Gate order i, f, g, o (torch)."""
    w_ih, w_hh = sd[f'{prefix}.weight_ih_l0{suffix}'], sd[f'{prefix}.weight_hh_l0{suffix}']
    b = sd[f'{prefix}.bias_ih_l0{suffix}'] + sd[f'{prefix}.bias_hh_l0{suffix}']
    n, t_len, _ = x.shape
    hid = w_hh.shape[1]
    h = x.new_zeros(n, hid)
    c = x.new_zeros(n, hid)
    xs = x @ w_ih.t() + b
    outs = [None] * t_len
    order = range(t_len - 1, -1, -1) if reverse else range(t_len)
    for t in order:
        g = xs[:, t] + h @ w_hh.t()
        i, f, gg, o = g.chunk(4, dim=1)
        c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
        h = torch.sigmoid(o) * torch.tanh(c)
        outs[t] = h
    return torch.stack(outs, 1)


def bilstm(sd, prefix, x):
    return torch.cat([lstm(sd, prefix, x), lstm(sd, prefix, x, reverse=True, suffix='_reverse')], -1)


def combine(x, dim):
    """style/model.py:796-815: mean over `dim` weighted by sqrt(1 + sum of squares over all other dims)."""
    dims = [i for i in range(x.dim()) if i != dim]
    norm = torch.sqrt(1.0 + (x ** 2).sum(dims, keepdim=True))
    return (x * norm).sum(dim) / norm.sum()


def combine_pair(a, b):
    return combine(torch.stack([a, b]), 0)


# ------------------------------------------------------------------------------------ encoders
def pitched_channels_encoder(sd, p, pitched, feats):
    """style/model.py:77-99 -> beats (1,C,B,nb,64), bars (1,B,128)."""
    _, C, B, nb = pitched.shape[:4]
    x = pitched.transpose(-1, -2).reshape(C * B * nb, 50, 56)            # (rows, frac*feat, note)
    x = F.conv1d(x, sd[p + '.beats_conv.module.weight'], sd[p + '.beats_conv.module.bias'], stride=7, padding=4)
    x1 = leaky(x).reshape(1, C, B, nb, -1)                                # (57, 8) flattened channel-major
    x2 = leaky(linear(sd, p + '.instruments_linear', feats)).view(1, C, 1, 1, -1).expand(1, C, B, nb, -1)
    x = leaky(linear(sd, p + '.linear', torch.cat([x1, x2], -1)))
    beats = lstm(sd, p + '.beats_lstm.module', x.reshape(C * B, nb, -1)).view(1, C, B, nb, -1)
    bars_in = combine(beats[:, :, :, -1], 1)                              # (1, B, 64)
    return beats, bilstm(sd, p + '.bars_lstm', bars_in)


def unpitched_channels_encoder(sd, p, unpitched):
    """style/model.py:128-141."""
    _, U, B, nb = unpitched.shape[:4]
    x = unpitched.transpose(-1, -2).reshape(1, U, B, nb, -1)              # (frac, feat, note) flattened
    x = leaky(linear(sd, p + '.linear', x))
    beats = lstm(sd, p + '.beats_lstm.module', x.reshape(U * B, nb, -1)).view(1, U, B, nb, -1)
    bars_in = combine(beats[:, :, :, -1], 1)
    return beats, bilstm(sd, p + '.bars_lstm', bars_in)


def pitched_rhythm_encoder(sd, p, beats, bars, pitched, feats, mode, bpm):
    """style/model.py:346-381 -> (1, B, nb, 10, 32)."""
    _, C, B, nb = pitched.shape[:4]
    shape = (1, C, B, nb, 10)
    x1 = leaky(linear(sd, p + '.beats_linear', beats)).unsqueeze(4).expand(*shape, -1)
    x2 = leaky(linear(sd, p + '.bars_linear', bars)).view(1, 1, B, 1, 1, -1).expand(*shape, -1)
    x3 = leaky(linear(sd, p + '.channels_linear', pitched.reshape(*shape, -1)))
    x4 = leaky(linear(sd, p + '.instruments_linear', feats)).view(1, C, 1, 1, 1, -1).expand(*shape, -1)
    x5 = leaky(linear(sd, p + '.mode_linear', mode)).view(1, 1, 1, 1, 1, -1).expand(*shape, -1)
    x6 = leaky(linear(sd, p + '.bpm_linear', bpm.unsqueeze(-1))).view(1, 1, 1, 1, 1, -1).expand(*shape, -1)
    x = leaky(linear(sd, p + '.linear', torch.cat([x1, x2, x3, x4, x5, x6], -1)))
    return combine(x, 1)


def unpitched_rhythm_encoder(sd, p, beats, bars, unpitched, bpm):
    """style/model.py:418-443."""
    _, U, B, nb = unpitched.shape[:4]
    shape = (1, U, B, nb, 10)
    x1 = leaky(linear(sd, p + '.beats_linear', beats)).unsqueeze(4).expand(*shape, -1)
    x2 = leaky(linear(sd, p + '.bars_linear', bars)).view(1, 1, B, 1, 1, -1).expand(*shape, -1)
    x3 = leaky(linear(sd, p + '.channels_linear', unpitched.reshape(*shape, -1)))
    x4 = leaky(linear(sd, p + '.bpm_linear', bpm.unsqueeze(-1))).view(1, 1, 1, 1, 1, -1).expand(*shape, -1)
    x = leaky(linear(sd, p + '.linear', torch.cat([x1, x2, x3, x4], -1)))
    return combine(x, 1)


def style_encoder(sd, p, bars, feats, mode, bpm):
    """style/model.py:179-200 -> (1, 256)."""
    C = feats.shape[1]
    x1 = lstm(sd, p + '.bars_lstm', bars)[:, -1].unsqueeze(1).expand(1, C, -1)
    x2 = leaky(linear(sd, p + '.instruments_linear', feats))
    x3 = leaky(linear(sd, p + '.mode_linear', mode)).unsqueeze(1).expand(1, C, -1)
    x4 = leaky(linear(sd, p + '.bpm_linear', bpm.unsqueeze(-1))).unsqueeze(1).expand(1, C, -1)
    x = leaky(linear(sd, p + '.linear', torch.cat([x1, x2, x3, x4], -1)))
    return combine(x, 1)


def _octave_degree(sd, p, y, n_feat):
    """octave_linear / scale_degree_linear + outer add (model.py:270-286, 644-660) -> (..., 56, n_feat)."""
    o = leaky(linear(sd, p + '.octave_linear', y))
    o = o.view(*o.shape[:-1], 8, 1, n_feat)
    d = leaky(linear(sd, p + '.scale_degree_linear', y))
    d = d.view(*d.shape[:-1], 1, 7, n_feat)
    x = leaky(o + d)
    return x.reshape(*x.shape[:-3], 56, n_feat)


def melody_encoder(sd, p, beats, bars, pitched, feats):
    """style/model.py:252-297 -> (1, B, nb, 10, 56, 8)."""
    _, C, B, nb = pitched.shape[:4]
    shape = (1, C, B, nb, 10)
    x1 = leaky(linear(sd, p + '.beats_linear', beats)).unsqueeze(-2).expand(*shape, -1)
    x2 = leaky(linear(sd, p + '.bars_linear', bars)).view(1, 1, B, 1, 1, -1).expand(*shape, -1)
    x3 = leaky(linear(sd, p + '.instruments_linear', feats)).view(1, C, 1, 1, 1, -1).expand(*shape, -1)
    y = torch.cat([x1, x2, x3], -1)
    a = _octave_degree(sd, p, y, 8)
    b = leaky(linear(sd, p + '.channels_linear', pitched))
    x = leaky(linear(sd, p + '.linear', torch.cat([a, b], -1)))
    return combine(x, 1)


def song_info_model(sd, p, style, rhythm):
    """style/model.py:513-562 -> (instruments (1,41), mode (1,2), bpm (1,))."""
    _, B, nb = rhythm.shape[:3]
    x = lstm(sd, p + '.beats_lstm.module', rhythm.reshape(B, nb, -1))[:, -1].unsqueeze(0)   # (1, B, 9)
    rf = lstm(sd, p + '.bars_lstm', x)[:, -1]                                                 # (1, 8)

    def head(name):
        a = leaky(linear(sd, f'{p}.style_{name}_linear', style))
        b = leaky(linear(sd, f'{p}.rhythm_{name}_linear', rf))
        return linear(sd, f'{p}.{name}_linear', torch.cat([a, b], -1))

    bpm = torch.sigmoid(head('bpm')[:, 0]) * BPM_RANGE + BPM_MIN
    return head('instruments'), head('mode'), bpm


def pitched_style_applier(sd, p, style, melody, rhythm, feats):
    """style/model.py:624-675 -> (1, C, B, nb, 10, 56, 5)."""
    C = feats.shape[1]
    _, B, nb = rhythm.shape[:3]
    shape = (1, C, B, nb, 10)
    x1 = leaky(linear(sd, p + '.style_linear', style)).view(1, 1, 1, 1, 1, -1).expand(*shape, -1)
    x2 = leaky(linear(sd, p + '.rhythm_linear', rhythm)).unsqueeze(1).expand(*shape, -1)
    x3 = leaky(linear(sd, p + '.instruments_linear', feats)).view(1, C, 1, 1, 1, -1).expand(*shape, -1)
    y = torch.cat([x1, x2, x3], -1)
    a = _octave_degree(sd, p, y, 30)
    b = leaky(linear(sd, p + '.melody_linear', melody)).unsqueeze(1).expand(*shape, 56, -1)
    x = linear(sd, p + '.linear', torch.cat([a, b], -1))
    return torch.cat([torch.sigmoid(x[..., :1]) * 6, torch.sigmoid(x[..., 1:])], -1)


def unpitched_style_applier(sd, p, style, rhythm):
    """style/model.py:703-724 -> (1, 1, B, nb, 10, 47, 2)."""
    _, B, nb = rhythm.shape[:3]
    x1 = leaky(linear(sd, p + '.style_linear', style)).view(1, 1, 1, 10, -1).expand(1, B, nb, 10, -1)
    x2 = leaky(linear(sd, p + '.rhythm_linear', rhythm))
    x = leaky(linear(sd, p + '.notes_linear', torch.cat([x1, x2], -1)))
    x = linear(sd, p + '.linear', x.view(1, B, nb, 10, 47, -1))
    return torch.cat([torch.sigmoid(x[..., :1]) * 6, torch.sigmoid(x[..., 1:])], -1).unsqueeze(1)


# ------------------------------------------------------------------------------------ whole model
def extract_style(sd, mode, bpm, pitched, feats, unpitched=None):
    """StyleTransferModel.extract_style (model.py:751-773)."""
    p_beats, p_bars = pitched_channels_encoder(sd, 'pitched_channels_encoder', pitched, feats)
    p_rhythm = pitched_rhythm_encoder(sd, 'pitched_rhythm_encoder', p_beats, p_bars, pitched, feats, mode, bpm)
    if unpitched is None:
        bars, rhythm = p_bars, p_rhythm
    else:
        u_beats, u_bars = unpitched_channels_encoder(sd, 'unpitched_channels_encoder', unpitched)
        u_rhythm = unpitched_rhythm_encoder(sd, 'unpitched_rhythm_encoder', u_beats, u_bars, unpitched, bpm)
        bars = combine_pair(p_bars, u_bars)
        rhythm = combine_pair(p_rhythm, u_rhythm)
    style = style_encoder(sd, 'style_encoder', bars, feats, mode, bpm)
    melody = melody_encoder(sd, 'melody_encoder', p_beats, p_bars, pitched, feats)
    return style, melody, rhythm


def apply_style(sd, style, melody, rhythm, feats, unpitched=False):
    x_p = pitched_style_applier(sd, 'pitched_style_applier', style, melody, rhythm, feats)
    x_u = unpitched_style_applier(sd, 'unpitched_style_applier', style, rhythm) if unpitched else None
    return x_p, x_u


def forward(sd, mode, bpm, pitched, feats, unpitched=None):
    """StyleTransferModel.forward (model.py:784-793)."""
    style, melody, rhythm = extract_style(sd, mode, bpm, pitched, feats, unpitched)
    info = song_info_model(sd, 'song_info_model', style, rhythm)
    x_p, x_u = apply_style(sd, style, melody, rhythm, feats, unpitched is not None)
    return info, x_p, x_u


# ------------------------------------------------------------------------------------ losses
def safe_div(num, den):
    """model.py:854-860."""
    if den.abs() < EPSILON:
        den = den - EPSILON if den < 0 else den + EPSILON
    return num / den


def quadratic_mean(tensors, weights=None):
    """get_mean(..., mean_type='quadratic') (utils/pytorch.py:74-94) with safe_sqrt (:68)."""
    n = len(tensors)
    if weights is None:
        weights = [1.0 / n] * n
    x = sum(w * t ** 2 for t, w in zip(tensors, weights))
    if x == 0:
        return torch.tensor(0., requires_grad=x.requires_grad) * x
    return torch.sqrt(x)


def channels_losses(pred, target, pitched=True):
    """get_channels_losses (model.py:909-921)."""
    tv, v = target[..., 1], pred[..., 1]
    mask = (tv > 0).float()
    tp, fp, fn = torch.min(v, tv).sum(), torch.relu(v - tv).sum(), torch.relu(tv - v).sum()
    precision, recall = safe_div(tp, tp + fp), safe_div(tp, tp + fn)
    notes = 1.0 - 2 * safe_div(precision * recall, precision + recall)
    n = mask.sum()
    velocity = (((tv - v) ** 2) * mask).sum() / n
    duration = ((((pred[..., 0] - target[..., 0].clamp(max=6)) / 6) ** 2) * mask).sum() / n
    if not pitched:
        return notes, velocity, duration
    bce = F.binary_cross_entropy(pred[..., 2:], target[..., 2:], reduction='none')
    accidentals = (bce * mask.unsqueeze(-1)).sum() / (n * 3)
    return notes, velocity, duration, accidentals


def combine_losses(notes, velocity, duration, accidentals=None):
    """model.py:924-932 (the weights carry gradient)."""
    notes2 = quadratic_mean([notes, velocity], [notes, 1 - notes])
    if accidentals is not None:
        return quadratic_mean([duration, accidentals, notes2])
    return quadratic_mean([duration, notes2])


def total_loss(instruments_pred, instruments_target, mode_pred, mode_target, bpm_pred, bpm_target,
               pitched_pred, pitched_target, unpitched_pred=None, unpitched_target=None, normalize=False):
    """get_total_loss (model.py:935-997) including its positional quirk: the caller passes (bpm_pred, bpm)
    in the mode_* slots and (mode_pred, mode) in the bpm_* slots (train-model.py:115-122), and
    get_song_info_loss unpacks (instruments, bpm, mode) from (instruments, mode, bpm) (:900-901)."""
    notes, velocity, duration, accidentals = channels_losses(pitched_pred, pitched_target)
    if normalize:
        accidentals = torch.tanh(accidentals)
    pitched = combine_losses(notes, velocity, duration, accidentals)
    out_p = dict(total=pitched, notes_loss=notes, velocity_loss=velocity, duration_loss=duration,
                 accidentals_loss=accidentals)
    if unpitched_target is not None:
        un, uv, ud = channels_losses(unpitched_pred, unpitched_target, pitched=False)
        unp = combine_losses(un, uv, ud)
        out_u = dict(total=unp, notes_loss=un, velocity_loss=uv, duration_loss=ud)
        channels = quadratic_mean([pitched, unp])
    else:
        out_u, channels = None, pitched
    in_instr, in_bpm, in_mode = instruments_pred, mode_pred, bpm_pred      # the unpack at :900
    t_instr, t_bpm, t_mode = instruments_target, mode_target, bpm_target   # and :901
    instruments = F.binary_cross_entropy_with_logits(in_instr, t_instr)
    mode = F.cross_entropy(in_mode, t_mode.argmax(1))
    bpm = ((in_bpm - t_bpm) / BPM_RANGE) ** 2
    if normalize:
        instruments, mode = torch.tanh(instruments), torch.tanh(mode)
    song = quadratic_mean([instruments, mode, bpm])
    return dict(total=quadratic_mean([channels, song]),
                channels_loss=dict(total=channels, pitched=out_p, unpitched=out_u),
                song_info_loss=dict(total=song, instruments_loss=instruments, mode_loss=mode, bpm_loss=bpm))


def used_instruments(feats, unpitched):
    """get_used_instruments (style/data.py:159-169) -> (1, 41)."""
    used = (feats[:, :, :40].sum(1) > 0).float()
    perc = torch.full((1, 1), 1.0 if unpitched is not None else 0.0, dtype=torch.float32, device=feats.device)
    return torch.cat([used, perc], 1)


def training_losses(sd, mode, bpm, pitched, feats, unpitched, bpm_target):
    """One iteration of train-model.py:111-122 (forward + get_total_loss with normalize=True)."""
    (instr, mode_pred, bpm_pred), x_p, x_u = forward(sd, mode, bpm, pitched, feats, unpitched)
    return total_loss(instr, used_instruments(feats, unpitched), bpm_pred, bpm_target, mode_pred, mode,
                      x_p, pitched, x_u, unpitched, normalize=True), (instr, mode_pred, bpm_pred, x_p, x_u)


# ------------------------------------------------------------------------------------ synthetic inputs
def synthetic_sample(seed, C=4, B=200, nb=4, drums=True, density=0.003, device='cpu'):
    """Bernoulli(density) occupancy, velocity U(0,1], duration U(0,4], one-hot
    accidentals; instruments = first C of sorted popular_instruments; mode major; bpm 120."""
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
    programs = sorted([0, 25, 48, 33, 1, 27, 49, 29, 35, 30, 50, 24, 5, 4, 32, 52, 26, 18, 28, 89, 65, 53, 61, 2,
                       17, 73, 54, 62, 16, 39, 34, 51, 90, 56, 66, 38, 11, 81, 3, 57])
    feats = instrument_features(programs[:C])
    mode = torch.tensor([[1., 0.]])
    bpm = torch.tensor([120.])
    to = lambda t: t.to(device) if t is not None else None
    return dict(mode=to(mode), bpm=to(bpm), pitched=to(p), feats=to(feats), unpitched=to(u), bpm_target=120)


_GROUPS = ['Piano', 'Chromatic Percussion', 'Organ', 'Guitar', 'Bass', 'Strings', 'Ensemble', 'Brass', 'Reed', 'Pipe',
           'Synth Lead', 'Synth Pad', 'Synth Effects', 'Ethnic', 'Percussive', 'Sound Effects']   # GM families of 8


def instrument_features(programs):
    """encode_instruments (style/data.py:122-127): one-hot over the 40 sorted popular programs ++ one-hot over
    their 11 sorted group names -> (1, C, 51)."""
    popular = sorted([0, 25, 48, 33, 1, 27, 49, 29, 35, 30, 50, 24, 5, 4, 32, 52, 26, 18, 28, 89, 65, 53, 61, 2, 17,
                      73, 54, 62, 16, 39, 34, 51, 90, 56, 66, 38, 11, 81, 3, 57])
    groups = sorted({_GROUPS[q // 8] for q in popular})
    x = torch.zeros(1, len(programs), 51)
    for i, q in enumerate(programs):
        x[0, i, popular.index(q)] = 1.0
        x[0, i, 40 + groups.index(_GROUPS[q // 8])] = 1.0
    return x


# ------------------------------------------------------------------------------------ weights
def _ms(*values, factor=1):
    """get_mean_size (style/model.py:31-33)."""
    import math
    return math.ceil(sum(values) / len(values) * factor)


def layer_specs(beat=64, bar=128, n_rhythm=8, style=256, melody=8, rhythm=32, instr=51, n_instr=41):
    """(state-dict prefix, kind, args) of every layer in the order train-model.py:62-85 constructs them;
    sizes follow the constructors in style/model.py (hard-coded values of train-model.py:54-60)."""
    L, S = [], lambda *a: L.append(a)
    p = 'pitched_channels_encoder'                                       # model.py:36-75
    c_out = _ms(50, beat)
    i_sz = _ms(instr, beat)
    S(p + '.beats_conv.module', 'conv', 50, c_out, 14)
    S(p + '.instruments_linear', 'lin', instr, i_sz)
    S(p + '.linear', 'lin', c_out * 8 + i_sz, beat)
    S(p + '.beats_lstm.module', 'lstm', beat, beat, False)
    S(p + '.bars_lstm', 'lstm', beat, bar // 2, True)
    p = 'unpitched_channels_encoder'                                     # :102-126
    S(p + '.linear', 'lin', 940, beat)
    S(p + '.beats_lstm.module', 'lstm', beat, beat, False)
    S(p + '.bars_lstm', 'lstm', beat, bar // 2, True)
    p = 'pitched_rhythm_encoder'                                         # :301-344
    sizes = [_ms(beat, rhythm), _ms(bar, rhythm, factor=.5), _ms(280, rhythm, factor=.1), _ms(instr, rhythm, factor=.5),
             _ms(2, rhythm, factor=.25), _ms(1, rhythm, factor=.25)]
    for name, i, o in zip(['beats', 'bars', 'channels', 'instruments', 'mode', 'bpm'], [beat, bar, 280, instr, 2, 1], sizes):
        S(f'{p}.{name}_linear', 'lin', i, o)
    S(p + '.linear', 'lin', sum(sizes), rhythm)
    p = 'unpitched_rhythm_encoder'                                       # :384-416
    sizes = [_ms(beat, rhythm), _ms(bar, rhythm, factor=.5), _ms(94, rhythm, factor=.25), _ms(1, rhythm, factor=.25)]
    for name, i, o in zip(['beats', 'bars', 'channels', 'bpm'], [beat, bar, 94, 1], sizes):
        S(f'{p}.{name}_linear', 'lin', i, o)
    S(p + '.linear', 'lin', sum(sizes), rhythm)
    p = 'style_encoder'                                                  # :144-177
    sizes = [_ms(bar, style), _ms(instr, style, factor=.25), _ms(2, style, factor=.1), _ms(style, 1, factor=.05)]
    S(p + '.bars_lstm', 'lstm', bar, sizes[0], False)
    S(p + '.instruments_linear', 'lin', instr, sizes[1])
    S(p + '.mode_linear', 'lin', 2, sizes[2])
    S(p + '.bpm_linear', 'lin', 1, sizes[3])
    S(p + '.linear', 'lin', sum(sizes), style)
    p = 'melody_encoder'                                                 # :203-250
    sizes = [_ms(beat, melody), _ms(bar, melody), _ms(instr, melody, factor=.25)]
    ch = _ms(5, melody)
    S(p + '.beats_linear', 'lin', beat, sizes[0])
    S(p + '.bars_linear', 'lin', bar, sizes[1])
    S(p + '.instruments_linear', 'lin', instr, sizes[2])
    S(p + '.octave_linear', 'lin', sum(sizes), melody * 8)
    S(p + '.scale_degree_linear', 'lin', sum(sizes), melody * 7)
    S(p + '.channels_linear', 'lin', 5, ch)
    S(p + '.linear', 'lin', melody + ch, melody)
    p = 'song_info_model'                                                # :446-511
    bl = _ms(10 * rhythm, n_rhythm, factor=.05)
    S(p + '.beats_lstm.module', 'lstm', 10 * rhythm, bl, False)
    S(p + '.bars_lstm', 'lstm', bl, n_rhythm, False)
    for name, fs, fr, out in [('instruments', .05, .25, n_instr), ('mode', .01, .1, 2), ('bpm', .01, .1, 1)]:
        tgt = n_instr if name == 'instruments' else (2 if name == 'mode' else 1)
        a, b = _ms(style, tgt, factor=fs), _ms(rhythm, tgt, factor=fr)
        S(f'{p}.style_{name}_linear', 'lin', style, a)
        S(f'{p}.rhythm_{name}_linear', 'lin', n_rhythm, b)
        S(f'{p}.{name}_linear', 'lin', a + b, out)
    p = 'pitched_style_applier'                                          # :582-622
    sizes = [_ms(style, 5, factor=.5), _ms(rhythm, 5, factor=.5), _ms(instr, 5, factor=.4)]
    ml = _ms(melody, 5, factor=3)
    S(p + '.style_linear', 'lin', style, sizes[0])
    S(p + '.rhythm_linear', 'lin', rhythm, sizes[1])
    S(p + '.instruments_linear', 'lin', instr, sizes[2])
    S(p + '.octave_linear', 'lin', sum(sizes), 30 * 8)
    S(p + '.scale_degree_linear', 'lin', sum(sizes), 30 * 7)
    S(p + '.melody_linear', 'lin', melody, ml)
    S(p + '.linear', 'lin', 30 + ml, 5)
    p = 'unpitched_style_applier'                                        # :678-701
    sl, rl = _ms(style, 2, factor=.5), _ms(rhythm, 2, factor=1)
    S(p + '.style_linear', 'lin', style, 10 * sl)
    S(p + '.rhythm_linear', 'lin', rhythm, rl)
    S(p + '.notes_linear', 'lin', sl + rl, 47 * 8)
    S(p + '.linear', 'lin', 8, 2)
    return L


def make_state_dict(seed=108, device='cpu'):
    """The weights train-model.py:52-85 creates: torch.manual_seed(seed), then torch's default initialisers
    for every layer in construction order (so the RNG stream is consumed identically).  NOTE the reference's
    state_dict key order follows StyleTransferModel.__init__ (model.py:737-749), not construction order."""
    import torch.nn as nn
    torch.manual_seed(seed)
    sd = {}
    for prefix, kind, *a in layer_specs():
        if kind == 'conv':
            m = nn.Conv1d(a[0], a[1], kernel_size=a[2], stride=7, padding=4)
        elif kind == 'lin':
            m = nn.Linear(a[0], a[1])
        else:
            m = nn.LSTM(input_size=a[0], hidden_size=a[1], num_layers=1, bidirectional=a[2], batch_first=True)
        for k, v in m.state_dict().items():
            sd[f'{prefix}.{k}'] = v.detach().to(device)
    return sd
