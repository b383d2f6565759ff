"""style/model.py of the reference with the network running in the fused CUDA kernels of libmst_b200.so.

Same classes, constructor signatures and parameter names as the reference (model.py:36-749), so
`torch.manual_seed(108)` + the construction order of train-model.py:62-85 gives bit-identical initial weights,
`model.state_dict()` / `torch.save(model)` keep working, and `torch.optim.Adam(model.parameters())` updates the very
buffer the kernels read: StyleTransferModel re-points every parameter (and its .grad) into ONE flat fp32 CUDA tensor
laid out by state-dict name as mst_model_param_* describes.

The reference's call pattern is kept (train-model.py:113-126):
    (instruments_pred, mode_pred, bpm_pred), pitched_pred, unpitched_pred = model(mode, bpm, pitched, feats, unpitched)
    losses = get_total_loss(...);  losses['total'].backward()
forward = mst_model_forward (activations kept in the workspace); get_total_loss = mst_model_loss on the predictions;
`.backward()` on the total = mst_model_backward on that workspace, which ADDS into the flat gradient buffer
(iter_size accumulation).  Only the hyper-parameters of train-model.py:54-60 are implemented (the kernels are
specialised for them); sub-modules cannot be called on their own.
"""
import math
import os

import torch
import torch.nn as nn

from mst_native import model as _mm

epsilon = 1e-7
n_note_features = 5
bpm_min, bpm_max = 50, 200
bpm_range = bpm_max - bpm_min
mean_type = 'quadratic'

device = 'cuda' if torch.cuda.is_available() else 'cpu'


def get_mean_size(*values, factor=1):
    return math.ceil(sum(values) / len(values) * factor)


class LSTM(nn.LSTM):
    """utils/pytorch.py:19 (batch-first LSTM); only its parameters are used here."""


class Distributed(nn.Module):
    """utils/pytorch.py:28 wrapper; keeps the `.module.` level of the reference's parameter names."""

    def __init__(self, module, depth=1):
        super().__init__()
        self.module = module
        self.depth = depth


class _Part(nn.Module):
    """A sub-module of the network: owns parameters with the reference's names; the computation is fused."""

    def forward(self, *args, **kwargs):
        raise NotImplementedError(f'{type(self).__name__} runs fused inside StyleTransferModel (forward / extract_style / '
                                  'predict_song_info / apply_style); it cannot be called on its own')

    @staticmethod
    def _check(**sizes):
        want = dict(beat_size=64, bar_size=128, instrument_size=51, rhythm_size=32, style_size=256, melody_size=8,
                    n_rhythm_features=8, n_instruments=41)
        for k, v in sizes.items():
            if v != want[k]:
                raise NotImplementedError(f'{k}={v}: the CUDA kernels implement train-model.py:54-60 ({k}={want[k]})')


def _lstm(i, h, bidirectional=False):
    return LSTM(input_size=i, hidden_size=h, num_layers=1, bidirectional=bidirectional, batch_first=True)


class PitchedChannelsEncoder(_Part):
    def __init__(self, beat_size, bar_size, instrument_size):
        super().__init__()
        self._check(beat_size=beat_size, bar_size=bar_size, instrument_size=instrument_size)
        c_out, i_sz = get_mean_size(50, beat_size), get_mean_size(instrument_size, beat_size)
        self.beats_conv = Distributed(nn.Conv1d(50, c_out, kernel_size=14, stride=7, padding=4), depth=3)
        self.instruments_linear = nn.Linear(instrument_size, i_sz)
        self.linear = nn.Linear(c_out * 8 + i_sz, beat_size)
        self.beats_lstm = Distributed(_lstm(beat_size, beat_size), depth=2)
        self.bars_lstm = _lstm(beat_size, bar_size // 2, True)


class UnpitchedChannelsEncoder(_Part):
    def __init__(self, beat_size, bar_size):
        super().__init__()
        self._check(beat_size=beat_size, bar_size=bar_size)
        self.linear = nn.Linear(940, beat_size)
        self.beats_lstm = Distributed(_lstm(beat_size, beat_size), depth=2)
        self.bars_lstm = _lstm(beat_size, bar_size // 2, True)


class StyleEncoder(_Part):
    def __init__(self, style_size, bar_size, instrument_size):
        super().__init__()
        self._check(style_size=style_size, bar_size=bar_size, instrument_size=instrument_size)
        sizes = [get_mean_size(bar_size, style_size), get_mean_size(instrument_size, style_size, factor=.25),
                 get_mean_size(2, style_size, factor=.1), get_mean_size(style_size, 1, factor=.05)]
        self.bars_lstm = _lstm(bar_size, sizes[0])
        self.instruments_linear = nn.Linear(instrument_size, sizes[1])
        self.mode_linear = nn.Linear(2, sizes[2])
        self.bpm_linear = nn.Linear(1, sizes[3])
        self.linear = nn.Linear(sum(sizes), style_size)


class MelodyEncoder(_Part):
    def __init__(self, melody_size, beat_size, bar_size, instrument_size):
        super().__init__()
        self._check(melody_size=melody_size, beat_size=beat_size, bar_size=bar_size, instrument_size=instrument_size)
        sizes = [get_mean_size(beat_size, melody_size), get_mean_size(bar_size, melody_size),
                 get_mean_size(instrument_size, melody_size, factor=.25)]
        ch = get_mean_size(n_note_features, melody_size)
        self.beats_linear = nn.Linear(beat_size, sizes[0])
        self.bars_linear = nn.Linear(bar_size, sizes[1])
        self.instruments_linear = nn.Linear(instrument_size, sizes[2])
        self.octave_linear = nn.Linear(sum(sizes), melody_size * 8)
        self.scale_degree_linear = nn.Linear(sum(sizes), melody_size * 7)
        self.channels_linear = nn.Linear(n_note_features, ch)
        self.linear = nn.Linear(melody_size + ch, melody_size)


class PitchedRhythmEncoder(_Part):
    def __init__(self, rhythm_size, beat_size, bar_size, instrument_size):
        super().__init__()
        self._check(rhythm_size=rhythm_size, beat_size=beat_size, bar_size=bar_size, instrument_size=instrument_size)
        sizes = [get_mean_size(beat_size, rhythm_size), get_mean_size(bar_size, rhythm_size, factor=.5),
                 get_mean_size(280, rhythm_size, factor=.1), get_mean_size(instrument_size, rhythm_size, factor=.5),
                 get_mean_size(2, rhythm_size, factor=.25), get_mean_size(1, rhythm_size, factor=.25)]
        self.beats_linear = nn.Linear(beat_size, sizes[0])
        self.bars_linear = nn.Linear(bar_size, sizes[1])
        self.channels_linear = nn.Linear(280, sizes[2])
        self.instruments_linear = nn.Linear(instrument_size, sizes[3])
        self.mode_linear = nn.Linear(2, sizes[4])
        self.bpm_linear = nn.Linear(1, sizes[5])
        self.linear = nn.Linear(sum(sizes), rhythm_size)


class UnpitchedRhythmEncoder(_Part):
    def __init__(self, rhythm_size, beat_size, bar_size):
        super().__init__()
        self._check(rhythm_size=rhythm_size, beat_size=beat_size, bar_size=bar_size)
        sizes = [get_mean_size(beat_size, rhythm_size), get_mean_size(bar_size, rhythm_size, factor=.5),
                 get_mean_size(94, rhythm_size, factor=.25), get_mean_size(1, rhythm_size, factor=.25)]
        self.beats_linear = nn.Linear(beat_size, sizes[0])
        self.bars_linear = nn.Linear(bar_size, sizes[1])
        self.channels_linear = nn.Linear(94, sizes[2])
        self.bpm_linear = nn.Linear(1, sizes[3])
        self.linear = nn.Linear(sum(sizes), rhythm_size)


class SongInfoModel(_Part):
    def __init__(self, n_rhythm_features, style_size, rhythm_size, n_instruments):
        super().__init__()
        self._check(n_rhythm_features=n_rhythm_features, style_size=style_size, rhythm_size=rhythm_size,
                    n_instruments=n_instruments)
        bl = get_mean_size(10 * rhythm_size, n_rhythm_features, factor=.05)
        self.beats_lstm = Distributed(_lstm(10 * rhythm_size, bl), depth=1)
        self.bars_lstm = _lstm(bl, n_rhythm_features)
        for name, fs, fr, out in [('instruments', .05, .25, n_instruments), ('mode', .01, .1, 2), ('bpm', .01, .1, 1)]:
            a, b = get_mean_size(style_size, out, factor=fs), get_mean_size(rhythm_size, out, factor=fr)
            setattr(self, f'style_{name}_linear', nn.Linear(style_size, a))
            setattr(self, f'rhythm_{name}_linear', nn.Linear(n_rhythm_features, b))
            setattr(self, f'{name}_linear', nn.Linear(a + b, out))


class PitchedStyleApplier(_Part):
    def __init__(self, style_size, melody_size, rhythm_size, instrument_size):
        super().__init__()
        self._check(style_size=style_size, melody_size=melody_size, rhythm_size=rhythm_size, instrument_size=instrument_size)
        sizes = [get_mean_size(style_size, n_note_features, factor=.5), get_mean_size(rhythm_size, n_note_features, factor=.5),
                 get_mean_size(instrument_size, n_note_features, factor=.4)]
        ml = get_mean_size(melody_size, n_note_features, factor=3)
        self.style_linear = nn.Linear(style_size, sizes[0])
        self.rhythm_linear = nn.Linear(rhythm_size, sizes[1])
        self.instruments_linear = nn.Linear(instrument_size, sizes[2])
        self.octave_linear = nn.Linear(sum(sizes), 30 * 8)
        self.scale_degree_linear = nn.Linear(sum(sizes), 30 * 7)
        self.melody_linear = nn.Linear(melody_size, ml)
        self.linear = nn.Linear(30 + ml, n_note_features)


class UnpitchedStyleApplier(_Part):
    def __init__(self, style_size, rhythm_size):
        super().__init__()
        self._check(style_size=style_size, rhythm_size=rhythm_size)
        sl, rl = get_mean_size(style_size, 2, factor=.5), get_mean_size(rhythm_size, 2, factor=1)
        self.style_linear = nn.Linear(style_size, 10 * sl)
        self.rhythm_linear = nn.Linear(rhythm_size, rl)
        self.notes_linear = nn.Linear(sl + rl, 47 * 8)
        self.linear = nn.Linear(8, 2)


class _ForwardContext:
    """What one model(...) call leaves behind for get_total_loss / backward: the workspace and its inputs."""

    def __init__(self, model, ws, inputs, src_ptr=0):
        self.model, self.ws, self.inputs, self.src_ptr = model, ws, inputs, src_ptr


class _TotalLoss(torch.autograd.Function):
    """losses['total']: its backward runs mst_model_backward, which adds the parameter gradients into the flat buffer
    the parameters' .grad tensors are views of (autograd itself has nothing to propagate)."""

    @staticmethod
    def forward(ctx, anchor, handle):
        ctx.handle = handle
        return handle['host'][0, :1].clone()          # shape (1,) like the reference's total (bpm_loss keeps its batch dim)

    @staticmethod
    def backward(ctx, grad_out):
        h = ctx.handle
        fc = h['fc']
        model = fc.model
        model._attach_grads()
        seed = float(grad_out.reshape(-1)[0])
        grads = model._flat_grads if seed == 1.0 else torch.zeros_like(model._flat_grads)
        mode, bpm, pitched, feats, unpitched = fc.inputs
        _mm.backward_from_forward(fc.ws, model._flat, grads, pitched, unpitched, feats, mode, bpm, h['instruments_target'],
                                  h['bpm_target'], normalize=h['normalize'])
        if seed != 1.0:
            model._flat_grads.add_(grads, alpha=seed)
        return None, None


class StyleTransferModel(nn.Module):
    def __init__(
            self, pitched_channels_encoder, unpitched_channels_encoder,
            style_encoder, melody_encoder,
            pitched_rhythm_encoder, unpitched_rhythm_encoder,
            song_info_model,
            pitched_style_applier, unpitched_style_applier):
        super().__init__()
        self.pitched_channels_encoder = pitched_channels_encoder
        self.unpitched_channels_encoder = unpitched_channels_encoder
        self.style_encoder = style_encoder
        self.melody_encoder = melody_encoder
        self.pitched_rhythm_encoder = pitched_rhythm_encoder
        self.unpitched_rhythm_encoder = unpitched_rhythm_encoder
        self.song_info_model = song_info_model
        self.pitched_style_applier = pitched_style_applier
        self.unpitched_style_applier = unpitched_style_applier
        self._flat = self._flat_grads = None
        self._workspaces = {}
        self._anchor = None

    # ---- flat parameter buffer -------------------------------------------------------------------------------
    def _flatten(self):
        """Move every parameter into the library's flat layout (once; idempotent)."""
        if self._flat is not None and self._flat.is_cuda and self._params_in_flat():
            return
        if not torch.cuda.is_available():
            raise RuntimeError('StyleTransferModel needs a CUDA device: the B200 path has no CPU fallback')
        named = dict(self.named_parameters())
        table = _mm.param_table()
        assert set(named) == {n for n, _, _ in table}, 'parameter names differ from the library table'
        flat = torch.empty(_mm.total_params(), dtype=torch.float32, device='cuda')
        for name, off, numel in table:
            p = named[name]
            flat[off:off + numel] = p.detach().reshape(-1).to('cuda', torch.float32)
            p.data = flat[off:off + numel].view(p.shape)
        self._flat = flat
        self._flat_grads = torch.zeros_like(flat)
        self._anchor = torch.zeros(1, device='cuda', requires_grad=True)

    def _params_in_flat(self):
        """True while every parameter still aliases the flat buffer (a Module._apply such as .cpu() / .cuda() / .half()
        reallocates them: the kernels must not keep reading the stale buffer while the optimiser updates new tensors)."""
        lo = self._flat.data_ptr()
        hi = lo + self._flat.numel() * 4
        return all(p.is_cuda and p.dtype == torch.float32 and lo <= p.data_ptr() < hi for p in self.parameters())

    def _attach_grads(self):
        """Parameters' .grad = views of the flat gradient buffer.  optimizer.zero_grad() (set_to_none) drops them:
        re-attaching after that starts from a zeroed buffer, which is what zero_grad means."""
        named = dict(self.named_parameters())
        if all(p.grad is None for p in named.values()):
            self._flat_grads.zero_()
        for name, off, numel in _mm.param_table():
            p = named[name]
            view = self._flat_grads[off:off + numel].view(p.shape)
            if p.grad is None or p.grad.data_ptr() != view.data_ptr():
                if p.grad is not None:
                    view.copy_(p.grad)
                p.grad = view

    def __getstate__(self):
        state = self.__dict__.copy()
        state['_workspaces'] = {}
        return state

    def __setstate__(self, state):
        # a torch.save(model) pickle of the REFERENCE carries none of the private attributes: default them, the first
        # forward then moves the parameters into the flat buffer
        self.__dict__.update(state)
        for k, v in (('_flat', None), ('_flat_grads', None), ('_workspaces', {}), ('_anchor', None)):
            self.__dict__.setdefault(k, v)

    def _workspace(self, C, B, nb, drums, training):
        key = (C, B, nb, bool(drums), bool(training))
        if key in self._workspaces:
            self._workspaces[key] = self._workspaces.pop(key)       # most recently used last
        else:
            # songs differ in length, but train-model.py:101 crops to 800 // C bars, so a few shapes recur: keep the last few
            # (their activation arenas and captured graphs) instead of one
            keep = int(os.environ.get('MST_WORKSPACE_CACHE', 4))
            while len(self._workspaces) >= max(1, keep):
                self._workspaces.pop(next(iter(self._workspaces)))
            # one CUDA graph per entry point and shape (MST_NO_GRAPH=1 launches kernel by kernel); MST_PRECISION picks the mode
            self._workspaces[key] = _mm.Workspace(1, C, B, nb, drums, training=training, graph=not os.environ.get('MST_NO_GRAPH'))
        return self._workspaces[key]

    @staticmethod
    def _prep(mode, bpm, pitched, feats, unpitched):
        f = lambda t: None if t is None else t.detach().to('cuda', torch.float32).contiguous()
        return f(mode), f(bpm).reshape(-1), f(pitched), f(feats), f(unpitched)

    # ---- the four entry points of the reference (model.py:751-793) ---------------------------------------------
    def _run_forward(self, mode, bpm, pitched_channels, instruments_features, unpitched_channels, training):
        self._flatten()
        inputs = self._prep(mode, bpm, pitched_channels, instruments_features, unpitched_channels)
        mode, bpm, pitched, feats, unpitched = inputs
        _, C, B, nb = pitched.shape[:4]
        ws = self._workspace(C, B, nb, unpitched is not None, training)
        out = ws.forward(self._flat, pitched, unpitched, feats, mode, bpm)
        return out, _ForwardContext(self, ws, ws.last_inputs, pitched.data_ptr())

    def extract_style(self, mode, bpm, pitched_channels, instruments_features, unpitched_channels=None):
        out, _ = self._run_forward(mode, bpm, pitched_channels, instruments_features, unpitched_channels, False)
        return out['style'].clone(), out['melody'].clone(), out['rhythm'].clone()

    def _predict_apply(self, style, melody, rhythm, instruments_features, unpitched):
        self._flatten()
        f = lambda t: t.detach().to('cuda', torch.float32).contiguous()
        style, melody, rhythm, feats = f(style), f(melody), f(rhythm), f(instruments_features)
        _, B, nb = rhythm.shape[:3]
        ws = self._workspace(feats.shape[1], B, nb, unpitched, False)
        return ws.predict_apply(self._flat, style, melody, rhythm, feats)

    def predict_song_info(self, style, rhythm):
        self._flatten()
        _, B, nb = rhythm.shape[:3]
        melody = torch.zeros(1, B, nb, 10, 56, 8, device='cuda')
        feats = torch.zeros(1, 1, 51, device='cuda')
        out = self._predict_apply(style, melody, rhythm, feats, False)
        return out['instruments'].clone(), out['mode'].clone(), out['bpm'].clone()

    def apply_style(self, style, melody, rhythm, instruments_features, unpitched=False):
        out = self._predict_apply(style, melody, rhythm, instruments_features, bool(unpitched))
        return out['pitched'].clone(), (out['unpitched'].clone() if unpitched else None)

    def forward(self, mode, bpm, pitched_channels, instruments_features, unpitched_channels=None):
        out, fc = self._run_forward(mode, bpm, pitched_channels, instruments_features, unpitched_channels,
                                    torch.is_grad_enabled())
        song_info = (out['instruments'], out['mode'], out['bpm'])
        x_pitched, x_unpitched = out['pitched'], out['unpitched']
        for t in song_info + (x_pitched,) + ((x_unpitched,) if x_unpitched is not None else ()):
            t._mst_forward = fc                           # lets get_total_loss find the workspace
        return song_info, x_pitched, x_unpitched


def combine(x, dim):
    """combine (model.py:796-815) for callers that use it on their own tensors (plain torch; the network's own
    combines are fused kernels)."""
    dims = [i for i in range(x.dim()) if i != dim]
    norm = torch.sqrt(1.0 + (x ** 2).sum(dims, keepdim=True))
    return (x * norm).sum(dim) / norm.sum()


def hard_output(x):
    """hard_output (model.py:818-832): in-place velocity gate (> .01) and accidental arg-max gate (> .1).
    decode_midi applies the same gates inside the decode kernel; this is the stand-alone tensor version."""
    velocity = x[..., 1]
    velocity *= (velocity > .01).float()
    if x.shape[-1] > 2:
        acc = x[..., 2:]
        acc_max = acc.max(-1, keepdim=True)[0]
        x = torch.cat([x[..., :2], ((acc == acc_max) & (acc > .1)).float()], -1)
    return x


_LEAVES = _mm.LOSS_LEAVES


def get_total_loss(instruments_pred, instruments_target, mode_pred, mode_target, bpm_pred, bpm_target,
                   pitched_pred, pitched_target, unpitched_pred=None, unpitched_target=None, normalize=False):
    """get_total_loss (model.py:935-997), positional quirk included: the caller passes the bpm pair in the mode_* slots
    and the mode pair in the bpm_* slots (train-model.py:115-122), get_song_info_loss swaps them back (:900-901)."""
    bpm_in, bpm_t, mode_in, mode_t = mode_pred, mode_target, bpm_pred, bpm_target          # undo the swap
    dev = pitched_pred.device
    N = pitched_pred.shape[0]
    bpm_t = torch.as_tensor(bpm_t, dtype=torch.float32, device=dev).reshape(-1).expand(N).contiguous()
    use_u = unpitched_target is not None
    leaves = _mm.total_loss(pitched_pred.detach(), pitched_target, unpitched_pred.detach() if use_u else None,
                            unpitched_target if use_u else None, instruments_pred.detach(), instruments_target,
                            mode_in.detach(), mode_t, bpm_in.detach().reshape(-1), bpm_t, normalize=normalize)
    fc = getattr(pitched_pred, '_mst_forward', None)
    # ONE device -> host copy of the (N, 16) leaf buffer: every leaf of the returned dict is a host scalar, so the
    # reference's `float(x)` per leaf (train-model.py:129, ~15 device syncs there) costs nothing more
    host = leaves.cpu()
    row = host[0]
    get = lambda k: row[_LEAVES[k]]
    total = row[:1]
    if fc is not None and fc.ws.training and torch.is_grad_enabled():
        same = fc.src_ptr == pitched_target.data_ptr() or torch.equal(fc.inputs[2], pitched_target)
        if same and (fc.inputs[4] is not None) == use_u:
            handle = dict(fc=fc, leaves=leaves, host=host, instruments_target=instruments_target.detach().float().contiguous(),
                          bpm_target=bpm_t, normalize=bool(normalize))
            total = _TotalLoss.apply(fc.model._anchor, handle)
    out_p = dict(total=get('channels_loss/pitched/total'), notes_loss=get('channels_loss/pitched/notes_loss'),
                 velocity_loss=get('channels_loss/pitched/velocity_loss'),
                 duration_loss=get('channels_loss/pitched/duration_loss'),
                 accidentals_loss=get('channels_loss/pitched/accidentals_loss'))
    out_u = None
    if use_u:
        out_u = dict(total=get('channels_loss/unpitched/total'), notes_loss=get('channels_loss/unpitched/notes_loss'),
                     velocity_loss=get('channels_loss/unpitched/velocity_loss'),
                     duration_loss=get('channels_loss/unpitched/duration_loss'))
    return {
        'total': total,
        'channels_loss': {'total': get('channels_loss/total'), 'pitched': out_p, 'unpitched': out_u},
        'song_info_loss': {'total': get('song_info_loss/total'), 'instruments_loss': get('song_info_loss/instruments_loss'),
                           'mode_loss': get('song_info_loss/mode_loss'), 'bpm_loss': get('song_info_loss/bpm_loss')},
    }
