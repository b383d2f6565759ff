"""ctypes binding of the model entry points of libmst_b200.so (mst_model_*, include/mst_b200.h).

Parameters live in one flat fp32 CUDA buffer whose layout the library describes by state-dict name
(mst_model_param_*), so a reference `state_dict()` maps onto it by name."""
import ctypes

import numpy as np

from . import lib as _lib


class ModelShape(ctypes.Structure):
    _fields_ = [('n_samples', ctypes.c_int32), ('n_channels', ctypes.c_int32), ('n_bars', ctypes.c_int32),
                ('n_beats', ctypes.c_int32), ('has_unpitched', ctypes.c_int32), ('precision', ctypes.c_int32)]


def precision_id(precision):
    """'fp32' / 'bf16' / None (= the MST_PRECISION environment variable, default fp32) -> MST_PRECISION_*."""
    import os
    if precision is None:
        precision = os.environ.get('MST_PRECISION', 'fp32')
    if precision in (0, 1):
        return int(precision)
    try:
        return {'fp32': 0, 'f32': 0, 'bf16': 1}[str(precision).lower()]
    except KeyError:
        raise ValueError(f'unknown precision {precision!r}: fp32 or bf16')


class ModelOutputs(ctypes.Structure):
    _fields_ = [(k, ctypes.c_void_p) for k in ('instruments', 'mode', 'bpm', 'pitched', 'unpitched', 'style',
                                               'melody', 'rhythm')]


class LossIn(ctypes.Structure):
    _fields_ = [(k, ctypes.c_void_p) for k in (
        'pitched_pred', 'pitched_target', 'unpitched_pred', 'unpitched_target', 'instruments_logits',
        'instruments_target', 'mode_logits', 'mode_target', 'bpm_pred', 'bpm_target')] + [
        ('normalize', ctypes.c_int32), ('reserved', ctypes.c_int32)]


EXPORTS = ['mst_model_param_count', 'mst_model_param_name', 'mst_model_param_numel', 'mst_model_param_offset',
           'mst_model_workspace_bytes', 'mst_model_forward', 'mst_model_train_step', 'mst_model_predict_apply',
           'mst_model_song_info', 'mst_model_backward',
           'mst_model_loss_workspace_bytes',
           'mst_model_loss', 'mst_model_loss_noalloc', 'mst_adam_step', 'mst_model_gemm']

N_LOSS_LEAVES = 16
# index of every leaf of get_total_loss's nested dict in the (N, 16) loss buffer (include/mst_b200.h MST_LOSS_*)
LOSS_LEAVES = {
    'total': 0, 'channels_loss/total': 1, 'channels_loss/pitched/total': 2, 'channels_loss/pitched/notes_loss': 3,
    'channels_loss/pitched/velocity_loss': 4, 'channels_loss/pitched/duration_loss': 5,
    'channels_loss/pitched/accidentals_loss': 6, 'channels_loss/unpitched/total': 7,
    'channels_loss/unpitched/notes_loss': 8, 'channels_loss/unpitched/velocity_loss': 9,
    'channels_loss/unpitched/duration_loss': 10, 'song_info_loss/total': 11, 'song_info_loss/instruments_loss': 12,
    'song_info_loss/mode_loss': 13, 'song_info_loss/bpm_loss': 14}

_ready = False
_loss_bufs = {}


def _L():
    global _ready
    L = _lib.load()
    if not _ready:
        L.mst_model_param_count.restype = ctypes.c_int
        L.mst_model_param_name.restype = ctypes.c_char_p
        L.mst_model_param_name.argtypes = [ctypes.c_int]
        L.mst_model_param_numel.restype = ctypes.c_longlong
        L.mst_model_param_numel.argtypes = [ctypes.c_int]
        L.mst_model_param_offset.restype = ctypes.c_longlong
        L.mst_model_param_offset.argtypes = [ctypes.c_int]
        L.mst_model_workspace_bytes.restype = ctypes.c_size_t
        L.mst_model_workspace_bytes.argtypes = [ctypes.POINTER(ModelShape), ctypes.c_int]
        L.mst_model_forward.restype = ctypes.c_int
        L.mst_model_forward.argtypes = [ctypes.c_void_p, ctypes.POINTER(ModelShape)] + [ctypes.c_void_p] * 5 + [
            ctypes.POINTER(ModelOutputs), ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p,
            ctypes.POINTER(ctypes.c_int)]
        L.mst_model_train_step.restype = ctypes.c_int
        L.mst_model_train_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ModelShape)] + \
            [ctypes.c_void_p] * 7 + [ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ModelOutputs), ctypes.c_void_p,
                                     ctypes.c_size_t, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
        L.mst_model_predict_apply.restype = ctypes.c_int
        L.mst_model_predict_apply.argtypes = [ctypes.c_void_p, ctypes.POINTER(ModelShape)] + [ctypes.c_void_p] * 4 + [
            ctypes.POINTER(ModelOutputs), ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
        L.mst_model_song_info.restype = ctypes.c_int
        L.mst_model_song_info.argtypes = [ctypes.c_void_p, ctypes.POINTER(ModelShape)] + [ctypes.c_void_p] * 2 + [
            ctypes.POINTER(ModelOutputs), ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
        L.mst_model_backward.restype = ctypes.c_int
        L.mst_model_backward.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ModelShape)] + \
            [ctypes.c_void_p] * 7 + [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p,
                                     ctypes.POINTER(ctypes.c_int)]
        L.mst_model_loss_workspace_bytes.restype = ctypes.c_size_t
        L.mst_model_loss_workspace_bytes.argtypes = [ctypes.POINTER(ModelShape)]
        L.mst_model_loss.restype = ctypes.c_int
        L.mst_model_loss.argtypes = [ctypes.POINTER(ModelShape), ctypes.POINTER(LossIn), ctypes.c_void_p, ctypes.c_void_p,
                                     ctypes.c_size_t, ctypes.c_void_p]
        L.mst_model_loss_noalloc.restype = ctypes.c_int
        L.mst_model_loss_noalloc.argtypes = [ctypes.POINTER(ModelShape), ctypes.POINTER(LossIn), ctypes.c_void_p, ctypes.c_void_p]
        L.mst_adam_step.restype = ctypes.c_int
        L.mst_adam_step.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_longlong, ctypes.c_int] + [ctypes.c_float] * 5 + [
            ctypes.c_int, ctypes.c_void_p]
        L.mst_model_gemm.restype = ctypes.c_int
        L.mst_model_gemm.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong,
                                     ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_int,
                                     ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                     ctypes.c_void_p]
        _ready = True
    return L


def param_table():
    """[(name, offset, numel)] of the flat parameter buffer."""
    L = _L()
    n = L.mst_model_param_count()
    return [(L.mst_model_param_name(i).decode(), int(L.mst_model_param_offset(i)), int(L.mst_model_param_numel(i)))
            for i in range(n)]


def total_params():
    L = _L()
    return int(L.mst_model_param_offset(L.mst_model_param_count()))


def flatten_state_dict(sd, device='cuda'):
    """state_dict (reference names) -> flat fp32 tensor in the library's layout."""
    import torch
    flat = torch.empty(total_params(), dtype=torch.float32, device=device)
    for name, off, numel in param_table():
        t = sd[name]
        assert t.numel() == numel, (name, tuple(t.shape), numel)
        flat[off:off + numel] = t.detach().reshape(-1).to(device=device, dtype=torch.float32)
    return flat


def unflatten(flat, like_sd):
    """flat buffer -> {name: tensor view shaped like like_sd[name]}."""
    return {name: flat[off:off + numel].view(like_sd[name].shape) for name, off, numel in param_table()}


class Workspace:
    """Activation arena for one batch shape (reused across steps)."""

    def __init__(self, N, C, B, nb, has_unpitched, training=False, device=None, precision=None, graph=False):
        import torch
        _lib.require_gpu()
        self.shape = ModelShape(int(N), int(C), int(B), int(nb), int(bool(has_unpitched)), precision_id(precision))
        self.training = bool(training)
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        self.bytes = _L().mst_model_workspace_bytes(ctypes.byref(self.shape), int(self.training))
        self.buf = torch.empty(self.bytes, dtype=torch.uint8, device=self.device)
        self.launches = 0
        # CUDA graphs (one per entry point, captured on first use when graph=True): a step is ~200-425 short launches, at
        # N = 1 (the reference's own batch, train-model.py:97-154) half of the step time is launch overhead
        self.graph = bool(graph)
        self._graphs = {}          # entry name -> (torch.cuda.CUDAGraph, result, n_launches)
        self._uses = {}
        self._static = {}          # fixed-address copies of the inputs the captured kernels read

    def _stage(self, name, t):
        """Copy `t` into this workspace's fixed-address input buffer `name` (graphs bake addresses in)."""
        import torch
        if t is None:
            return None
        buf = self._static.get(name)
        if buf is None or buf.shape != t.shape:
            buf = self._static[name] = torch.empty_like(t, dtype=torch.float32, device=self.device).contiguous()
        if buf.data_ptr() != t.data_ptr():
            buf.copy_(t, non_blocking=True)
        return buf

    def _run(self, name, fn):
        """Run `fn(stream)` eagerly the first time (its launches execute), capture it once, replay afterwards."""
        import torch
        if not self.graph:
            return fn(torch.cuda.current_stream(self.device))
        hit = self._graphs.get(name)
        if hit is None:
            uses = self._uses[name] = self._uses.get(name, 0) + 1
            if uses < 2:                                           # a shape seen once is not worth a capture
                return fn(torch.cuda.current_stream(self.device))
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                result = fn(torch.cuda.current_stream(self.device))    # recorded, not executed
            hit = self._graphs[name] = (g, result, self.launches)
        g, result, self.launches = hit
        g.replay()
        return result

    def _view(self, ptr, shape):
        import torch
        if not ptr:
            return None
        off = ptr - self.buf.data_ptr()
        n = int(np.prod(shape))
        return self.buf[off:off + 4 * n].view(torch.float32).view(*shape)

    def forward(self, flat_params, pitched, unpitched, feats, mode, bpm, stream=None):
        """Inputs are contiguous float32 CUDA tensors with a leading sample dim N.  Returns a dict of views into
        the workspace (valid until the next call)."""
        import torch
        s = self.shape
        N, C, B, nb = s.n_samples, s.n_channels, s.n_bars, s.n_beats
        assert tuple(pitched.shape) == (N, C, B, nb, 10, 56, 5) and pitched.is_contiguous()
        assert tuple(feats.shape) == (N, C, 51) and tuple(mode.shape) == (N, 2) and tuple(bpm.shape) == (N,)
        if s.has_unpitched:
            assert unpitched is not None and unpitched.numel() == N * B * nb * 10 * 47 * 2 and unpitched.is_contiguous()
        feats, mode, bpm = feats.contiguous(), mode.contiguous(), bpm.contiguous()
        if self.graph and stream is None:
            pitched, unpitched = self._stage('pitched', pitched), self._stage('unpitched', unpitched if s.has_unpitched else None)
            feats, mode, bpm = self._stage('feats', feats), self._stage('mode', mode), self._stage('bpm', bpm)
            self._static['params_ptr'] = flat_params.data_ptr()

        self.last_inputs = (mode, bpm, pitched, feats, unpitched)      # what the kernels read (staged copies under a graph)

        def call(st):
            outs = ModelOutputs()
            nl = ctypes.c_int(0)
            _lib.check(_L().mst_model_forward(
                flat_params.data_ptr(), ctypes.byref(s), pitched.data_ptr(),
                unpitched.data_ptr() if s.has_unpitched else None, feats.data_ptr(), mode.data_ptr(), bpm.data_ptr(),
                ctypes.byref(outs), self.buf.data_ptr(), self.bytes, int(self.training), st.cuda_stream, ctypes.byref(nl)))
            self.launches = nl.value
            return self._outs(outs)
        if stream is not None:
            return call(stream)
        return self._run(f'forward@{flat_params.data_ptr()}', call)

    def _outs(self, outs):
        s = self.shape
        N, C, B, nb = s.n_samples, s.n_channels, s.n_bars, s.n_beats
        return dict(
            instruments=self._view(outs.instruments, (N, 41)), mode=self._view(outs.mode, (N, 2)),
            bpm=self._view(outs.bpm, (N,)), pitched=self._view(outs.pitched, (N, C, B, nb, 10, 56, 5)),
            unpitched=self._view(outs.unpitched, (N, 1, B, nb, 10, 47, 2)) if s.has_unpitched else None,
            style=self._view(outs.style, (N, 256)), melody=self._view(outs.melody, (N, B, nb, 10, 56, 8)),
            rhythm=self._view(outs.rhythm, (N, B, nb, 10, 32)))

    def predict_apply(self, flat_params, style, melody, rhythm, feats, stream=None):
        """predict_song_info + apply_style from given style / melody / rhythm tensors (contiguous float32 CUDA)."""
        import torch
        s = self.shape
        N, C, B, nb = s.n_samples, s.n_channels, s.n_bars, s.n_beats
        assert tuple(style.shape) == (N, 256) and tuple(melody.shape) == (N, B, nb, 10, 56, 8)
        assert tuple(rhythm.shape) == (N, B, nb, 10, 32) and tuple(feats.shape) == (N, C, 51)
        outs = ModelOutputs()
        nl = ctypes.c_int(0)
        st = torch.cuda.current_stream(self.device) if stream is None else stream
        _lib.check(_L().mst_model_predict_apply(
            flat_params.data_ptr(), ctypes.byref(s), style.contiguous().data_ptr(), melody.contiguous().data_ptr(),
            rhythm.contiguous().data_ptr(), feats.contiguous().data_ptr(), ctypes.byref(outs), self.buf.data_ptr(), self.bytes,
            st.cuda_stream, ctypes.byref(nl)))
        self.launches = nl.value
        return dict(
            instruments=self._view(outs.instruments, (N, 41)), mode=self._view(outs.mode, (N, 2)),
            bpm=self._view(outs.bpm, (N,)), pitched=self._view(outs.pitched, (N, C, B, nb, 10, 56, 5)),
            unpitched=self._view(outs.unpitched, (N, 1, B, nb, 10, 47, 2)) if s.has_unpitched else None)

    def song_info(self, flat_params, style, rhythm, stream=None):
        """predict_song_info from given style (N,256) / rhythm (N,B,nb,10,32): dict of workspace views instruments (N,41),
        mode (N,2), bpm (N)."""
        import torch
        s = self.shape
        N, B, nb = s.n_samples, s.n_bars, s.n_beats
        assert tuple(style.shape) == (N, 256) and tuple(rhythm.shape) == (N, B, nb, 10, 32)
        outs = ModelOutputs()
        nl = ctypes.c_int(0)
        st = torch.cuda.current_stream(self.device) if stream is None else stream
        _lib.check(_L().mst_model_song_info(flat_params.data_ptr(), ctypes.byref(s), style.contiguous().data_ptr(),
                                            rhythm.contiguous().data_ptr(), ctypes.byref(outs), self.buf.data_ptr(), self.bytes,
                                            st.cuda_stream, ctypes.byref(nl)))
        self.launches = nl.value
        return dict(instruments=self._view(outs.instruments, (N, 41)), mode=self._view(outs.mode, (N, 2)),
                    bpm=self._view(outs.bpm, (N,)))

    def train_step(self, flat_params, flat_grads, pitched, unpitched, feats, mode, bpm, instruments_target,
                   bpm_target, losses=None, normalize=True, stream=None):
        """forward + get_total_loss + backward for N independent samples; gradients are ADDED into flat_grads.
        Returns (losses (N, 16) device tensor, outputs dict of workspace views)."""
        import torch
        assert self.training, 'Workspace(training=True) is required'
        s = self.shape
        N, C, B, nb = s.n_samples, s.n_channels, s.n_bars, s.n_beats
        assert tuple(pitched.shape) == (N, C, B, nb, 10, 56, 5) and pitched.is_contiguous()
        assert tuple(feats.shape) == (N, C, 51) and tuple(mode.shape) == (N, 2) and tuple(bpm.shape) == (N,)
        assert tuple(instruments_target.shape) == (N, 41) and tuple(bpm_target.shape) == (N,)
        if s.has_unpitched:
            assert unpitched is not None and unpitched.numel() == N * B * nb * 10 * 47 * 2 and unpitched.is_contiguous()
        if losses is None:
            losses = torch.empty(N, N_LOSS_LEAVES, dtype=torch.float32, device=self.device)
        feats, mode, bpm = feats.contiguous(), mode.contiguous(), bpm.contiguous()
        it, bt = instruments_target.contiguous(), bpm_target.contiguous()
        use_graph = self.graph and stream is None
        out_losses = losses
        if use_graph:
            pitched, unpitched = self._stage('pitched', pitched), self._stage('unpitched', unpitched if s.has_unpitched else None)
            feats, mode, bpm = self._stage('feats', feats), self._stage('mode', mode), self._stage('bpm', bpm)
            it, bt = self._stage('it', it), self._stage('bt', bt)
            losses = self._static.setdefault('losses', torch.empty(N, N_LOSS_LEAVES, dtype=torch.float32, device=self.device))

        def call(st):
            outs = ModelOutputs()
            nl = ctypes.c_int(0)
            _lib.check(_L().mst_model_train_step(
                flat_params.data_ptr(), flat_grads.data_ptr(), ctypes.byref(s), pitched.data_ptr(),
                unpitched.data_ptr() if s.has_unpitched else None, feats.data_ptr(), mode.data_ptr(), bpm.data_ptr(),
                it.data_ptr(), bt.data_ptr(), int(bool(normalize)), losses.data_ptr(), ctypes.byref(outs),
                self.buf.data_ptr(), self.bytes, st.cuda_stream, ctypes.byref(nl)))
            self.launches = nl.value
            return self._outs(outs)
        if not use_graph:
            return losses, call(torch.cuda.current_stream(self.device) if stream is None else stream)
        outs = self._run(f'train_step@{flat_params.data_ptr()},{flat_grads.data_ptr()},{int(bool(normalize))}', call)
        if out_losses.data_ptr() != losses.data_ptr():
            out_losses.copy_(losses, non_blocking=True)
        return out_losses, outs


def backward_from_forward(ws, flat_params, flat_grads, pitched, unpitched, feats, mode, bpm, instruments_target, bpm_target,
                          normalize=True, losses=None, stream=None):
    """get_total_loss + backward on the workspace `ws` that Workspace.forward (training=True) just filled."""
    import torch
    assert ws.training
    s = ws.shape
    if losses is None:
        losses = torch.empty(s.n_samples, N_LOSS_LEAVES, dtype=torch.float32, device=ws.device)
    it, bt = instruments_target.contiguous(), bpm_target.contiguous()
    use_graph = ws.graph and stream is None
    out_losses = losses
    if use_graph:
        # the forward of this step staged its inputs at fixed addresses; targets and the leaf buffer get theirs here
        pitched, unpitched = ws._stage('pitched', pitched), ws._stage('unpitched', unpitched if s.has_unpitched else None)
        feats, mode, bpm = ws._stage('feats', feats), ws._stage('mode', mode), ws._stage('bpm', bpm)
        it, bt = ws._stage('it', it), ws._stage('bt', bt)
        losses = ws._static.setdefault('bwd_losses', torch.empty(s.n_samples, N_LOSS_LEAVES, dtype=torch.float32, device=ws.device))

    def call(st):
        nl = ctypes.c_int(0)
        _lib.check(_L().mst_model_backward(
            flat_params.data_ptr(), flat_grads.data_ptr(), ctypes.byref(s), pitched.data_ptr(),
            unpitched.data_ptr() if s.has_unpitched else None, feats.data_ptr(), mode.data_ptr(), bpm.data_ptr(),
            it.data_ptr(), bt.data_ptr(), int(bool(normalize)), losses.data_ptr(), ws.buf.data_ptr(),
            ws.bytes, st.cuda_stream, ctypes.byref(nl)))
        ws.launches = nl.value
        return None
    if not use_graph:
        call(torch.cuda.current_stream(ws.device) if stream is None else stream)
        return losses
    ws._run(f'backward@{flat_params.data_ptr()},{flat_grads.data_ptr()},{int(bool(normalize))}', call)
    if out_losses.data_ptr() != losses.data_ptr():
        out_losses.copy_(losses, non_blocking=True)
    return out_losses


def total_loss(pitched_pred, pitched_target, unpitched_pred, unpitched_target, instruments_logits, instruments_target,
               mode_logits, mode_target, bpm_pred, bpm_target, normalize=False):
    """get_total_loss on device tensors with a leading sample dim -> (N, 16) loss leaves."""
    import torch
    _lib.require_gpu()
    N, C, B, nb = pitched_pred.shape[:4]
    shape = ModelShape(int(N), int(C), int(B), int(nb), int(unpitched_pred is not None), 0)
    f = lambda t: None if t is None else t.contiguous().float()
    keep = [f(t) for t in (pitched_pred, pitched_target, unpitched_pred, unpitched_target, instruments_logits,
                           instruments_target, mode_logits, mode_target, bpm_pred, bpm_target)]
    li = LossIn(*[None if t is None else t.data_ptr() for t in keep], int(bool(normalize)), 0)
    wsb = _L().mst_model_loss_workspace_bytes(ctypes.byref(shape))
    key = (int(N), str(pitched_pred.device))
    if key not in _loss_bufs:
        _loss_bufs[key] = torch.empty(wsb, dtype=torch.uint8, device=pitched_pred.device)
    ws = _loss_bufs[key]
    out = torch.empty(int(N), N_LOSS_LEAVES, dtype=torch.float32, device=pitched_pred.device)
    _lib.check(_L().mst_model_loss(ctypes.byref(shape), ctypes.byref(li), out.data_ptr(), ws.data_ptr(), wsb,
                                   torch.cuda.current_stream(pitched_pred.device).cuda_stream))
    return out


LOSS_NOALLOC_MAX_SAMPLES = 256          # MST_LOSS_NOALLOC_MAX_SAMPLES


def total_loss_noalloc(out, pitched_pred, pitched_target, unpitched_pred, unpitched_target, instruments_logits,
                       instruments_target, mode_logits, mode_target, bpm_pred, bpm_target, normalize=False, stream=None):
    """`total_loss` through mst_model_loss_noalloc: nothing is allocated on the device, by the library or by this wrapper (the
    kernels' scratch is a static device array inside the library, at most LOSS_NOALLOC_MAX_SAMPLES samples).  Every tensor,
    including the (N, 16) float32 output `out`, is the caller's and must already be contiguous float32 on one device; calls
    share the scratch, so issue them on one stream (or order them).  No counterpart in the reference."""
    import torch
    _lib.require_gpu()
    tensors = (pitched_pred, pitched_target, unpitched_pred, unpitched_target, instruments_logits, instruments_target,
               mode_logits, mode_target, bpm_pred, bpm_target)
    for t in tensors + (out,):
        if t is not None and not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise ValueError('total_loss_noalloc takes contiguous float32 CUDA tensors (it does not copy or cast)')
    N, C, B, nb = pitched_pred.shape[:4]
    if tuple(out.shape) != (int(N), N_LOSS_LEAVES):
        raise ValueError(f'out must be ({int(N)}, {N_LOSS_LEAVES})')
    shape = ModelShape(int(N), int(C), int(B), int(nb), int(unpitched_pred is not None), 0)
    li = LossIn(*[None if t is None else t.data_ptr() for t in tensors], int(bool(normalize)), 0)
    st = torch.cuda.current_stream(out.device) if stream is None else stream
    _lib.check(_L().mst_model_loss_noalloc(ctypes.byref(shape), ctypes.byref(li), out.data_ptr(), st.cuda_stream))
    return out


def adam_step(params, grads, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, grad_scale=1.0, zero_grad=True, stream=None):
    """torch.optim.Adam.step on the flat buffers (in place)."""
    import torch
    _lib.require_gpu()
    st = torch.cuda.current_stream(params.device) if stream is None else stream
    _lib.check(_L().mst_adam_step(params.data_ptr(), grads.data_ptr(), m.data_ptr(), v.data_ptr(), params.numel(), int(step),
                                  float(lr), float(beta1), float(beta2), float(eps), float(grad_scale), int(bool(zero_grad)),
                                  st.cuda_stream))


def gemm(A, a_si, a_sk, B, b_sk, b_sj, C, c_si, M, N, K, bias=None, act=0, accumulate=0, split_k=1, kernel=0, stream=None):
    """mst_model_gemm on float32 CUDA tensors with explicit element strides (see include/mst_b200.h)."""
    import torch
    _lib.require_gpu()
    st = torch.cuda.current_stream(C.device) if stream is None else stream
    _lib.check(_L().mst_model_gemm(A.data_ptr(), a_si, a_sk, B.data_ptr(), b_sk, b_sj, C.data_ptr(), c_si, M, N, K,
                                   None if bias is None else bias.data_ptr(), act, accumulate, split_k, kernel,
                                   st.cuda_stream))
    return C
