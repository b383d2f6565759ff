"""Data-parallel training loop of the style-transfer network over the C ABI (train-model.py:97-154).

One process per GPU.  Every rank holds an identical replica of the flat parameter / Adam-state buffers and runs
mst_model_train_step on its own shard of the global batch (N independent batch-1 samples per rank); the flat
980 325-float gradient buffer is summed over ranks with ONE all-reduce (NCCL over NVLink on GPUs; gloo in the CPU
tests of the host logic) and every rank applies the same fused Adam + StepLR update (mst_adam_step), so the
replicas stay bit-identical without a parameter broadcast.  Sum (not mean) reproduces the reference's
iter_size gradient accumulation: a global batch of R*N samples == iter_size = R*N.
"""

def step_lr(base_lr, optimizer_steps_done, step_size=200, gamma=0.9):
    """torch.optim.lr_scheduler.StepLR(step_size, gamma) as train-model.py:90 uses it: lr of the NEXT optimizer
    step after `optimizer_steps_done` scheduler steps."""
    return base_lr * gamma ** (optimizer_steps_done // step_size)


def shard_samples(n_samples, world, rank):
    """Contiguous, balanced shard [lo, hi) of a global batch (first n_samples % world ranks get one more)."""
    base, extra = divmod(n_samples, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def balanced_assignment(costs, world):
    """Greedy longest-processing-time assignment of ragged samples (cost = C*B channel-bars) to ranks.
    Returns a list of index lists, one per rank; deterministic on every rank."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    loads = [0] * world
    out = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (loads[k], k))
        out[r].append(i)
        loads[r] += costs[i]
    return out


def allreduce_sum_(flat, group=None):
    """In-place sum of the flat gradient buffer over the process group (no-op for a single process)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


class Trainer:
    """Flat-buffer trainer: params / grads / Adam m, v as single fp32 CUDA tensors in the library's layout."""

    def __init__(self, state_dict, lr=0.01, step_size=200, gamma=0.9, betas=(0.9, 0.999), eps=1e-8, device=None, precision=None, graph=False):
        import torch
        from . import model as mm
        self.mm = mm
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        self.params = mm.flatten_state_dict(state_dict, device=self.device)
        self.grads = torch.zeros_like(self.params)
        self.m = torch.zeros_like(self.params)
        self.v = torch.zeros_like(self.params)
        self.base_lr, self.step_size, self.gamma, self.betas, self.eps = lr, step_size, gamma, betas, eps
        self.precision = mm.precision_id(precision)     # 'fp32' | 'bf16' (BASELINE configs[3]: bf16 dense layers)
        self.graph = bool(graph)                        # replay one CUDA graph per batch shape instead of ~425 launches
        self.optimizer_steps = 0
        self._ws = {}
        self.launches = 0

    def workspace(self, N, C, B, nb, drums):
        key = (N, C, B, nb, bool(drums))
        if key not in self._ws:
            self._ws[key] = self.mm.Workspace(N, C, B, nb, drums, training=True, device=self.device, precision=self.precision,
                                              graph=self.graph)
        return self._ws[key]

    def accumulate(self, pitched, unpitched, feats, mode, bpm, instruments_target, bpm_target, losses=None):
        """forward + loss + backward of N samples; gradients are added to self.grads.  Returns (N, 16) loss leaves."""
        N, C, B, nb = pitched.shape[:4]
        ws = self.workspace(N, C, B, nb, unpitched is not None)
        losses, _ = ws.train_step(self.params, self.grads, pitched, unpitched, feats, mode, bpm, instruments_target,
                                  bpm_target, losses=losses, normalize=True)
        self.launches = ws.launches
        return losses

    def optimizer_step(self, group=None):
        """all-reduce(sum) of the gradients over ranks, then Adam + zero_grad + StepLR on every rank."""
        allreduce_sum_(self.grads, group)
        lr = step_lr(self.base_lr, self.optimizer_steps, self.step_size, self.gamma)
        self.optimizer_steps += 1
        self.mm.adam_step(self.params, self.grads, self.m, self.v, self.optimizer_steps, lr, self.betas[0], self.betas[1],
                          self.eps, 1.0, True)
        self.launches += 1
        return lr

    def state_dict(self, like):
        return {k: v.clone() for k, v in self.mm.unflatten(self.params, like).items()}



def group_by_shape(shapes, indices=None):
    """{(C, B, nb, drums): [sample indices]} in first-appearance order: samples of one shape run as ONE batched step."""
    groups = {}
    for i in (range(len(shapes)) if indices is None else indices):
        groups.setdefault(tuple(shapes[i]), []).append(i)
    return groups


def ragged_plan(shapes, world, rank):
    """The part of a ragged global batch this rank trains: balanced_assignment by C*B channel-bars (deterministic on every
    rank, no communication), then grouped by shape.  shapes: [(C, B, nb, drums)] of the whole global batch."""
    mine = balanced_assignment([s[0] * s[1] for s in shapes], world)[rank]
    return group_by_shape(shapes, sorted(mine))


def _accumulate_ragged(self, samples, world=None, rank=None):
    """forward + loss + backward of this rank's share of a ragged list of samples (dicts with pitched (1,C,B,nb,10,56,5),
    unpitched or None, feats, mode, bpm, used, bpm_target): songs of different (C, B) as train-model.py:97-126 feeds them.
    Every rank passes the SAME list; the share is decided by ragged_plan.  Returns {sample index: (16,) loss leaves}."""
    import torch
    import torch.distributed as dist
    if world is None:
        on = dist.is_available() and dist.is_initialized()
        world, rank = (dist.get_world_size(), dist.get_rank()) if on else (1, 0)
    shapes = [(s['pitched'].shape[1], s['pitched'].shape[2], s['pitched'].shape[3], s['unpitched'] is not None) for s in samples]
    out = {}
    for shape, idx in ragged_plan(shapes, world, rank).items():
        cat = lambda k: torch.cat([samples[i][k] for i in idx]).to(self.device, torch.float32).contiguous()
        unp = cat('unpitched') if shape[3] else None
        bt = torch.tensor([float(samples[i]['bpm_target']) for i in idx], dtype=torch.float32, device=self.device)
        leaves = self.accumulate(cat('pitched'), unp, cat('feats'), cat('mode'), cat('bpm'), cat('used'), bt)
        for j, i in enumerate(idx):
            out[i] = leaves[j].clone()
    return out


Trainer.accumulate_ragged = _accumulate_ragged


def train_on_files(trainer, files, instruments, songs_per_step=2, min_n_messages=100, max_steps=None, crop_channel_bars=800,
                   encode_batch=16, on_step=None):
    """train-model.py:92-154 as a product loop on the CUDA path, data-parallel over the ranks of the default process group:
    files -> EventPipeline (C++ parse threads, encode with prepare_input's 800 // C crop on the device) -> ragged global
    batches of `songs_per_step` songs (iter_size) -> each rank trains the share ragged_plan gives it -> ONE gradient
    all-reduce + fused Adam per step.  Every rank encodes the same files (the encoder is ~1000x faster than a training
    step; it keeps the plan communication-free) and only trains its share.  Returns the number of optimizer steps."""
    import torch
    from style import data as sdata
    from . import abi, lib as _lib
    from .pipeline import EventPipeline
    pipe = EventPipeline(abi.make_filter(instruments, min_n_messages, crop_channel_bars=crop_channel_bars), batch_files=encode_batch)
    steps, batch = 0, []
    try:
        for path, r, arena in pipe.run(files):
            if int(r['status']) != abi.SONG_OK:
                continue
            pitched, unpitched = _lib.song_grids(r, arena)
            if float(pitched.sum()) == 0:                                         # train-model.py:105-106
                continue
            if unpitched is not None and float(unpitched.sum()) == 0:             # :107-109
                unpitched = None
            ins = [int(x) for x in r['pitched_instruments'][:int(r['n_pitched'])]]
            feats = torch.as_tensor(sdata.encode_instruments(ins), dtype=torch.float32, device=trainer.device).unsqueeze(0)
            batch.append(dict(
                pitched=pitched.unsqueeze(0).clone(), unpitched=None if unpitched is None else unpitched.unsqueeze(0).clone(), feats=feats,
                mode=torch.tensor([[1., 0.] if int(r['mode']) == 0 else [0., 1.]], device=trainer.device),
                bpm=torch.tensor([float(r['bpm'])], device=trainer.device), used=sdata.get_used_instruments(feats, unpitched),
                bpm_target=float(r['bpm']), path=path))
            if len(batch) == songs_per_step:
                leaves = trainer.accumulate_ragged(batch)
                lr = trainer.optimizer_step()
                steps += 1
                if on_step is not None:
                    on_step(steps, batch, leaves, lr)
                batch = []
                if max_steps is not None and steps >= max_steps:
                    break
    finally:
        pipe.close()
    return steps
