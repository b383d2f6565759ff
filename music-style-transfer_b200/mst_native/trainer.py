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

