"""style/utils/pytorch.py of the reference: tensor helpers with the reference's names (utils/pytorch.py:7-94).

`LSTM` and `Distributed` are the classes the drop-in model builds its parameter containers from (style/model.py);
they are re-exported here because `torch.save(model)` pickles of the REFERENCE name them by this module path, so a
reference snapshot unpickles onto the CUDA-backed classes.  The functional helpers are plain torch glue for callers
that use them on their own tensors (the network's own reshapes / concatenations are fused into the kernels)."""
import math

import torch

from style.model import LSTM, Distributed     # noqa: F401  (pickle compatibility: style.utils.pytorch.LSTM / Distributed)


def squash_dims(tensor, dim_begin, dim_end=None):
    """Merge dims [dim_begin, dim_end) into one (utils/pytorch.py:7-16)."""
    nd = tensor.dim()
    if dim_end is None:
        dim_end = nd
    if dim_begin < 0:
        dim_begin, dim_end = dim_begin + nd, dim_end + nd
    shape = list(tensor.shape)
    return tensor.view(*shape[:dim_begin], math.prod(shape[dim_begin:dim_end]), *shape[dim_end:])


def cat_with_broadcast(tensors, dim=0):
    """torch.cat after expanding every other dim to the common (maximum) size (utils/pytorch.py:54-65)."""
    assert len(tensors) and all(t.dim() == tensors[0].dim() for t in tensors)
    target = [max(t.shape[i] for t in tensors) for i in range(tensors[0].dim())]
    parts = []
    for t in tensors:
        shape = list(target)
        shape[dim] = t.shape[dim]
        parts.append(t.expand(*shape))
    return torch.cat(parts, dim=dim)


def safe_sqrt(x):
    """sqrt with a zero (not NaN) gradient at 0 (utils/pytorch.py:68-71)."""
    if x == 0:
        return torch.tensor(0., requires_grad=x.requires_grad) * x
    return torch.sqrt(x)


def get_mean(tensors, weights=None, mean_type='arithmetic'):
    """Weighted arithmetic / harmonic / geometric / quadratic mean of a list of tensors (utils/pytorch.py:74-94)."""
    n = len(tensors)
    if weights is None:
        weights = [1.0 / n] * n
    if mean_type == 'arithmetic':
        return sum(w * t for t, w in zip(tensors, weights))
    if mean_type == 'harmonic':
        return 1 / get_mean([1 / t for t in tensors], weights=weights)
    if mean_type == 'geometric':
        return torch.stack(tensors).prod(0) ** (1 / n)
    if mean_type == 'quadratic':
        return safe_sqrt(get_mean([t ** 2 for t in tensors], weights=weights))
    raise ValueError(f'Unsupported mean type: {mean_type}')
