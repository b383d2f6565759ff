"""Stub for `flatten_dict` (used only by the reference's logging: train-model.py:4,148;
style/utils/data.py:5,13).  Test infrastructure only."""


def flatten(d, reducer='tuple', _prefix=()):
    out = {}
    for k, v in d.items():
        key = _prefix + (k,)
        if isinstance(v, dict):
            out.update(flatten(v, reducer='tuple', _prefix=key))
        else:
            out[key] = v
    if _prefix or reducer == 'tuple':
        return out
    sep = {'path': '/', 'underscore': '_', 'dot': '.'}[reducer]
    return {sep.join(map(str, k)): v for k, v in out.items()}
