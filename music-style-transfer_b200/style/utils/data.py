"""style.utils.data of the reference: CSV logging of the flattened loss dict (utils/data.py:27-46)."""
import csv
import os


def assert_dir(path):
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)


def save_to_csv(path, **row):
    assert_dir(path)
    new = not os.path.exists(path)
    with open(path, 'a', newline='') as f:
        w = csv.DictWriter(f, fieldnames=list(row))
        if new:
            w.writeheader()
        w.writerow(row)


def flatten_dict(d, prefix=''):
    """flatten_dict.flatten(d, reducer='underscore') for the nested loss dict (train-model.py:148)."""
    out = {}
    for k, v in d.items():
        key = f'{prefix}_{k}' if prefix else k
        if isinstance(v, dict):
            out.update(flatten_dict(v, key))
        else:
            out[key] = v
    return out
