"""Host helpers train-model.py imports from style.utils.misc (utils/misc.py of the reference): file iteration,
the EMA progress bar, recursive dict_map, directory creation."""
import glob
import os
from collections import defaultdict


def iter_all_files(path, pattern='**/*'):
    for x in glob.iglob(os.path.join(path, pattern), recursive=True):
        if os.path.isfile(x):
            yield x


def make_dirs(path):
    if path:
        os.makedirs(path, exist_ok=True)


def dict_map(f, d, recursive=False):
    return {k: (dict_map(f, v, recursive) if recursive and isinstance(v, dict) else f(v)) for k, v in d.items()}


class ProgressBar:
    """tqdm bar whose postfix shows an exponential moving average (momentum .99) of the values added."""

    def __init__(self, n_iterations=None, momentum=.99):
        from tqdm import tqdm
        self.n_iterations, self.momentum = n_iterations, momentum
        self.pbar = tqdm(total=n_iterations)
        self.values_sum, self.values_seen, self.avg_values = defaultdict(float), defaultdict(float), {}

    def add(self, n, **values):
        self.pbar.update(n)
        self.update_values(n, **values)
        if self.pbar.n == self.n_iterations:
            self.pbar.close()

    def update_values(self, n, **values):
        for k, v in values.items():
            if v is None:
                continue
            self.values_sum[k] = self.values_sum[k] * self.momentum + v * n
            self.values_seen[k] = self.values_seen[k] * self.momentum + n
        self.avg_values = {k: self.values_sum[k] / self.values_seen[k] for k in self.values_sum}
        self.pbar.set_postfix({k: f'{v:.2f}' for k, v in self.avg_values.items()})
