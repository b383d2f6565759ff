"""`metrics.ndcg` is called at style/scales.py:209 but its value never enters `loss`
(style/scales.py:188), so a constant keeps the reference's scale choice unchanged."""


def ndcg(*args, **kwargs):
    return 0.0
