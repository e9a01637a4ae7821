"""CPU restatement of the metric side.  TEST INFRASTRUCTURE ONLY.

* ``accuracy``        restates ``/root/reference/torchbooster/metrics.py:27``.
* ``RunningAverage``  restates ``/root/reference/torchbooster/metrics.py:55-75``.
  Both are pinned by ``tests/golden/metrics_reference.npz``, produced by importing the
  reference module itself (``tests/golden/make_golden.py``).
* ``label_rank`` / ``topk_correct`` / ``confusion`` have **no reference counterpart**
  (SURVEY.md F1).  They are defined by the rank rule of SURVEY.md section 8(a) M3/M4 -- chosen so
  that k=1 coincides with ``argmax`` (first maximal index, NaN maximal) -- and are cross-checked
  against ``torch.topk`` / ``torch.bincount`` on tie-free data in ``tests/test_oracle.py``.
"""
from __future__ import annotations

import numpy as np
import torch


def accuracy(logits: torch.Tensor, labels: torch.Tensor, dim: int = -1) -> torch.Tensor:
    return (logits.argmax(dim=dim) == labels).sum() / logits.size(0)


class RunningAverage:
    def __init__(self) -> None:
        self.current = 0
        self.value = 0.

    def update(self, value: float) -> None:
        self.value = self.value * self.current + value
        self.current += 1
        self.value = self.value / self.current


def _order_key(logits: np.ndarray) -> np.ndarray:
    """Total pre-order used by the rank rule: NaN above +inf (all NaN equal), -0.0 == +0.0."""
    x = np.asarray(logits, dtype=np.float64)  # exact for fp16/bf16/fp32 inputs
    return np.where(np.isnan(x), np.inf, x), np.isnan(x)


def label_rank(logits, labels) -> np.ndarray:
    """``rank_n = #{j : l[n,j] > l[n,y]  or  (l[n,j] == l[n,y] and j < y)}``, NaN treated as
    greater than every number and equal to every other NaN.  Labels outside ``[0, C)`` get rank C."""
    if isinstance(logits, torch.Tensor):
        logits = logits.detach().cpu().to(torch.float64).numpy()
    labels = np.asarray(labels.cpu() if isinstance(labels, torch.Tensor) else labels).astype(np.int64)
    n, c = logits.shape
    val, isnan = _order_key(logits)
    # two-level key: (isnan, value)
    ok = (labels >= 0) & (labels < c)
    y = np.where(ok, labels, 0)
    rows = np.arange(n)
    vy, ny = val[rows, y][:, None], isnan[rows, y][:, None]
    greater = (isnan & ~ny) | ((isnan == ny) & ~isnan & (val > vy))
    equal = (isnan & ny) | (~isnan & ~ny & (val == vy))
    before = np.arange(c)[None, :] < y[:, None]
    rank = (greater | (equal & before)).sum(axis=1)
    return np.where(ok, rank, c).astype(np.int64)


def argmax_first(logits) -> np.ndarray:
    """First maximal index, NaN maximal -- the documented behaviour of ``torch.argmax`` on CPU."""
    if isinstance(logits, torch.Tensor):
        logits = logits.detach().cpu().to(torch.float64).numpy()
    val, isnan = _order_key(logits)
    anynan = isnan.any(axis=1)
    first_nan = isnan.argmax(axis=1)
    return np.where(anynan, first_nan, val.argmax(axis=1)).astype(np.int64)


def topk_correct(logits, labels, k: int) -> int:
    return int((label_rank(logits, labels) < k).sum())


def confusion(logits, labels, num_classes: int) -> np.ndarray:
    """``cm[y, argmax] += 1``; rows are true labels.  Samples with labels outside ``[0, C)`` are skipped."""
    labels = np.asarray(labels.cpu() if isinstance(labels, torch.Tensor) else labels).astype(np.int64)
    pred = argmax_first(logits)
    ok = (labels >= 0) & (labels < num_classes)
    flat = labels[ok] * num_classes + pred[ok]
    return np.bincount(flat, minlength=num_classes * num_classes).reshape(num_classes, num_classes).astype(np.int64)
