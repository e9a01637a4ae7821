"""Metric side of the hot path -- drop-ins for ``torchbooster/metrics.py`` plus the device
accumulators the north star names.

Drop-in surface (same names, arguments and results as the reference):
``accuracy`` (``torchbooster/metrics.py:11-27``), ``Accuracy`` (``:30-52``),
``RunningAverage`` (``:55-75``).

Additions (no reference counterpart): ``MetricState`` -- top-1 / top-k correct
counts and a confusion matrix accumulated on the device by one CUDA kernel per batch
(``tb_metrics_update``), reduced across ranks with one ``all_reduce`` of a packed int64 buffer;
``top_k_accuracy``; ``confusion_matrix``; ``DeviceRunningAverage``.

Tie rule.  ``rank(y) = #{j : l[j] > l[y] or (l[j] == l[y] and j < y)}``, NaN maximal; a sample is
top-k correct when ``rank(y) < k``.  k = 1 coincides with ``argmax`` (first maximal index), so
``MetricState.top1`` equals the reference's accuracy count bit for bit.  ``torch.topk`` breaks
ties in an implementation-defined order, so on exact ties at the k-th value membership may
differ from ``torch.topk``; without such ties it is identical.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import Tensor
from torch.nn import Module

from . import _native as N


def _rows(logits: Tensor, labels: Tensor, dim: int):
    """View ``logits`` as ``[M, C]`` with the class axis last and ``labels`` as ``[M]`` int64."""
    if logits.dim() == 0:
        raise IndexError("Dimension out of range")
    if dim < 0:
        dim += logits.dim()
    if dim != logits.dim() - 1:
        logits = logits.movedim(dim, -1)
    lead = logits.shape[:-1]
    c = logits.shape[-1]
    labels = torch.broadcast_to(labels, lead) if labels.shape != lead else labels
    logits2 = logits.reshape(-1, c)
    if logits2.stride(-1) != 1:
        logits2 = logits2.contiguous()
    labels1 = labels.reshape(-1).to(device=logits.device, dtype=torch.int64).contiguous()
    return logits2, labels1


def _launch_update(logits2: Tensor, labels1: Tensor, k: int, counters: Tensor, confusion: Optional[Tensor]) -> None:
    N.require_cuda(logits2, "logits")
    if logits2.dtype not in (torch.float32, torch.bfloat16, torch.float16):
        logits2 = logits2.float()
    m, c = logits2.shape
    with torch.cuda.device(logits2.device):
        rc = N.lib().tb_metrics_update(
            logits2.data_ptr(), N.dtype_code(logits2.dtype), m, c, logits2.stride(0) if m > 1 else c,
            labels1.data_ptr(), k, counters.data_ptr(),
            None if confusion is None else confusion.data_ptr(), N.current_stream_ptr(logits2.device))
    N.check(rc)


def accuracy(logits: Tensor, labels: Tensor, dim: int = -1) -> Tensor:
    """Mean batch accuracy, a 0-dim float32 tensor on ``logits.device`` -- the reference's
    ``(logits.argmax(dim=dim) == labels).sum() / logits.size(0)`` (``torchbooster/metrics.py:27``),
    computed by one kernel and without a host sync.  As in the reference the denominator is
    ``logits.size(0)`` whatever the rank of ``logits``.

    Parameters
    ----------
    logits, labels: Tensor
        ``(num_samples, num_classes)`` and ``(num_samples,)``
    dim: int (default: -1)
        dimension the argmax is taken over
    """
    logits2, labels1 = _rows(logits, labels, dim)
    counters = torch.zeros(4, dtype=torch.int64, device=logits2.device)
    _launch_update(logits2, labels1, 1, counters, None)
    return counters[1] / logits.size(0)


class Accuracy(Module):
    """Module form of :func:`accuracy` (``torchbooster/metrics.py:30-52``)."""

    def __init__(self) -> None:
        super().__init__()

    def forward(self, logits: Tensor, labels: Tensor, dim: int = -1) -> Tensor:
        return accuracy(logits, labels, dim=dim)


class RunningAverage:
    """Sequential running mean on the host, ``old <- (old * t + new) / (t + 1)`` in float64
    exactly as ``torchbooster/metrics.py:55-75`` (attributes ``current`` and ``value``)."""

    def __init__(self) -> None:
        self.current = 0
        self.value = 0.

    def update(self, value: float) -> None:
        total = self.value * self.current + value
        self.current += 1
        self.value = total / self.current


class DeviceRunningAverage:
    """Running mean kept on the device: no ``.item()`` per step.

    The reference feeds ``RunningAverage`` with ``loss.item()`` / ``acc.item()`` every step
    (``examples/img_cls/resnet/resnet.py:66-67``), i.e. two host syncs per step.  Here ``update``
    enqueues a tiny kernel adding ``sum(x)`` and ``numel(x)`` to a float64 ``[sum, count]`` state;
    ``value`` syncs only when read.  The result is the plain mean ``sum / count`` (float64), which
    differs from the sequential recurrence by rounding only; use ``RunningAverage`` for the
    bit-exact drop-in.
    """

    def __init__(self, device=None) -> None:
        self.device = N.cuda_device(device)
        self.state = torch.zeros(2, dtype=torch.float64, device=self.device)

    def update(self, value: Tensor) -> None:
        N.require_cuda(value, "value")
        v = value.detach().reshape(-1)
        if v.dtype not in (torch.float32, torch.bfloat16, torch.float16):
            v = v.float()
        if not v.is_contiguous():
            v = v.contiguous()
        with torch.cuda.device(self.device):
            N.check(N.lib().tb_running_sum_update(v.data_ptr(), N.dtype_code(v.dtype), v.numel(),
                                                  self.state.data_ptr(), N.current_stream_ptr(self.device)))

    @property
    def current(self) -> int:
        return int(self.state[1].item())

    @property
    def value(self) -> float:
        s, n = self.state.tolist()
        return s / n if n else 0.

    def reset(self) -> None:
        self.state.zero_()


class MetricState:
    """Device-resident classification metrics: ``n``, top-1 and top-k correct counts and the
    ``[C, C]`` confusion matrix (rows = true label, columns = argmax), in ONE packed int64 buffer
    ``[n, correct_1, correct_k, invalid_labels, cm...]`` so that the cross-rank reduction is a
    single ``all_reduce(SUM)``; plus a float64 ``[loss_sum, loss_count]`` pair.
    """

    HEADER = 4

    def __init__(self, num_classes: int, k: int = 5, device=None, confusion: bool = True) -> None:
        if num_classes < 1 or k < 1:
            raise ValueError("num_classes and k must be >= 1")
        self.num_classes, self.k, self.with_confusion = int(num_classes), int(k), bool(confusion)
        self.device = N.cuda_device(device, allow_cpu=True)   # CPU state: reduction tests only; update() needs CUDA
        n = self.HEADER + (self.num_classes ** 2 if confusion else 0)
        self.buffer = torch.zeros(n, dtype=torch.int64, device=self.device)
        self.loss = torch.zeros(2, dtype=torch.float64, device=self.device)
        self.rows_fed = 0          # host-side count of the rows scored since the last reset (bounds every bin)
        self._wire = None          # int32 transport buffer of the exchange, allocated at its first use

    @property
    def counters(self) -> Tensor:
        return self.buffer[:self.HEADER]

    @property
    def confusion(self) -> Optional[Tensor]:
        if not self.with_confusion:
            return None
        return self.buffer[self.HEADER:].view(self.num_classes, self.num_classes)

    def reset(self) -> None:
        self.buffer.zero_()
        self.loss.zero_()
        self.rows_fed = 0

    def note_rows(self, rows: int) -> None:
        """Tell the state about rows that were scored into ``buffer`` directly through the C ABI (``update`` counts
        its own): the int32 transport of ``all_reduce`` checks this count against its exactness limit."""
        self.rows_fed += int(rows)

    def update(self, logits: Tensor, labels: Tensor, dim: int = -1) -> None:
        """Accumulate one batch (enqueue only; no host sync)."""
        logits2, labels1 = _rows(logits, labels, dim)
        if logits2.shape[1] != self.num_classes:
            raise ValueError(f"logits have {logits2.shape[1]} classes, state has {self.num_classes}")
        if logits2.device != self.device:
            raise ValueError(f"logits on {logits2.device}, state on {self.device}")
        cm = self.buffer[self.HEADER:] if self.with_confusion else None
        _launch_update(logits2, labels1, self.k, self.buffer, cm)
        self.rows_fed += int(logits2.shape[0])

    def update_loss(self, loss: Tensor) -> None:
        """Add per-step (0-dim) or per-sample losses to the running sum (enqueue only)."""
        N.require_cuda(loss, "loss")
        v = loss.detach().reshape(-1)
        if v.dtype not in (torch.float32, torch.bfloat16, torch.float16):
            v = v.float()
        if not v.is_contiguous():
            v = v.contiguous()
        with torch.cuda.device(self.device):
            N.check(N.lib().tb_running_sum_update(v.data_ptr(), N.dtype_code(v.dtype), v.numel(),
                                                  self.loss.data_ptr(), N.current_stream_ptr(self.device)))

    def all_reduce(self, group=None, transport: str = "int64") -> None:
        """Sum the state over all ranks (``torchbooster_b200.distributed.all_reduce_metrics``).  ``transport="int32"``
        halves the bytes on the wire (same totals; needs fewer than 2**31 rows scored by all ranks together)."""
        from .distributed import all_reduce_metrics
        all_reduce_metrics(self, group=group, transport=transport)

    def wire(self) -> Tensor:
        """The int32 transport buffer ``[12 + bins]`` (see ``tb_metric_state_pack_i32``)."""
        if self._wire is None:
            self._wire = torch.empty(12 + self.buffer.numel() - self.HEADER, dtype=torch.int32, device=self.device)
        return self._wire

    def pack_wire(self) -> Tensor:
        N.require_cuda(self.buffer, "MetricState")
        w = self.wire()
        with torch.cuda.device(self.device):
            N.check(N.lib().tb_metric_state_pack_i32(self.buffer.data_ptr(), self.buffer.numel() - self.HEADER, w.data_ptr(),
                                                     N.current_stream_ptr(self.device)))
        return w

    def unpack_wire(self) -> None:
        w = self.wire()
        with torch.cuda.device(self.device):
            N.check(N.lib().tb_metric_state_unpack_i32(w.data_ptr(), self.buffer.numel() - self.HEADER, self.buffer.data_ptr(),
                                                       N.current_stream_ptr(self.device)))

    def compute(self) -> dict:
        """Read the state back (one sync)."""
        head = self.buffer[:self.HEADER].tolist()
        n, c1, ck, invalid = head
        loss_sum, loss_n = self.loss.tolist()
        return {
            "n": n, "correct_1": c1, f"correct_{self.k}": ck, "invalid_labels": invalid,
            "top1": c1 / n if n else 0.0, f"top{self.k}": ck / n if n else 0.0,
            "confusion": self.confusion.clone() if self.with_confusion else None,
            "loss": loss_sum / loss_n if loss_n else 0.0,
        }


def top_k_accuracy(logits: Tensor, labels: Tensor, k: int = 5, dim: int = -1) -> Tensor:
    """Fraction of samples whose label ranks below ``k`` (rank rule above); 0-dim float32 tensor."""
    logits2, labels1 = _rows(logits, labels, dim)
    counters = torch.zeros(4, dtype=torch.int64, device=logits2.device)
    _launch_update(logits2, labels1, k, counters, None)
    return counters[2] / logits.size(0)


def confusion_matrix(logits: Tensor, labels: Tensor, num_classes: Optional[int] = None, dim: int = -1) -> Tensor:
    """``cm[y, argmax] += 1`` as an int64 ``[C, C]`` tensor (``torch.bincount(y*C+pred).view(C,C)``)."""
    logits2, labels1 = _rows(logits, labels, dim)
    c = logits2.shape[1]
    if num_classes is not None and num_classes != c:
        raise ValueError(f"num_classes={num_classes} but logits have {c} classes")
    counters = torch.zeros(4, dtype=torch.int64, device=logits2.device)
    cm = torch.zeros(c * c, dtype=torch.int64, device=logits2.device)
    _launch_update(logits2, labels1, 1, counters, cm)
    return cm.view(c, c)
