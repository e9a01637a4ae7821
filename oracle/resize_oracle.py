"""Integer restatement of Pillow's 8-bit antialiased bilinear resample.  TEST INFRASTRUCTURE ONLY.

Follows the algorithm torchvision reaches through ``F.resize`` on a PIL image
(``torchvision/transforms/_functional_pil.py:242-253`` -> ``Image.resize(..., BILINEAR)`` ->
libImaging ``ImagingResample``), which is what ``T.Resize(size)`` means at
``/root/reference/examples/img_stt/adain/adain.py:170`` and ``online/online.py:172``.
Pillow is a third-party dependency of the reference (unpinned; 12.2.0 in this image); its
published algorithm (libImaging/Resample.c: ``precompute_coeffs``, ``normalize_coeffs_8bpc``,
``ImagingResampleHorizontal_8bpc``, ``ImagingResampleVertical_8bpc``) is restated here and
pinned bit-for-bit against Pillow itself in ``tests/test_oracle.py``.
"""
from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2  # 22


def resized_output_size(h: int, w: int, size) -> tuple[int, int]:
    """torchvision ``_compute_resized_output_size`` (``functional.py:353-384``)."""
    if isinstance(size, int) or len(size) == 1:
        s = size if isinstance(size, int) else size[0]
        short, long = (w, h) if w <= h else (h, w)
        new_short, new_long = s, int(s * long / short)
        new_w, new_h = (new_short, new_long) if w <= h else (new_long, new_short)
        return new_h, new_w
    return int(size[0]), int(size[1])


def bilinear_taps(in_size: int, out_size: int):
    """``precompute_coeffs`` + ``normalize_coeffs_8bpc`` for the triangle filter (support 1.0)
    over the full box ``[0, in_size)``.

    Returns ``(ksize, bounds[out,2] (xmin, n), kk[out, ksize] int32)``.
    """
    scale = in_size / out_size
    filterscale = max(scale, 1.0)
    support = 1.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = np.zeros(ksize, np.float64)
        ww = 0.0
        for x in range(xmax):
            a = (x + xmin - center + 0.5) * ss
            if a < 0.0:
                a = -a
            v = 1.0 - a if a < 1.0 else 0.0
            w[x] = v
            ww += v
        for x in range(xmax):
            if ww != 0.0:
                w[x] /= ww
        for x in range(ksize):
            v = w[x]
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
        bounds[xx] = (xmin, xmax)
    return ksize, bounds, kk


def _clip8(acc: np.ndarray) -> np.ndarray:
    return np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)


def _pass_axis0(img: np.ndarray, out_size: int) -> np.ndarray:
    """One resample pass along axis 0 of ``img`` (uint8, any trailing shape)."""
    ksize, bounds, kk = bilinear_taps(img.shape[0], out_size)
    out = np.empty((out_size,) + img.shape[1:], np.uint8)
    src = img.astype(np.int64)
    for xx in range(out_size):
        xmin, n = bounds[xx]
        acc = np.full(img.shape[1:], 1 << (PRECISION_BITS - 1), np.int64)
        for x in range(n):
            acc += src[xmin + x] * int(kk[xx, x])
        out[xx] = _clip8(acc)
    return out


def resize_u8(img: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """``ImagingResample`` for ``[H, W, C]`` / ``[H, W]`` uint8: horizontal pass first into a
    **uint8-rounded** intermediate, then the vertical pass.  A pass whose size is unchanged is
    skipped (``need_horizontal`` / ``need_vertical``), and ``Image.resize`` returns a copy when
    nothing changes."""
    assert img.dtype == np.uint8
    h, w = img.shape[:2]
    cur = img
    if out_w != w:
        cur = np.swapaxes(_pass_axis0(np.swapaxes(cur, 0, 1), out_w), 0, 1)
    if out_h != h:
        cur = _pass_axis0(cur, out_h)
    return np.ascontiguousarray(cur)
