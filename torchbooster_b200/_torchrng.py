"""Replaying torch's CPU generator (mt19937) word by word.

``torch.manual_seed(s)`` seeds an mt19937 exactly like ``numpy.random.RandomState(s)``; the draws
the torchvision transforms make each spend one 32-bit word:

* ``torch.randint(0, n, (1,))``           -> ``word % n``                      (n < 2**32)
* ``torch.rand(1)``                       -> ``fl32(word & (2**24 - 1)) * 2**-24``
* ``torch.empty(1).uniform_(a, b)``       -> ``x * fl32(b - a) + a`` in float32 with ``x`` as above.
  The multiply-add is ONE fused operation in torch's AVX2 / AVX512 kernels and two roundings in
  its DEFAULT (no vector extension) build of the same kernel, so the result depends on
  ``torch.backends.cpu.get_cpu_capability()``; ``uniform`` follows whichever this process runs.
"""
from __future__ import annotations

import math
from fractions import Fraction

import numpy as np
import torch

MASK24 = np.uint64((1 << 24) - 1)
INV24 = np.float32(2.0 ** -24)


def fused_multiply_add() -> bool:
    """Does torch's CPU ``uniform_`` kernel contract ``x * (b - a) + a`` on this host?"""
    return str(torch.backends.cpu.get_cpu_capability()).upper() != "DEFAULT"


def unit_float32(words: np.ndarray) -> np.ndarray:
    return (np.asarray(words, np.uint64) & MASK24).astype(np.float32) * INV24


def _odd_power(v: float):
    """v = m * 2**e with m an odd integer (or 0)."""
    fr = Fraction(float(v))
    m, e = fr.numerator, -int(math.log2(fr.denominator))
    while m and m % 2 == 0:
        m //= 2
        e += 1
    return m, e


def uniform(words: np.ndarray, a: float, b: float, fused=None) -> np.ndarray:
    """float32 results of ``torch.empty(n).uniform_(a, b)`` element by element for the given words."""
    words = np.asarray(words, np.uint64)
    lo, span = np.float32(a), np.float32(np.float32(b) - np.float32(a))
    if fused is None:
        fused = fused_multiply_add()
    if not fused:
        return (unit_float32(words) * span).astype(np.float32) + lo
    # fused: round(k * 2**-24 * span + lo) once.  Exact integer numerator over a power-of-two
    # denominator, then one correctly rounded int64 -> float32 conversion.
    k = (words & MASK24).astype(np.int64)
    (ms, es), (ml, el) = _odd_power(float(span)), _odd_power(float(lo))
    es -= 24
    s = max(-es, -el, 0)
    if abs(ms) << (es + s + 24) >= 1 << 62 or abs(ml) << (el + s) >= 1 << 62:
        out = np.empty(words.shape, np.float32)        # exponents too far apart for int64: exact rationals
        flat = out.reshape(-1)
        for i, kk in enumerate(k.reshape(-1).tolist()):
            flat[i] = _round_fraction_to_float32(Fraction(kk) * Fraction(float(span)) / (1 << 24) + Fraction(float(lo)))
        return out
    num = k * np.int64(ms << (es + s)) + np.int64(ml << (el + s))
    return (num.astype(np.float32) * np.float32(2.0 ** -s)).astype(np.float32)


def _round_fraction_to_float32(fr: Fraction) -> np.float32:
    if fr == 0:
        return np.float32(0.0)
    sign, fr = (-1, -fr) if fr < 0 else (1, fr)
    e = fr.numerator.bit_length() - fr.denominator.bit_length()
    if Fraction(2) ** e > fr:
        e -= 1
    q = max(e - 23, -149)                                # unit in the last place (subnormal floor)
    scaled = fr / Fraction(2) ** q
    n = scaled.numerator // scaled.denominator
    rem = scaled - n
    if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and n % 2):
        n += 1
    return np.float32(sign * math.ldexp(n, q))
