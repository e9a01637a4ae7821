"""Host side of ``T.RandomRotation`` / ``T.RandAugment`` on the GPU: which operation every sample
gets, and its integer encoding for ``tb_augment_u8_nhwc_to_nchw``.

The reference's CIFAR pipeline (``examples/img_cls/resnet/resnet.py:96-103``) runs, per sample on
a PIL image, ``RandomCrop -> RandomHorizontalFlip -> RandomRotation(15) -> RandAugment() ->
ToTensor -> Normalize``.  The pixel work of all six stages happens in one kernel; what is left for
the host is the part that is a handful of scalars per sample:

* ``AugmentParams``  replays the draws torchvision makes from torch's CPU generator
  (``RandomRotation.get_params``: one ``uniform_``; ``RandAugment.forward``,
  ``torchvision/transforms/autoaugment.py:340-365``: ``randint(14)`` and, for signed operations,
  ``randint(2)``), after the crop / flip draws of the same sample;
* ``encode_op``      turns (name, magnitude) into the record the kernel reads: the affine family
  becomes Pillow's six 16.16 fixed-point coefficients (matrix arithmetic in float64 with the very
  expressions ``Image.rotate`` / ``F.affine`` use), the enhancers become their float32 blend
  factor, posterize a bit mask, solarize an integer threshold.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _native as N
from . import _torchrng

Op = Tuple[str, float]

# operation codes of include/torchbooster_b200.h (TB_AUG_*)
IDENTITY, AFFINE, BRIGHTNESS, COLOR, CONTRAST, SHARPNESS, POSTERIZE, SOLARIZE, AUTOCONTRAST, EQUALIZE, INVERT = range(11)
OP_WORDS = 8            # int32 words per record: code, six parameters, reserved

# RandAugment._augmentation_space, in its dictionary order: name -> signed
RAND_AUGMENT_SPACE = (("Identity", False), ("ShearX", True), ("ShearY", True), ("TranslateX", True), ("TranslateY", True),
                      ("Rotate", True), ("Brightness", True), ("Color", True), ("Contrast", True), ("Sharpness", True),
                      ("Posterize", False), ("Solarize", False), ("AutoContrast", False), ("Equalize", False))
_ENHANCERS = {"Brightness": BRIGHTNESS, "Color": COLOR, "Contrast": CONTRAST, "Sharpness": SHARPNESS}


def rand_augment_magnitudes(height: int, width: int, magnitude: int = 9, num_bins: int = 31) -> dict:
    """The magnitude each operation receives (before the sign draw), as float32-exact Python floats:
    bin ``magnitude`` of the ``torch.linspace`` tables of ``RandAugment._augmentation_space``."""
    pick = lambda t: float(t[magnitude].item())
    lin = lambda hi: pick(torch.linspace(0.0, hi, num_bins))
    return {"Identity": 0.0, "ShearX": lin(0.3), "ShearY": lin(0.3),
            "TranslateX": lin(150.0 / 331.0 * width), "TranslateY": lin(150.0 / 331.0 * height),
            "Rotate": lin(30.0), "Brightness": lin(0.9), "Color": lin(0.9), "Contrast": lin(0.9), "Sharpness": lin(0.9),
            "Posterize": pick(8 - (torch.arange(num_bins) / ((num_bins - 1) / 4)).round().int()),
            "Solarize": pick(torch.linspace(255.0, 0.0, num_bins)), "AutoContrast": 0.0, "Equalize": 0.0}


# ------------------------------------------------------------------------------------------------
# encoding
# ------------------------------------------------------------------------------------------------
def _fix(v: float) -> int:
    return int(math.floor(v * 65536.0 + 0.5))


def _affine_record(a: Sequence[float], height: int, width: int) -> List[int]:
    """Pillow's nearest-neighbour affine in 16.16 fixed point: source x = (a2' + x a0 + y a1) >> 16 with the
    half-pixel centre folded into the offsets (libImaging/Geometry.c, affine_fixed)."""
    if a[1] == 0 and a[3] == 0 and not (a[0] == 1 and a[4] == 1 and float(a[2]).is_integer() and float(a[5]).is_integer()):
        raise NotImplementedError("axis-aligned scaling (Pillow's ImagingScaleAffine path) is only built for integer shifts")
    for x, y in ((0, 0), (width, height), (0, height), (width, 0)):
        if not (abs(x * a[0] + y * a[1] + a[2]) < 32768.0 and abs(x * a[3] + y * a[4] + a[5]) < 32768.0):
            raise NotImplementedError("affine coordinates beyond Pillow's fixed-point range")
    rec = [AFFINE, _fix(a[0]), _fix(a[1]), _fix(a[2] + a[0] * 0.5 + a[1] * 0.5),
           _fix(a[3]), _fix(a[4]), _fix(a[5] + a[3] * 0.5 + a[4] * 0.5), 0]
    return [rec[0]] + [((v + 2 ** 31) % 2 ** 32) - 2 ** 31 for v in rec[1:]]


def _rotation_record(angle: float, height: int, width: int) -> List[int]:
    """``Image.rotate(angle)`` with expand False, centre (w/2, h/2): multiples of 90 degrees are exact
    transpositions, anything else the rounded-to-15-decimals matrix."""
    one, half = 65536, 32768
    a = angle % 360.0
    if a == 0:
        return [IDENTITY] + [0] * 7
    if a == 180:
        return [AFFINE, -one, 0, (width - 1) * one + half, 0, -one, (height - 1) * one + half, 0]
    if a == 90 and width == height:
        return [AFFINE, 0, -one, (width - 1) * one + half, one, 0, half, 0]
    if a == 270 and width == height:
        return [AFFINE, 0, one, half, -one, 0, (height - 1) * one + half, 0]
    cx, cy = width / 2.0, height / 2.0
    r = -math.radians(a)
    m = [round(math.cos(r), 15), round(math.sin(r), 15), 0.0, round(-math.sin(r), 15), round(math.cos(r), 15), 0.0]
    m[2], m[5] = m[0] * -cx + m[1] * -cy + m[2], m[3] * -cx + m[4] * -cy + m[5]
    m[2] += cx
    m[5] += cy
    return _affine_record(m, height, width)


def _inverse_affine(center, angle, translate, scale, shear) -> List[float]:
    """The matrix ``F.affine`` hands to Pillow (torchvision ``functional.py:1006-1063``)."""
    rot, sx, sy = math.radians(angle), math.radians(shear[0]), math.radians(shear[1])
    (cx, cy), (tx, ty) = center, translate
    a = math.cos(rot - sy) / math.cos(sy)
    b = -math.cos(rot - sy) * math.tan(sx) / math.cos(sy) - math.sin(rot)
    c = math.sin(rot - sy) / math.cos(sy)
    d = -math.sin(rot - sy) * math.tan(sx) / math.cos(sy) + math.cos(rot)
    m = [v / scale for v in (d, -b, 0.0, -c, a, 0.0)]
    m[2] += m[0] * (-cx - tx) + m[1] * (-cy - ty)
    m[5] += m[3] * (-cx - tx) + m[4] * (-cy - ty)
    m[2] += cx
    m[5] += cy
    return m


def _float_bits(v: float) -> int:
    return int(np.float32(v).view(np.int32))


def encode_op(name: str, magnitude: float, height: int, width: int, channels: int) -> List[int]:
    """One operation -> 8 int32 words.  Names: ``RandomRotation`` (angle in degrees) and the
    operations of ``torchvision.transforms.autoaugment._apply_op`` with its magnitudes."""
    m = float(magnitude)
    if name == "Identity":
        return [IDENTITY] + [0] * 7
    if name in ("RandomRotation", "Rotate"):
        return _rotation_record(m, height, width)
    if name == "ShearX":
        return _affine_record(_inverse_affine([0, 0], 0.0, [0, 0], 1.0, [math.degrees(math.atan(m)), 0.0]), height, width)
    if name == "ShearY":
        return _affine_record(_inverse_affine([0, 0], 0.0, [0, 0], 1.0, [0.0, math.degrees(math.atan(m))]), height, width)
    if name == "TranslateX":
        return _affine_record(_inverse_affine([width * 0.5, height * 0.5], 0.0, [int(m), 0], 1.0, [0.0, 0.0]), height, width)
    if name == "TranslateY":
        return _affine_record(_inverse_affine([width * 0.5, height * 0.5], 0.0, [0, int(m)], 1.0, [0.0, 0.0]), height, width)
    if name in _ENHANCERS:
        if name == "Color" and channels == 1:
            return [IDENTITY] + [0] * 7         # a grey image is its own degenerate
        return [_ENHANCERS[name], _float_bits(1.0 + m)] + [0] * 6       # ImagingBlend takes a C float
    if name == "Posterize":
        return [POSTERIZE, (0xFF << (8 - int(m))) & 0xFF] + [0] * 6
    if name == "Solarize":
        return [SOLARIZE, max(0, min(256, math.ceil(m)))] + [0] * 6     # i < threshold  <=>  i < ceil(threshold)
    if name == "AutoContrast":
        return [AUTOCONTRAST] + [0] * 7
    if name == "Equalize":
        return [EQUALIZE] + [0] * 7
    if name == "Invert":
        return [INVERT] + [0] * 7
    raise ValueError(f"The provided operator {name} is not recognized.")     # autoaugment.py:88-89


def encode_ops(ops: Sequence[Sequence[Op]], height: int, width: int, channels: int) -> np.ndarray:
    """``ops[b]`` = the (name, magnitude) list of sample ``b`` -> int32 ``[B, n_ops, 8]`` (short lists are
    padded with Identity)."""
    n = max((len(o) for o in ops), default=0)
    out = np.zeros((len(ops), max(n, 1), OP_WORDS), np.int32)
    cache = {}
    for b, mine in enumerate(ops):
        for i, (name, magnitude) in enumerate(mine):
            key = (name, float(magnitude))
            rec = cache.get(key)
            if rec is None:
                rec = cache[key] = np.array(encode_op(name, magnitude, height, width, channels), np.int64).astype(np.int32)
            out[b, i] = rec
    return out


# ------------------------------------------------------------------------------------------------
# replay of torchvision's draws
# ------------------------------------------------------------------------------------------------
class AugmentParams:
    """Replays, for each sample in turn, the draws of
    ``RandomCrop -> RandomHorizontalFlip -> RandomRotation(degrees) -> RandAugment(num_ops, magnitude, bins)``
    from torch's CPU generator after ``torch.manual_seed(seed)`` (one worker process's order).
    Stages that are not in the pipeline draw nothing."""

    def __init__(self, seed: int, degrees=None, rand_augment: Optional[Tuple[int, int, int]] = (2, 9, 31)) -> None:
        self._rs = np.random.RandomState(int(seed) & 0xFFFFFFFF)
        if degrees is not None:
            if np.isscalar(degrees):
                if degrees < 0:
                    raise ValueError("If degrees is a single number, it must be positive.")   # transforms.py _setup_angle
                degrees = (-degrees, degrees)
            degrees = (float(degrees[0]), float(degrees[1]))
        self.degrees = degrees
        self.rand_augment = None if rand_augment is None else tuple(int(v) for v in rand_augment)
        self._words = np.zeros(0, np.uint64)

    def _take(self, n: int) -> np.ndarray:
        if self._words.shape[0] < n:
            more = self._rs.randint(0, 2 ** 32, size=max(n - self._words.shape[0], 4096), dtype=np.uint64)
            self._words = np.concatenate([self._words, more])
        return self._words

    def draw_records(self, batch: int, h: int, w: int, channels: int, crop: Optional[Tuple[int, int]],
                     padding: Tuple[int, int] = (0, 0), flip_p: Optional[float] = None):
        """The same walk as ``draw`` followed by ``encode_ops``, in one native pass
        (``tb_augment_params_walk``): returns ``(top, left, flip, records int32[B, n, 8])``."""
        px, py = padding
        ph, pw = h + 2 * py, w + 2 * px
        if crop is not None and (ph < crop[0] or pw < crop[1]):
            raise ValueError(f"Required crop size {tuple(crop)} is larger than input image size {(ph, pw)}")
        need_crop = crop is not None and not (ph == crop[0] and pw == crop[1])
        out_h, out_w = crop if crop is not None else (ph, pw)
        n_ops, magnitude, bins = self.rand_augment if self.rand_augment is not None else (0, 0, 0)
        per = (self.degrees is not None) + n_ops
        most = (2 if need_crop else 0) + (flip_p is not None) + (self.degrees is not None) + 2 * n_ops
        words = np.ascontiguousarray(self._take(batch * most))
        key = (out_h, out_w, channels, magnitude, bins)
        if n_ops and getattr(self, "_table_key", None) != key:
            mags = rand_augment_magnitudes(out_h, out_w, magnitude, bins)
            self._table = np.array([[encode_op(name, sign * mags[name], out_h, out_w, channels) for sign in (1.0, -1.0)]
                                    for name, _ in RAND_AUGMENT_SPACE], np.int64).astype(np.int32)
            self._signed = np.array([s for _, s in RAND_AUGMENT_SPACE], np.uint8)
            self._table_key = key
        top, left, flip = np.zeros(batch, np.int32), np.zeros(batch, np.int32), np.zeros(batch, np.uint8)
        rec = np.zeros((batch, max(per, 1), OP_WORDS), np.int32)
        lo, hi = self.degrees if self.degrees is not None else (0.0, 0.0)
        ptr = lambda a: a.ctypes.data
        used = N.lib().tb_augment_params_walk(
            ptr(words), words.shape[0], batch, ph - crop[0] + 1 if need_crop else 0, pw - crop[1] + 1 if need_crop else 0,
            -1.0 if flip_p is None else float(flip_p), int(self.degrees is not None), float(np.float32(lo)),
            float(np.float32(np.float32(hi) - np.float32(lo))), int(_torchrng.fused_multiply_add()), out_h, out_w,
            n_ops, len(RAND_AUGMENT_SPACE), ptr(self._table) if n_ops else None, ptr(self._signed) if n_ops else None,
            ptr(top), ptr(left), ptr(flip), ptr(rec) if per else None)
        if used < 0:
            N.check(int(used))
        self._words = words[used:]
        return top, left, flip, rec

    def draw(self, batch: int, h: int, w: int, crop: Optional[Tuple[int, int]], padding: Tuple[int, int] = (0, 0),
             flip_p: Optional[float] = None):
        """``h, w``: unpadded source size.  Returns ``(top int32[B], left int32[B], flip uint8[B], ops)`` with
        ``ops[b]`` the (name, magnitude) list of sample ``b``."""
        px, py = padding
        ph, pw = h + 2 * py, w + 2 * px
        if crop is not None and (ph < crop[0] or pw < crop[1]):
            raise ValueError(f"Required crop size {tuple(crop)} is larger than input image size {(ph, pw)}")
        need_crop = crop is not None and not (ph == crop[0] and pw == crop[1])
        out_h, out_w = crop if crop is not None else (ph, pw)
        n_ops, magnitude, bins = self.rand_augment if self.rand_augment is not None else (0, 0, 0)
        most = (2 if need_crop else 0) + (flip_p is not None) + (self.degrees is not None) + 2 * n_ops
        words = self._take(batch * most)
        mags = rand_augment_magnitudes(out_h, out_w, magnitude, bins) if n_ops else {}
        # every word's value under each reading it might get; the walk below only picks
        angle = _torchrng.uniform(words[:batch * most], *self.degrees) if self.degrees is not None else None
        unit = _torchrng.unit_float32(words[:batch * most]) if flip_p is not None else None
        wl = words[:batch * most].tolist()
        top, left, flip = np.zeros(batch, np.int32), np.zeros(batch, np.int32), np.zeros(batch, np.uint8)
        ops: List[List[Op]] = []
        rh, rw = (ph - crop[0] + 1, pw - crop[1] + 1) if need_crop else (1, 1)
        fp = None if flip_p is None else np.float32(flip_p)
        pos = 0
        for b in range(batch):
            if need_crop:
                top[b], left[b] = wl[pos] % rh, wl[pos + 1] % rw
                pos += 2
            if fp is not None:
                flip[b] = unit[pos] < fp
                pos += 1
            mine: List[Op] = []
            if angle is not None:
                mine.append(("RandomRotation", float(angle[pos])))
                pos += 1
            for _ in range(n_ops):
                name, signed = RAND_AUGMENT_SPACE[wl[pos] % len(RAND_AUGMENT_SPACE)]
                pos += 1
                m = mags[name]
                if signed:
                    if wl[pos] % 2:
                        m *= -1.0
                    pos += 1
                mine.append((name, m))
            ops.append(mine)
        self._words = words[pos:]
        return top, left, flip, ops


class ChainParams:
    """Replay for pipelines built around ``T.RandomResizedCrop`` -- per sample, in this order:
    ``RandomRotation(pre_rotation)`` -> ``RandomResizedCrop.get_params`` -> ``RandomHorizontalFlip`` ->
    ``RandomRotation(rotation)`` -> ``RandAugment`` (``examples/img_stt/online/online.py:173`` is
    pre_rotation=180, the resized crop, RandAugment).  Stages that are absent draw nothing.  A plain
    sequential walk in Python: these pipelines run at batch sizes (8 in ``online.yml``) where that is free.
    """

    def __init__(self, seed: int, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0), pre_rotation=None, rotation=None,
                 rand_augment: Optional[Tuple[int, int, int]] = None) -> None:
        self._rs = np.random.RandomState(int(seed) & 0xFFFFFFFF)
        self.scale = (float(scale[0]), float(scale[1]))
        self.ratio = (float(ratio[0]), float(ratio[1]))
        lr = torch.log(torch.tensor(self.ratio))                    # float32, as torchvision computes it
        self._log_ratio = (float(lr[0].item()), float(lr[1].item()))
        span = lambda d: None if d is None else ((-float(d), float(d)) if np.isscalar(d) else (float(d[0]), float(d[1])))
        self.pre_rotation, self.rotation = span(pre_rotation), span(rotation)
        self.rand_augment = None if rand_augment is None else tuple(int(v) for v in rand_augment)
        self._words: List[int] = []

    def _word(self) -> int:
        if not self._words:
            self._words = self._rs.randint(0, 2 ** 32, size=4096, dtype=np.uint64).tolist()[::-1]
        return self._words.pop()

    def _uniform(self, lo: float, hi: float) -> float:
        return float(_torchrng.uniform(np.array([self._word()], np.uint64), lo, hi)[0])

    def _box(self, height: int, width: int):
        area = height * width
        for _ in range(10):
            target_area = area * self._uniform(*self.scale)
            aspect = float(torch.exp(torch.tensor(np.float32(self._uniform(*self._log_ratio)))).item())
            w = int(round(math.sqrt(target_area * aspect)))
            h = int(round(math.sqrt(target_area / aspect)))
            if 0 < w <= width and 0 < h <= height:
                i = self._word() % (height - h + 1)
                j = self._word() % (width - w + 1)
                return i, j, h, w
        in_ratio = float(width) / float(height)
        if in_ratio < min(self.ratio):
            w = width
            h = int(round(w / min(self.ratio)))
        elif in_ratio > max(self.ratio):
            h = height
            w = int(round(h * max(self.ratio)))
        else:
            w, h = width, height
        return (height - h) // 2, (width - w) // 2, h, w

    def draw(self, batch: int, height: int, width: int, out_h: int, out_w: int, flip_p: Optional[float] = None):
        """Returns ``(box int32[B,4], flip uint8[B], ops, pre)``: ``ops[b]`` the (name, magnitude) list applied to
        the resized crop, ``pre[b]`` the one applied to the source image (``None`` when the stage is absent)."""
        box, flip = np.zeros((batch, 4), np.int32), np.zeros(batch, np.uint8)
        pre = [] if self.pre_rotation is not None else None
        ops = [] if (self.rotation is not None or self.rand_augment is not None) else None
        n_ops, magnitude, bins = self.rand_augment if self.rand_augment is not None else (0, 0, 0)
        mags = rand_augment_magnitudes(out_h, out_w, magnitude, bins) if n_ops else {}
        for b in range(batch):
            if pre is not None:
                pre.append([("RandomRotation", self._uniform(*self.pre_rotation))])
            box[b] = self._box(height, width)
            if flip_p is not None:
                flip[b] = _torchrng.unit_float32(np.array([self._word()], np.uint64))[0] < np.float32(flip_p)
            if ops is not None:
                mine: List[Op] = []
                if self.rotation is not None:
                    mine.append(("RandomRotation", self._uniform(*self.rotation)))
                for _ in range(n_ops):
                    name, signed = RAND_AUGMENT_SPACE[self._word() % len(RAND_AUGMENT_SPACE)]
                    m = mags[name]
                    if signed and self._word() % 2:
                        m *= -1.0
                    mine.append((name, m))
                ops.append(mine)
        return box, flip, ops, pre
