"""RandomRotation / RandAugment on PIL images: torchvision itself, and an integer restatement.
TEST INFRASTRUCTURE ONLY.

The reference's CIFAR pipeline (``/root/reference/examples/img_cls/resnet/resnet.py:96-103``) is
``RandomCrop(32, 4, reflect) -> RandomHorizontalFlip -> RandomRotation(15) -> RandAugment() ->
ToTensor -> Normalize``, per sample, on PIL images.  Two statements of the two middle stages:

* ``apply_ops``  -- literally torchvision: ``F.rotate`` for ``T.RandomRotation.forward``
  (``torchvision/transforms/transforms.py`` RandomRotation) and ``autoaugment._apply_op`` for each
  RandAugment operation (``torchvision/transforms/autoaugment.py:13-90``), NEAREST, fill None.
* ``restated_op`` -- what Pillow 12.2.0's C code does for each of them, in numpy (the statement the
  CUDA kernel is written against); pinned to ``apply_ops`` in ``tests/test_oracle.py``:
  nearest affine = 16.16 fixed point stepped from the pixel centre (``Geometry.c`` affine_fixed;
  pure translations take the scale path, which agrees with it for integer shifts);
  enhance = ``ImagingBlend`` in float32 ``deg + alpha * (img - deg)``, truncated, clipped when
  alpha is outside [0, 1]; grey = ``(19595 R + 38470 G + 7471 B + 0x8000) >> 16``;
  SMOOTH = 3x3 (1 1 1 / 1 5 1 / 1 1 1) / 13 + 0.5 truncated, border pixels copied;
  autocontrast / equalize = ``ImageOps`` tables from per-band histograms.

``draw_augment`` runs torchvision's own draws under ``torch.manual_seed`` (the order one worker
process makes them) with ``_apply_op`` intercepted to record (name, magnitude).
"""
from __future__ import annotations

import math
from typing import List, Sequence, Tuple

import numpy as np
import torch
import torchvision.transforms as T
import torchvision.transforms.functional as F
from PIL import Image
from torchvision.transforms import InterpolationMode
from torchvision.transforms import autoaugment as _aa

Op = Tuple[str, float]

RAND_AUGMENT_OPS = ("Identity", "ShearX", "ShearY", "TranslateX", "TranslateY", "Rotate", "Brightness", "Color",
                    "Contrast", "Sharpness", "Posterize", "Solarize", "AutoContrast", "Equalize")


def _to_pil(img: np.ndarray) -> Image.Image:
    return Image.fromarray(img[:, :, 0], mode="L") if img.shape[2] == 1 else Image.fromarray(img)


def apply_ops_pil(pil: Image.Image, ops: Sequence[Op]) -> Image.Image:
    for name, magnitude in ops:
        if name == "RandomRotation":        # T.RandomRotation.forward: F.rotate(img, angle, NEAREST, False, None, 0)
            pil = F.rotate(pil, float(magnitude), InterpolationMode.NEAREST, False, None, 0)
        else:
            pil = _aa._apply_op(pil, name, float(magnitude), interpolation=InterpolationMode.NEAREST, fill=None)
    return pil


def apply_ops(img: np.ndarray, ops: Sequence[Op]) -> np.ndarray:
    """``[H,W,C]`` uint8 -> ``[H,W,C]`` uint8 through torchvision's PIL backend."""
    H, W, C = img.shape
    return np.asarray(apply_ops_pil(_to_pil(img), ops)).reshape(H, W, C).copy()


def draw_augment(seed: int, batch: int, h: int, w: int, th: int, tw: int, padding: int = 0, p: float = 0.5,
                 degrees=None, num_ops: int = 2, magnitude: int = 9, num_magnitude_bins: int = 31,
                 rand_augment: bool = True):
    """Per sample, in pipeline order: RandomCrop.get_params, ``rand(1) < p``,
    RandomRotation.get_params, RandAugment.forward's draws.  Returns ``(top, left, flip, ops)``
    with ``ops[b]`` the list of (name, magnitude) that sample receives."""
    state = torch.get_rng_state()
    recorded: List[Op] = []
    original = _aa._apply_op

    def record(img, op_name, magnitude, interpolation, fill):
        recorded.append((op_name, float(magnitude)))
        return img

    try:
        torch.manual_seed(seed)
        _aa._apply_op = record
        padded = Image.new("L", (w + 2 * padding, h + 2 * padding))
        cropped = Image.new("RGB", (tw, th))
        ra = T.RandAugment(num_ops, magnitude, num_magnitude_bins) if rand_augment else None
        top, left, flip, ops = np.zeros(batch, np.int32), np.zeros(batch, np.int32), np.zeros(batch, np.uint8), []
        for b in range(batch):
            i, j, _, _ = T.RandomCrop.get_params(padded, (th, tw))
            top[b], left[b] = i, j
            flip[b] = 1 if bool(torch.rand(1) < p) else 0
            mine: List[Op] = []
            if degrees is not None:
                lo, hi = (-degrees, degrees) if np.isscalar(degrees) else degrees
                mine.append(("RandomRotation", T.RandomRotation.get_params([float(lo), float(hi)])))
            if ra is not None:
                recorded.clear()
                ra(cropped)
                mine.extend(recorded)
            ops.append(mine)
    finally:
        _aa._apply_op = original
        torch.set_rng_state(state)
    return top, left, flip, ops


# ------------------------------------------------------------------------------------------------
# restatement of Pillow's arithmetic
# ------------------------------------------------------------------------------------------------
def _fix(v: float) -> int:
    return int(math.floor(v * 65536.0 + 0.5))


def fixed_coefficients(a: Sequence[float]) -> Tuple[int, ...]:
    """16.16 coefficients of Pillow's nearest affine: the offsets carry the half-pixel centre."""
    return (_fix(a[0]), _fix(a[1]), _fix(a[2] + a[0] * 0.5 + a[1] * 0.5),
            _fix(a[3]), _fix(a[4]), _fix(a[5] + a[3] * 0.5 + a[4] * 0.5))


def affine_nearest(img: np.ndarray, a: Sequence[float]) -> np.ndarray:
    H, W, _ = img.shape
    if a[1] == 0 and a[3] == 0:      # Pillow's scale path; only unit scale with integer shifts occurs here
        assert a[0] == 1 and a[4] == 1 and float(a[2]).is_integer() and float(a[5]).is_integer()
    a0, a1, a2, a3, a4, a5 = fixed_coefficients(a)
    y, x = np.mgrid[0:H, 0:W].astype(np.int64)
    xin, yin = (a2 + x * a0 + y * a1) >> 16, (a5 + x * a3 + y * a4) >> 16
    ok = (xin >= 0) & (xin < W) & (yin >= 0) & (yin < H)
    out = np.zeros_like(img)
    out[ok] = img[yin[ok], xin[ok]]
    return out


def rotation_matrix(angle: float, w: int, h: int) -> List[float]:
    """``Image.rotate`` (expand False, centre = (w/2, h/2)): entries rounded to 15 decimals first."""
    cx, cy = w / 2.0, h / 2.0
    r = -math.radians(angle % 360.0)
    m = [round(math.cos(r), 15), round(math.sin(r), 15), 0.0, round(-math.sin(r), 15), round(math.cos(r), 15), 0.0]
    m[2], m[5] = m[0] * -cx + m[1] * -cy + m[2], m[3] * -cx + m[4] * -cy + m[5]
    m[2] += cx
    m[5] += cy
    return m


def inverse_affine_matrix(center, angle, translate, scale, shear) -> List[float]:
    """``torchvision.transforms.functional._get_inverse_affine_matrix`` (``functional.py:1006-1063``)."""
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


def rotate(img: np.ndarray, angle: float) -> np.ndarray:
    H, W, _ = img.shape
    a = angle % 360.0
    if a == 0:
        return img.copy()
    if a == 180:
        return img[::-1, ::-1].copy()
    if a in (90, 270) and H == W:
        return np.rot90(img, 1 if a == 90 else 3).copy()
    return affine_nearest(img, rotation_matrix(angle, W, H))


def blend(degenerate: np.ndarray, img: np.ndarray, alpha: float) -> np.ndarray:
    alpha = np.float32(alpha)
    t = degenerate.astype(np.float32) + alpha * (img.astype(np.int32) - degenerate.astype(np.int32)).astype(np.float32)
    assert t.dtype == np.float32
    if 0 <= alpha <= 1:
        return t.astype(np.int32).astype(np.uint8)
    return np.where(t <= 0, 0, np.where(t >= 255, 255, t.astype(np.int32))).astype(np.uint8)


def grey(img: np.ndarray) -> np.ndarray:
    if img.shape[2] == 1:
        return img[..., 0]
    r, g, b = (img[..., i].astype(np.int64) for i in range(3))
    return ((r * 19595 + g * 38470 + b * 7471 + 0x8000) >> 16).astype(np.uint8)


def smooth(img: np.ndarray) -> np.ndarray:
    H, W, C = img.shape
    out = img.copy()
    if H < 3 or W < 3:
        return out
    k = (np.array([1, 1, 1, 1, 5, 1, 1, 1, 1], np.float32) / np.float32(13)).astype(np.float32)
    f = img.astype(np.float32)
    ss = np.full((H - 2, W - 2, C), 0.5, np.float32)
    for ki, r in enumerate((2, 1, 0)):
        row = f[r:r + H - 2]
        ss = ss + (row[:, 0:W - 2] * k[3 * ki] + row[:, 1:W - 1] * k[3 * ki + 1] + row[:, 2:W] * k[3 * ki + 2])
    out[1:-1, 1:-1] = np.clip(np.floor(ss), 0, 255).astype(np.uint8)
    return out


def grey_mean(img: np.ndarray) -> int:
    """``int(ImageStat.Stat(img.convert("L")).mean[0] + 0.5)``."""
    g = grey(img).astype(np.int64)
    return int(float(int(g.sum())) / float(g.size) + 0.5)


def autocontrast_table(hist: Sequence[int]) -> np.ndarray:
    nz = [i for i in range(256) if hist[i]]
    lo, hi = nz[0], nz[-1]
    if hi <= lo:
        return np.arange(256, dtype=np.uint8)
    scale = 255.0 / (hi - lo)
    offset = -lo * scale
    return np.array([min(255, max(0, int(i * scale + offset))) for i in range(256)], np.uint8)


def equalize_table(hist: Sequence[int]) -> np.ndarray:
    h = [int(v) for v in hist]
    present = [v for v in h if v]
    if len(present) <= 1:
        return np.arange(256, dtype=np.uint8)
    step = (sum(present) - present[-1]) // 255
    if not step:
        return np.arange(256, dtype=np.uint8)
    n, table = step // 2, []
    for i in range(256):
        table.append(min(n // step, 255))
        n += h[i]
    return np.array(table, np.uint8)


def restated_op(img: np.ndarray, name: str, magnitude: float) -> np.ndarray:
    H, W, C = img.shape
    if name == "Identity":
        return img
    if name == "RandomRotation" or name == "Rotate":
        return rotate(img, magnitude)
    if name == "ShearX":
        return affine_nearest(img, inverse_affine_matrix([0, 0], 0.0, [0, 0], 1.0, [math.degrees(math.atan(magnitude)), 0.0]))
    if name == "ShearY":
        return affine_nearest(img, inverse_affine_matrix([0, 0], 0.0, [0, 0], 1.0, [0.0, math.degrees(math.atan(magnitude))]))
    if name == "TranslateX":
        return affine_nearest(img, inverse_affine_matrix([W * 0.5, H * 0.5], 0.0, [int(magnitude), 0], 1.0, [0.0, 0.0]))
    if name == "TranslateY":
        return affine_nearest(img, inverse_affine_matrix([W * 0.5, H * 0.5], 0.0, [0, int(magnitude)], 1.0, [0.0, 0.0]))
    if name == "Brightness":
        return blend(np.zeros_like(img), img, 1.0 + magnitude)
    if name == "Color":
        return img if C == 1 else blend(np.repeat(grey(img)[..., None], C, 2), img, 1.0 + magnitude)
    if name == "Contrast":
        return blend(np.full_like(img, grey_mean(img)), img, 1.0 + magnitude)
    if name == "Sharpness":
        return blend(smooth(img), img, 1.0 + magnitude)
    if name == "Posterize":
        return img & np.uint8((0xFF << (8 - int(magnitude))) & 0xFF)
    if name == "Solarize":
        return np.where(img < magnitude, img, 255 - img).astype(np.uint8)
    if name == "Invert":
        return (255 - img).astype(np.uint8)
    if name in ("AutoContrast", "Equalize"):
        make = autocontrast_table if name == "AutoContrast" else equalize_table
        out = img.copy()
        for c in range(C):
            out[..., c] = make(np.bincount(img[..., c].ravel(), minlength=256))[img[..., c]]
        return out
    raise KeyError(name)


def restated_ops(img: np.ndarray, ops: Sequence[Op]) -> np.ndarray:
    for name, magnitude in ops:
        img = restated_op(img, name, float(magnitude))
    return img
