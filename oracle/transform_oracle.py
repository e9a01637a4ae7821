"""Per-sample torchvision/PIL restatement of the batch transform.  TEST INFRASTRUCTURE ONLY.

This *is* the reference's data path: TorchBooster's examples hand PIL images to
``torchvision.transforms`` pipelines inside DataLoader workers
(``/root/reference/examples/img_cls/resnet/resnet.py:94-107``,
``examples/img_stt/adain/adain.py:167-171``, ``examples/img_gen/gan/gan.py:99,135``) and collate
with ``torch.stack`` (``torchbooster/config.py:371-379``).  Here the random parameters are
explicit inputs so that the CUDA path and the oracle see the same crop/flip decisions.
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np
import torch
import torchvision.transforms.functional as F
from PIL import Image


def _to_pil(img: np.ndarray) -> Image.Image:
    if img.ndim == 3 and img.shape[2] == 1:
        return Image.fromarray(img[:, :, 0], mode="L")
    return Image.fromarray(img)


def transform_sample(
    img: np.ndarray,
    *,
    resize=None,
    center_crop=None,
    padding: int = 0,
    padding_mode: str = "constant",
    crop: Optional[Sequence[int]] = None,   # (top, left, h, w) in padded coordinates
    resized_crop: Optional[Sequence[int]] = None,   # (top, left, h, w, out_h, out_w): T.RandomResizedCrop.forward
    flip: bool = False,
    ops=None,                                       # [(name, magnitude)]: RandomRotation / RandAugment operations
    pre_ops=None,                                   # the same, applied to the source image first (online.py:173)
    mean=None,
    std=None,
    one_minus: bool = False,
    dtype: torch.dtype = torch.float32,
) -> torch.Tensor:
    """Resize -> CenterCrop -> pad -> crop -> hflip -> [rotation, RandAugment ops] -> ToTensor ->
    Normalize -> (1 - X) -> cast.

    Each stage is the torchvision functional the matching ``T.*`` class calls:
    ``T.Resize`` (``transforms.py:354``), ``T.CenterCrop`` (``:381``), ``T.RandomCrop.forward``
    (``:657-690``: ``F.pad`` then ``F.crop(i, j, h, w)``), ``T.RandomHorizontalFlip.forward``
    (``:712-722``), ``T.ToTensor`` (``functional.py:127-178``), ``T.Normalize``
    (``_functional_tensor.py:905-928``).
    """
    pil = _to_pil(img)
    if pre_ops:
        from .augment_oracle import apply_ops_pil
        pil = apply_ops_pil(pil, pre_ops)
    if resize is not None:
        pil = F.resize(pil, resize)
    if center_crop is not None:
        pil = F.center_crop(pil, center_crop)
    if padding:
        pil = F.pad(pil, padding, 0, padding_mode)
    if crop is not None:
        pil = F.crop(pil, *[int(v) for v in crop])
    if resized_crop is not None:
        i, j, h, w, oh, ow = [int(v) for v in resized_crop]
        pil = F.resized_crop(pil, i, j, h, w, [oh, ow])
    if flip:
        pil = F.hflip(pil)
    if ops:
        from .augment_oracle import apply_ops_pil       # resnet.py:99-100: T.RandomRotation, T.RandAugment
        pil = apply_ops_pil(pil, ops)
    t = F.to_tensor(pil)
    if mean is not None:
        t = F.normalize(t, list(mean), list(std))
    if one_minus:
        t = 1.0 - t     # gan.py:99 / vae.py:107
    return t.to(dtype)


def transform_batch(batch: np.ndarray, *, crop_top=None, crop_left=None, out_size=None,
                    flip=None, boxes=None, ops=None, pre_ops=None, **kw) -> torch.Tensor:
    """``[B,H,W,C]`` uint8 -> ``[B,C,h,w]``; default collate = ``torch.stack``."""
    if isinstance(batch, torch.Tensor):
        batch = batch.cpu().numpy()
    out = []
    for b in range(batch.shape[0]):
        crop = None
        if out_size is not None and crop_top is not None:
            crop = (int(crop_top[b]), int(crop_left[b]), int(out_size[0]), int(out_size[1]))
        f = bool(flip[b]) if flip is not None else False
        rc = None
        if boxes is not None:
            rc = tuple(int(v) for v in boxes[b]) + (int(out_size[0]), int(out_size[1]))
        out.append(transform_sample(batch[b], crop=crop, resized_crop=rc, flip=f, ops=None if ops is None else ops[b],
                                    pre_ops=None if pre_ops is None else pre_ops[b], **kw))
    return torch.stack(out)


def index_map_sample(img: np.ndarray, lut: np.ndarray, *, top: int, left: int, out_h: int, out_w: int,
                     flip: bool, pad_y: int = 0, pad_x: int = 0, pad_mode: str = "constant") -> np.ndarray:
    """Independent second statement of the fused index map: one gather per
    output element, then a per-channel 256-entry table.  Used to cross-check the torchvision
    path and for shapes PIL cannot hold (C=2)."""
    H, W, C = img.shape
    ys = np.arange(out_h)[:, None]
    xs = np.arange(out_w)[None, :]
    if flip:
        xs = out_w - 1 - xs
    u = top + ys - pad_y + 0 * xs
    v = left + xs - pad_x + 0 * ys

    def fold(i, n):
        inside = (i >= 0) & (i < n)
        if pad_mode == "reflect":
            j = np.where(i < 0, -i, np.where(i >= n, 2 * (n - 1) - i, i))
        elif pad_mode == "symmetric":
            j = np.where(i < 0, -i - 1, np.where(i >= n, 2 * n - 1 - i, i))
        elif pad_mode == "edge":
            j = np.clip(i, 0, n - 1)
        else:
            j = np.clip(i, 0, n - 1)
            return j, inside
        return j, np.ones_like(inside)

    uu, iu = fold(u, H)
    vv, iv = fold(v, W)
    px = img[uu, vv, :]
    px = np.where((iu & iv)[..., None], px, 0)
    out = np.empty((C, out_h, out_w), lut.dtype)
    for c in range(C):
        out[c] = lut[c][px[..., c]]
    return out


def to_image(x: torch.Tensor, mean=None, std=None, mode=None) -> np.ndarray:
    """The reference's ``utransform`` (``examples/img_stt/adain/adain.py:168-171``; ``online.py:169-174`` and
    ``offline.py:108-111`` with the min-max stretch), literally: ``T.Compose([T.Normalize(mean, std),
    T.Lambda(clamp | hdr), T.ToPILImage()])`` on one ``[C,H,W]`` float tensor -> ``[H,W,C]`` uint8."""
    import torchvision.transforms as T
    stages = []
    if mean is not None:
        stages.append(T.Normalize(mean=mean, std=std))
    if mode == "clamp":
        stages.append(T.Lambda(lambda t: t.clamp(0, 1)))
    elif mode == "stretch":
        stages.append(T.Lambda(lambda t: (t - t.min()) / (t.max() - t.min())))
    stages.append(T.ToPILImage())
    arr = np.asarray(T.Compose(stages)(x.detach().cpu()))
    return arr.reshape(x.shape[1], x.shape[2], x.shape[0]).copy()
