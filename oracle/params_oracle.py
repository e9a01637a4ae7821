"""torchvision's own random crop / flip draws.  TEST INFRASTRUCTURE ONLY.

Per sample the reference pipeline (``/root/reference/examples/img_cls/resnet/resnet.py:97-98``)
draws, in this order, from torch's global CPU generator:
``T.RandomCrop.get_params`` -> ``torch.randint(0, h - th + 1, (1,))``, ``torch.randint(0, w - tw + 1, (1,))``
(``torchvision/transforms/transforms.py:634-655``; nothing is drawn when the sizes are equal) and
``T.RandomHorizontalFlip.forward`` -> ``torch.rand(1) < p`` (``:712-722``).
"""
from __future__ import annotations

import numpy as np
import torch
import torchvision.transforms as T
from PIL import Image


def draw_crop_flip(seed: int, batch: int, h: int, w: int, th: int, tw: int, padding: int = 0, p: float = 0.5):
    """Literal per-sample torchvision draws under ``torch.manual_seed(seed)``.

    ``h, w`` are the unpadded image sizes.  Returns ``(top, left, flip)`` numpy arrays; ``top``
    and ``left`` are in padded coordinates, exactly what ``F.crop`` receives.
    """
    state = torch.get_rng_state()
    try:
        torch.manual_seed(seed)
        img = Image.new("L", (w + 2 * padding, h + 2 * padding))
        top = np.zeros(batch, np.int32)
        left = np.zeros(batch, np.int32)
        flip = np.zeros(batch, np.uint8)
        for b in range(batch):
            i, j, _, _ = T.RandomCrop.get_params(img, (th, tw))
            top[b], left[b] = i, j
            flip[b] = 1 if bool(torch.rand(1) < p) else 0
    finally:
        torch.set_rng_state(state)
    return top, left, flip


def draw_resized_crop_flip(seed: int, batch: int, h: int, w: int, scale=(0.08, 1.0), ratio=(3 / 4, 4 / 3), p: float = 0.5):
    """Literal per-sample ``T.RandomResizedCrop.get_params`` then ``torch.rand(1) < p`` draws
    (``examples/img_stt/online/online.py:173`` builds ``T.RandomResizedCrop(conf.size)``)."""
    state = torch.get_rng_state()
    try:
        torch.manual_seed(seed)
        img = Image.new("L", (w, h))
        box = np.zeros((batch, 4), np.int32)
        flip = np.zeros(batch, np.uint8)
        for b in range(batch):
            box[b] = T.RandomResizedCrop.get_params(img, list(scale), list(ratio))
            flip[b] = 1 if bool(torch.rand(1) < p) else 0
    finally:
        torch.set_rng_state(state)
    return box, flip
