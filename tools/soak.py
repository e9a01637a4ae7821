#!/usr/bin/env python
"""Randomised soak of every CUDA path against the oracle (torchvision / numpy restatements): far more
seeds and shapes than the test-suite holds.  `python tools/soak.py --seconds 120 [--seed 0]`.
Prints one line per failure with everything needed to replay it, and a summary."""
import argparse
import sys
import time
import traceback
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import augment_oracle as AO          # noqa: E402
from oracle import metrics_oracle as MO          # noqa: E402
from oracle import transform_oracle as TO        # noqa: E402
from torchbooster_b200 import augment as A       # noqa: E402
from torchbooster_b200.metrics import MetricState  # noqa: E402
from torchbooster_b200.transforms import BatchTransform, ToImage, TransformSpec  # noqa: E402

DTYPES = [torch.float32, torch.bfloat16, torch.float16]
MODES = ["constant", "reflect", "edge", "symmetric"]
NAMES = list(AO.RAND_AUGMENT_OPS) + ["RandomRotation", "Invert"]


def same(a, b):
    a, b = a.cpu().contiguous(), b.cpu().contiguous()
    view = {torch.float32: torch.int32, torch.bfloat16: torch.int16, torch.float16: torch.int16, torch.uint8: torch.uint8}[a.dtype]
    return a.dtype == b.dtype and a.shape == b.shape and torch.equal(a.view(view), b.view(view))


def images(rng, B, H, W, C):
    kind = rng.integers(0, 4)
    if kind == 0:
        return rng.integers(0, 256, (B, H, W, C), dtype=np.uint8)
    if kind == 1:
        return (rng.integers(0, 256, (B, 1, 1, C)) + rng.integers(-6, 7, (B, H, W, C))).clip(0, 255).astype(np.uint8)
    if kind == 2:
        return np.broadcast_to(rng.integers(0, 256, (B, 1, 1, C), dtype=np.uint8), (B, H, W, C)).copy()
    return (rng.integers(0, 2, (B, H, W, C)) * 255).astype(np.uint8)


def mean_std(C):
    return ((0.485, 0.456, 0.406)[:C], (0.229, 0.224, 0.225)[:C])


def case_transform(rng):
    C = int(rng.choice([1, 3, 3, 3, 4, 2]))
    H, W, B = int(rng.integers(1, 90)), int(rng.integers(1, 90)), int(rng.integers(1, 9))
    pad = int(rng.integers(0, min(H, W, 6)))
    mode = MODES[int(rng.integers(0, 4))]
    if mode == "reflect":
        pad = min(pad, H - 1, W - 1)
    th, tw = int(rng.integers(1, H + 2 * pad + 1)), int(rng.integers(1, W + 2 * pad + 1))
    dtype = DTYPES[int(rng.integers(0, 3))]
    batch = images(rng, B, H, W, C)
    top = rng.integers(0, H + 2 * pad - th + 1, B).astype(np.int32)
    left = rng.integers(0, W + 2 * pad - tw + 1, B).astype(np.int32)
    flip = rng.integers(0, 2, B).astype(np.uint8)
    if C == 2:
        return None         # PIL has no two-channel mode: covered by the index-map tests
    m, s = ((0.5,) * C, (0.25,) * C) if C == 4 else mean_std(C)
    spec = TransformSpec(crop=(th, tw), padding=pad, padding_mode=mode, flip_p=.5, mean=m, std=s, dtype=dtype)
    bt = BatchTransform(spec, device="cuda")
    bt.flags = int(rng.choice([0, 1, 2]))
    got = bt(torch.from_numpy(batch).cuda(), params=(top, left, flip))
    want = TO.transform_batch(batch, crop_top=top, crop_left=left, out_size=(th, tw), flip=flip, padding=pad, padding_mode=mode,
                              mean=m, std=s, dtype=dtype)
    return same(got, want), dict(kind="transform", C=C, H=H, W=W, B=B, pad=pad, mode=mode, th=th, tw=tw, dtype=str(dtype), flags=bt.flags)


def case_resize(rng):
    C = int(rng.choice([1, 3, 3]))
    H, W, B = int(rng.integers(2, 200)), int(rng.integers(2, 200)), int(rng.integers(1, 5))
    if rng.integers(0, 2):
        W = 16 * int(rng.integers(1, 40))            # 16-byte aligned rows: the ring kernel's territory (C == 3)
        B = int(rng.integers(1, 40))
    rs = (int(rng.integers(1, 150)), int(rng.integers(1, 150))) if rng.integers(0, 2) else int(rng.integers(2, 150))
    if rng.integers(0, 4) == 0 and H % 4 == 0 and W % 4 == 0:
        f = int(rng.choice([2, 4]))
        rs = (H // f, W // f)                        # exact 2x / 4x: taps with a single byte plane (their own code path)
    cc = None if rng.integers(0, 3) == 0 else (int(rng.integers(1, 120)), int(rng.integers(1, 120)))
    dtype = DTYPES[int(rng.integers(0, 3))]
    batch = images(rng, B, H, W, C)
    m, s = mean_std(C)
    spec = TransformSpec(resize=rs, center_crop=cc, mean=m, std=s, dtype=dtype)
    bt = BatchTransform(spec, device="cuda")
    bt.flags = int(rng.choice([0, 0, 1, 16, 32, 1 << 8, 2 << 8, 3 << 8, 4]))      # default, generic, two-phase, slide, ring groups, no PDL
    got = bt(torch.from_numpy(batch).cuda())
    want = TO.transform_batch(batch, resize=rs, center_crop=cc, mean=m, std=s, dtype=dtype)
    return same(got, want), dict(kind="resize", C=C, H=H, W=W, B=B, rs=rs, cc=cc, dtype=str(dtype), flags=bt.flags)


def case_rrc(rng):
    C = int(rng.choice([1, 3, 3]))
    H, W, B = int(rng.integers(4, 260)), int(rng.integers(4, 260)), int(rng.integers(1, 7))
    size = (int(rng.integers(1, 120)), int(rng.integers(1, 120)))
    seed = int(rng.integers(0, 2 ** 31))
    dtype = DTYPES[int(rng.integers(0, 3))]
    batch = images(rng, B, H, W, C)
    m, s = mean_std(C)
    from oracle import params_oracle as PO
    boxes, flip = PO.draw_resized_crop_flip(seed, B, H, W)
    spec = TransformSpec(random_resized_crop=size, flip_p=.5, mean=m, std=s, dtype=dtype)
    bt = BatchTransform(spec, device="cuda", seed=seed)
    bt.flags = int(rng.choice([0, 16]))
    got = bt(torch.from_numpy(batch).cuda())
    want = TO.transform_batch(batch, boxes=boxes, out_size=size, flip=flip, mean=m, std=s, dtype=dtype)
    return same(got, want), dict(kind="rrc", C=C, H=H, W=W, B=B, size=size, seed=seed, dtype=str(dtype), flags=bt.flags)


def case_ragged(rng):
    C = int(rng.choice([1, 3, 3]))
    B = int(rng.integers(1, 7))
    sizes = [(int(rng.integers(3, 220)), int(rng.integers(3, 220))) for _ in range(B)]
    imgs = [images(rng, 1, h, w, C)[0] for h, w in sizes]
    m, s = mean_std(C)
    dtype = DTYPES[int(rng.integers(0, 3))]
    if rng.integers(0, 2):
        rs, cc = int(rng.integers(2, 100)), int(rng.integers(1, 100))
        spec = TransformSpec(resize=rs, center_crop=cc, mean=m, std=s, dtype=dtype)
        got = BatchTransform(spec, device="cuda").ragged(imgs)
        want = torch.stack([TO.transform_sample(im, resize=rs, center_crop=cc, mean=m, std=s, dtype=dtype) for im in imgs])
        info = dict(kind="ragged_resize", C=C, sizes=sizes, rs=rs, cc=cc, dtype=str(dtype))
    else:
        import torchvision.transforms as T
        from PIL import Image
        size, seed = int(rng.integers(1, 100)), int(rng.integers(0, 2 ** 31))
        compose = T.Compose([T.RandomResizedCrop(size), T.RandomHorizontalFlip(), T.ToTensor(), T.Normalize(m, s)])
        state = torch.get_rng_state()
        torch.manual_seed(seed)
        want = torch.stack([compose(Image.fromarray(im if C == 3 else im[:, :, 0])) for im in imgs]).to(dtype)
        torch.set_rng_state(state)
        got = BatchTransform(TransformSpec(random_resized_crop=size, flip_p=.5, mean=m, std=s, dtype=dtype), device="cuda",
                             seed=seed).ragged(imgs)
        info = dict(kind="ragged_rrc", C=C, sizes=sizes, size=size, seed=seed, dtype=str(dtype))
    return same(got, want), info


def case_augment(rng):
    C = int(rng.choice([1, 3, 3]))
    H, W, B = int(rng.integers(2, 120)), int(rng.integers(2, 120)), int(rng.integers(1, 10))
    pad = int(rng.integers(0, min(H, W, 5)))
    mode = MODES[int(rng.integers(0, 4))]
    if mode == "reflect":
        pad = min(pad, H - 1, W - 1)
    th, tw = int(rng.integers(1, H + 2 * pad + 1)), int(rng.integers(1, W + 2 * pad + 1))
    dtype = DTYPES[int(rng.integers(0, 3))]
    batch = images(rng, B, H, W, C)
    top = rng.integers(0, H + 2 * pad - th + 1, B).astype(np.int32)
    left = rng.integers(0, W + 2 * pad - tw + 1, B).astype(np.int32)
    flip = rng.integers(0, 2, B).astype(np.uint8)
    ops = []
    for b in range(B):
        mags = A.rand_augment_magnitudes(th, tw, int(rng.integers(0, 31)))
        mine = []
        for _ in range(int(rng.integers(0, 7))):
            name = NAMES[int(rng.integers(0, len(NAMES)))]
            mag = float(np.float32(rng.uniform(-360, 360))) if name == "RandomRotation" else mags.get(name, 0.0)
            if name not in ("Posterize", "Solarize", "RandomRotation") and rng.integers(0, 2):
                mag = -mag
            mine.append((name, mag))
        ops.append(mine)
    m, s = mean_std(C)
    spec = TransformSpec(crop=(th, tw), padding=pad, padding_mode=mode, flip_p=.5, rotation=10, rand_augment=True, mean=m, std=s, dtype=dtype)
    bt = BatchTransform(spec, device="cuda")
    bt.flags = (int(rng.choice([0, 0, 2, 4, 8])) << 8) | (int(rng.choice([0, 2, 3, 4])) << 12)
    info = dict(kind="augment", C=C, H=H, W=W, B=B, pad=pad, mode=mode, th=th, tw=tw, dtype=str(dtype), flags=bt.flags, ops=ops)
    try:
        got = bt(torch.from_numpy(batch).cuda(), params=(top, left, flip, ops))
    except Exception as e:
        return False, dict(info, ops=None, error=repr(e))
    want = TO.transform_batch(batch, crop_top=top, crop_left=left, out_size=(th, tw), flip=flip, ops=ops, padding=pad,
                              padding_mode=mode, mean=m, std=s, dtype=dtype)
    return same(got, want), info


def case_twostage(rng):
    """Resize / CenterCrop, then RandomCrop with any padding mode, flip and (half of the time) operations: the resize kernel
    hands planar bytes to the augment kernel (TransformSpec.two_stage)."""
    from torchbooster_b200.transforms import resized_output_size
    C = int(rng.choice([1, 3, 3]))
    H, W, B = int(rng.integers(8, 160)), int(rng.integers(8, 160)), int(rng.integers(1, 8))
    rs = None if rng.integers(0, 4) == 0 else ((int(rng.integers(6, 120)), int(rng.integers(6, 120))) if rng.integers(0, 2) else int(rng.integers(6, 120)))
    cur_h, cur_w = resized_output_size(H, W, rs) if rs is not None else (H, W)
    cc = None
    if rs is None or rng.integers(0, 2):
        cc = (int(rng.integers(4, cur_h + 12)), int(rng.integers(4, cur_w + 12)))
        cur_h, cur_w = cc
    with_ops = bool(rng.integers(0, 2))
    pad = int(rng.integers(0 if with_ops else 1, 6))
    mode = MODES[int(rng.integers(0, 4))]
    if mode == "reflect":
        pad = min(pad, cur_h - 1, cur_w - 1)
    if mode == "symmetric":
        pad = min(pad, cur_h, cur_w)
    if not with_ops and pad == 0:
        return True, dict(kind="twostage", skipped=True)
    if not with_ops and mode == "constant" and cc is None:
        mode = "edge"                                 # (constant padding after a bare resize is the one-launch route)
    th, tw = int(rng.integers(1, cur_h + 2 * pad + 1)), int(rng.integers(1, cur_w + 2 * pad + 1))
    dtype = DTYPES[int(rng.integers(0, 3))]
    batch = images(rng, B, H, W, C)
    top = rng.integers(0, cur_h + 2 * pad - th + 1, B).astype(np.int32)
    left = rng.integers(0, cur_w + 2 * pad - tw + 1, B).astype(np.int32)
    flip = rng.integers(0, 2, B).astype(np.uint8)
    ops = None
    if with_ops:
        ops = []
        for b in range(B):
            mags = A.rand_augment_magnitudes(th, tw, int(rng.integers(0, 31)))
            mine = []
            for _ in range(int(rng.integers(0, 4))):
                name = NAMES[int(rng.integers(0, len(NAMES)))]
                mag = float(np.float32(rng.uniform(-360, 360))) if name == "RandomRotation" else mags.get(name, 0.0)
                if name not in ("Posterize", "Solarize", "RandomRotation") and rng.integers(0, 2):
                    mag = -mag
                mine.append((name, mag))
            ops.append(mine)
    m, s = mean_std(C)
    spec = TransformSpec(resize=rs, center_crop=cc, crop=(th, tw), padding=pad, padding_mode=mode, flip_p=.5,
                         rotation=10 if with_ops else None, rand_augment=True if with_ops else None, mean=m, std=s, dtype=dtype)
    info = dict(kind="twostage", C=C, H=H, W=W, B=B, rs=rs, cc=cc, pad=pad, mode=mode, th=th, tw=tw, dtype=str(dtype), ops=ops)
    if not spec.two_stage:
        return False, dict(info, error="not two_stage")
    try:
        got = BatchTransform(spec, device="cuda")(torch.from_numpy(batch).cuda(), params=(top, left, flip, ops) if with_ops else (top, left, flip))
    except Exception as e:
        return False, dict(info, ops=None, error=repr(e))
    want = TO.transform_batch(batch, resize=rs, center_crop=cc, crop_top=top, crop_left=left, out_size=(th, tw), flip=flip, ops=ops,
                              padding=pad, padding_mode=mode, mean=m, std=s, dtype=dtype)
    return same(got, want), info


def case_metrics(rng):
    N, C, k = int(rng.integers(1, 600)), int(rng.integers(1, 1300)), int(rng.integers(1, 8))
    if rng.integers(0, 2):
        C = 8 * int(rng.integers(1, 129))            # whole vectors, at most 1024 classes: the one-batch row kernel
    dtype = DTYPES[int(rng.integers(0, 3))]
    g = torch.Generator().manual_seed(int(rng.integers(0, 2 ** 31)))
    logits = torch.randn(N, C, generator=g)
    kind = int(rng.integers(0, 4))
    if kind == 1:
        logits = (logits * 2).round() / 2                      # many exact ties
    elif kind == 2:
        logits[torch.rand(N, C, generator=g) < 0.01] = float("nan")
        logits[torch.rand(N, C, generator=g) < 0.01] = float("inf")
        logits[torch.rand(N, C, generator=g) < 0.01] = -float("inf")
    elif kind == 3:
        logits = torch.zeros(N, C) * torch.where(torch.rand(N, C, generator=g) < 0.5, -1.0, 1.0)     # +0 / -0
    logits = logits.to(dtype)
    labels = torch.randint(0, C, (N,), generator=g)
    if kind == 1:
        labels[::2] = logits[::2].float().argmax(-1)           # labels on (tied) maxima
    st = MetricState(C, k=k, device="cuda")
    st.update(logits.cuda(), labels.cuda())
    res = st.compute()
    ok = (res["correct_1"] == MO.topk_correct(logits.float(), labels, 1) and res[f"correct_{k}"] == MO.topk_correct(logits.float(), labels, k)
          and np.array_equal(res["confusion"].cpu().numpy(), MO.confusion(logits.float(), labels, C)))
    return ok, dict(kind="metrics", N=N, C=C, k=k, dtype=str(dtype), variant=kind)


def case_image(rng):
    C = int(rng.choice([1, 3, 3]))
    H, W = int(rng.integers(1, 150)), int(rng.integers(1, 150))
    mode = [None, "clamp", "stretch"][int(rng.integers(0, 3))]
    g = torch.Generator().manual_seed(int(rng.integers(0, 2 ** 31)))
    m, s = (np.array(v) for v in mean_std(C))
    x = torch.randn((C, H, W), generator=g) * float(rng.uniform(0.1, 3))
    if mode is None:
        x = ((torch.rand((C, H, W), generator=g) * 0.99 + 0.005) - torch.tensor(m, dtype=torch.float32)[:, None, None]) / \
            torch.tensor(s, dtype=torch.float32)[:, None, None]
        x = x.clamp(min=float(((0.002 - m) / s).max()), max=float(((0.997 - m) / s).min()))
    got = ToImage.inverse_of(m, s, mode)(x.cuda())
    want = TO.to_image(x, mean=-m / s, std=1.0 / s, mode=mode)
    return np.array_equal(got.cpu().numpy(), want), dict(kind="to_image", C=C, H=H, W=W, mode=mode)


def case_stream(rng):
    """C == 3 with 16-byte aligned rows: the streaming (bulk-copy) kernel, device and pinned-host sources,
    strided batches, caller-provided outputs."""
    B = int(rng.integers(1, 40))
    H, W = int(rng.integers(8, 150)), 16 * int(rng.integers(1, 12))
    pad = int(rng.choice([0, 0, 0, 2, 4]))
    mode = MODES[int(rng.integers(0, 4))]
    dtype = DTYPES[int(rng.integers(0, 3))]
    px = 4 if dtype == torch.float32 else 8
    tw = px * int(rng.integers(1, (W + 2 * pad) // px + 1))
    th = int(rng.integers(1, H + 2 * pad + 1))
    batch = images(rng, B, H, W, 3)
    top = rng.integers(0, H + 2 * pad - th + 1, B).astype(np.int32)
    left = rng.integers(0, W + 2 * pad - tw + 1, B).astype(np.int32)
    flip = rng.integers(0, 2, B).astype(np.uint8)
    m, s = mean_std(3)
    one_minus = bool(rng.integers(0, 4) == 0)
    spec = TransformSpec(crop=(th, tw), padding=pad, padding_mode=mode, flip_p=.5, mean=m, std=s, one_minus=one_minus, dtype=dtype)
    bt = BatchTransform(spec, device="cuda")
    bt.flags = int(rng.choice([0, 2, 2, 6]))
    src_kind = int(rng.integers(0, 3))
    if src_kind == 0:
        src = torch.from_numpy(batch).cuda()
    elif src_kind == 1:
        src = torch.from_numpy(batch).pin_memory()                  # read in place over PCIe
    else:
        wide = torch.zeros((B, H + 3, W + 16, 3), dtype=torch.uint8, device="cuda")
        wide[:, :H, :W] = torch.from_numpy(batch).cuda()
        src = wide[:, :H, :W]                                       # batch stride and row pitch larger than dense
    out = torch.empty((B, 3, th, tw), dtype=dtype, device="cuda") if rng.integers(0, 2) else None
    info = dict(kind="stream", B=B, H=H, W=W, pad=pad, mode=mode, th=th, tw=tw, dtype=str(dtype), flags=bt.flags, src=src_kind,
                one_minus=one_minus)
    try:
        got = bt(src, params=(top, left, flip), out=out)
    except Exception as e:
        return False, dict(info, error=repr(e))
    want = TO.transform_batch(batch, crop_top=top, crop_left=left, out_size=(th, tw), flip=flip, padding=pad, padding_mode=mode,
                              mean=m, std=s, one_minus=one_minus, dtype=dtype)
    return same(got, want), info


def case_chain(rng):
    """pre_rotation -> RandomResizedCrop -> flip -> rotation -> RandAugment (online.py:173 and variants), seeded,
    against the literal Compose; and large images through the cluster path of the augment kernel."""
    import torchvision.transforms as T
    from PIL import Image
    C = 3
    if rng.integers(0, 2):
        B, H, W, size = int(rng.integers(1, 5)), 4 * int(rng.integers(10, 60)), 4 * int(rng.integers(10, 60)), int(rng.integers(8, 120))
        pre = None if rng.integers(0, 2) else float(rng.choice([15, 90, 180]))
        flip_p = None if rng.integers(0, 2) else 0.5
        rot = None if rng.integers(0, 2) else 20.0
        ra = bool(rng.integers(0, 2))
        if pre is None and rot is None and not ra:
            ra = True
        seed = int(rng.integers(0, 2 ** 31))
        batch = images(rng, B, H, W, C)
        m, s = mean_std(C)
        stages = ([T.RandomRotation(pre)] if pre is not None else []) + [T.RandomResizedCrop(size)] + \
            ([T.RandomHorizontalFlip(flip_p)] if flip_p is not None else []) + ([T.RandomRotation(rot)] if rot is not None else []) + \
            ([T.RandAugment()] if ra else []) + [T.ToTensor(), T.Normalize(m, s)]
        compose = T.Compose(stages)
        state = torch.get_rng_state()
        torch.manual_seed(seed)
        want = torch.stack([compose(Image.fromarray(batch[b])) for b in range(B)])
        torch.set_rng_state(state)
        spec = TransformSpec(random_resized_crop=size, pre_rotation=pre, flip_p=flip_p, rotation=rot, rand_augment=ra, mean=m, std=s)
        info = dict(kind="chain", B=B, H=H, W=W, size=size, pre=pre, flip_p=flip_p, rot=rot, ra=ra, seed=seed)
        try:
            got = BatchTransform(spec, device="cuda", seed=seed)(torch.from_numpy(batch).cuda())
        except Exception as e:
            return False, dict(info, error=repr(e))
        return same(got, want), info
    B, H, W = int(rng.integers(1, 4)), int(rng.integers(100, 420)), int(rng.integers(100, 420))
    batch = images(rng, B, H, W, C)
    m, s = mean_std(C)
    ops = []
    for b in range(B):
        mags = A.rand_augment_magnitudes(H, W, int(rng.integers(0, 31)))
        mine = []
        for _ in range(int(rng.integers(1, 5))):
            name = NAMES[int(rng.integers(0, len(NAMES)))]
            mag = float(np.float32(rng.uniform(-360, 360))) if name == "RandomRotation" else mags.get(name, 0.0)
            if name not in ("Posterize", "Solarize", "RandomRotation") and rng.integers(0, 2):
                mag = -mag
            mine.append((name, mag))
        ops.append(mine)
    spec = TransformSpec(rand_augment=True, mean=m, std=s)
    info = dict(kind="augment_large", B=B, H=H, W=W, ops=ops)
    try:
        got = BatchTransform(spec, device="cuda")(torch.from_numpy(batch).cuda(), params=(None, None, None, ops))
    except Exception as e:
        return False, dict(info, ops=None, error=repr(e))
    return same(got, TO.transform_batch(batch, ops=ops, mean=m, std=s)), info


def case_loader(rng):
    """LMDB file -> BatchLoader (gathered into pinned slots, or the registered map read in place) -> transform,
    ragged last batch, labels; against the oracle on the same sampler order."""
    import tempfile
    from oracle import params_oracle as PO
    from torchbooster_b200.lmdb import LMDBReader, write_image_lmdb
    from torchbooster_b200.loader import BatchLoader
    n, bs = int(rng.integers(1, 120)), int(rng.integers(1, 40))
    H, W = int(rng.integers(4, 70)), 16 * int(rng.integers(1, 6)) if rng.integers(0, 2) else int(rng.integers(4, 70))
    th, tw = int(rng.integers(1, H + 1)), int(rng.integers(1, W + 1))
    in_place = bool(rng.integers(0, 2))
    seed = int(rng.integers(0, 2 ** 31))
    imgs = images(rng, n, H, W, 3)
    order = rng.permutation(n)
    m, s = mean_std(3)
    dtype = DTYPES[int(rng.integers(0, 3))]
    info = dict(kind="loader", n=n, bs=bs, H=H, W=W, th=th, tw=tw, in_place=in_place, seed=seed, dtype=str(dtype))
    with tempfile.TemporaryDirectory() as tmp:
        write_image_lmdb(tmp + "/db", imgs, labels=np.arange(n) % 7)
        reader = LMDBReader(tmp + "/db")
        spec = TransformSpec(crop=(th, tw), flip_p=.5, mean=m, std=s, dtype=dtype)
        try:
            loader = BatchLoader(reader, (H, W, 3), bs, sampler=order.tolist(), labels=reader.read_labels(),
                                 transform=BatchTransform(spec, device="cuda", seed=seed), in_place=in_place, threads=3)
            outs = [(x.clone(), y.clone()) for x, y in loader]
            torch.cuda.synchronize()
        except NotImplementedError as e:
            if in_place and "registration" in str(e):
                return None
            return False, dict(info, error=repr(e))
        except Exception as e:
            return False, dict(info, error=repr(e))
        finally:
            reader.close()
    top, left, flip = PO.draw_crop_flip(seed, n, H, W, th, tw, 0)
    want = TO.transform_batch(imgs[order], crop_top=top, crop_left=left, out_size=(th, tw), flip=flip, mean=m, std=s, dtype=dtype)
    ok = same(torch.cat([x for x, _ in outs]), want) and torch.cat([y for _, y in outs]).cpu().tolist() == [int(i) % 7 for i in order]
    return ok, info


CASES = [case_transform, case_resize, case_rrc, case_ragged, case_augment, case_twostage, case_metrics, case_image, case_stream, case_chain, case_loader]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=60)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--only", default="", help="comma separated case names (resize, metrics, ...)")
    a = ap.parse_args()
    cases = [c for c in CASES if not a.only or c.__name__[5:] in a.only.split(",")]
    rng = np.random.default_rng(a.seed)
    t0 = time.time()
    done, fails, classes = {}, 0, {}
    i = 0
    while time.time() - t0 < a.seconds:
        fn = cases[i % len(cases)]
        sub = int(rng.integers(0, 2 ** 31))
        i += 1
        try:
            r = fn(np.random.default_rng(sub))
            if r is None:
                continue
            ok, info = r
        except Exception as e:      # an exception is a finding too
            ok, info = False, dict(kind=fn.__name__, error=repr(e), trace=traceback.format_exc()[-600:])
        done[fn.__name__] = done.get(fn.__name__, 0) + 1
        if not ok:
            fails += 1
            key = (info.get("kind"), str(info.get("error", "mismatch"))[:60])
            classes[key] = classes.get(key, 0) + 1
            if classes[key] <= 3:
                print("FAIL", dict(sub_seed=sub, **info), flush=True)
    torch.cuda.synchronize()
    for key, n in sorted(classes.items(), key=lambda kv: -kv[1]):
        print("CLASS", n, key, flush=True)
    print("soak:", done, "failures:", fails, flush=True)
    return 1 if fails else 0


if __name__ == "__main__":
    sys.exit(main())
