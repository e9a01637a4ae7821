"""Regenerates the golden fixtures in this directory.  Run from the repo root, in the build
container (needs /root/reference and torchvision/Pillow):

    python tests/golden/make_golden.py

* metrics_reference.npz  -- outputs of the REFERENCE ITSELF: ``torchbooster.metrics.accuracy`` and
  ``RunningAverage`` imported from /root/reference.
* augment_torchvision.npz -- the reference's CIFAR training transform (resnet.py:96-103) run as the
  very ``T.Compose`` it builds, sample after sample under ``torch.manual_seed``, plus every
  RandAugment operation on its own through ``autoaugment._apply_op``.
* transform_torchvision.npz, resize_pillow.npz, params_torchvision.npz -- outputs of the
  third-party code the reference's examples call (torchvision PIL backend / Pillow), which is
  where the transform arithmetic lives.
Library versions are stored in every file.
"""
import sys
from pathlib import Path

import numpy as np
import PIL
import torch
import torchvision
import torchvision.transforms as T
import torchvision.transforms.functional as F
from PIL import Image

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent))
VERSIONS = np.array([f"torch {torch.__version__}", f"torchvision {torchvision.__version__}",
                     f"Pillow {PIL.__version__}", f"numpy {np.__version__}"])

IMAGENET = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
CIFAR = ((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))


def metrics_reference():
    sys.path.insert(0, "/root/reference")
    from torchbooster.metrics import RunningAverage, accuracy  # the reference's own code
    rng = np.random.default_rng(7)
    out = {"versions": VERSIONS}
    nan = float("nan")
    rows = np.array([[1, 3, 3, 0], [2, 2, 2, 2], [nan, 5, 1, 0], [0, 1, nan, nan],
                     [-0.0, 0.0, -1, -2], [0.0, -0.0, -1, -2], [-np.inf, -np.inf, -np.inf, -np.inf],
                     [np.inf, nan, np.inf, 1]], np.float32)
    out["kat_logits"] = rows
    out["kat_argmax"] = torch.from_numpy(rows).argmax(dim=-1).numpy()
    for lab in range(4):
        labels = torch.full((rows.shape[0],), lab, dtype=torch.int64)
        out[f"kat_accuracy_label{lab}"] = accuracy(torch.from_numpy(rows), labels).numpy()
    logits = rng.standard_normal((257, 10)).astype(np.float32)
    labels = rng.integers(0, 10, 257)
    out["a_logits"], out["a_labels"] = logits, labels
    out["a_accuracy"] = accuracy(torch.from_numpy(logits), torch.from_numpy(labels)).numpy()
    # bf16-valued logits: many exact ties
    lt = torch.from_numpy(rng.standard_normal((300, 1000)).astype(np.float32)).to(torch.bfloat16)
    lb = rng.integers(0, 1000, 300)
    out["b_logits_bf16_bits"] = lt.view(torch.int16).numpy()
    out["b_labels"] = lb
    out["b_accuracy"] = accuracy(lt, torch.from_numpy(lb)).float().numpy()
    # >2-D logits: denominator stays size(0)
    l3 = rng.standard_normal((2, 5, 4)).astype(np.float32)
    y3 = torch.from_numpy(l3).argmax(-1)
    out["c_logits"], out["c_labels"] = l3, y3.numpy()
    out["c_accuracy"] = accuracy(torch.from_numpy(l3), y3).numpy()
    # dim=0 variant
    l4 = rng.standard_normal((7, 6)).astype(np.float32)
    y4 = rng.integers(0, 7, 6)
    out["d_logits"], out["d_labels"] = l4, y4
    out["d_accuracy_dim0"] = accuracy(torch.from_numpy(l4), torch.from_numpy(y4), dim=0).numpy()
    seq = rng.standard_normal(64)
    ra = RunningAverage()
    vals = []
    for v in seq:
        ra.update(float(v))
        vals.append(ra.value)
    out["ra_inputs"], out["ra_values"] = seq, np.array(vals, np.float64)
    r2 = RunningAverage()
    for v in (.1, .2, .3):
        r2.update(v)
    out["ra_123"] = np.array([r2.value, r2.current], np.float64)
    np.savez_compressed(HERE / "metrics_reference.npz", **out)


def transform_torchvision():
    rng = np.random.default_rng(11)
    out = {"versions": VERSIONS}
    # (name, shape, kwargs) -- small versions of BASELINE.json's configs
    batch_a = rng.integers(0, 256, (4, 32, 32, 3), dtype=np.uint8)    # C1 CIFAR: reflect pad 4, crop 32, flip
    top = np.array([0, 8, 3, 5]); left = np.array([8, 0, 4, 1]); flip = np.array([1, 0, 1, 0])
    res = []
    for b in range(4):
        im = F.pad(Image.fromarray(batch_a[b]), 4, 0, "reflect")
        im = F.crop(im, int(top[b]), int(left[b]), 32, 32)
        if flip[b]:
            im = F.hflip(im)
        res.append(F.normalize(F.to_tensor(im), CIFAR[0], CIFAR[1]))
    out["c1_in"], out["c1_top"], out["c1_left"], out["c1_flip"] = batch_a, top, left, flip
    out["c1_out"] = torch.stack(res).numpy()

    batch_b = rng.integers(0, 256, (3, 48, 64, 3), dtype=np.uint8)    # C2-like: crop 40x56, flip, bf16
    top = np.array([0, 8, 5]); left = np.array([8, 0, 3]); flip = np.array([0, 1, 1])
    res = []
    for b in range(3):
        im = F.crop(Image.fromarray(batch_b[b]), int(top[b]), int(left[b]), 40, 56)
        if flip[b]:
            im = F.hflip(im)
        res.append(F.normalize(F.to_tensor(im), IMAGENET[0], IMAGENET[1]).to(torch.bfloat16))
    out["c2_in"], out["c2_top"], out["c2_left"], out["c2_flip"] = batch_b, top, left, flip
    out["c2_out_bf16_bits"] = torch.stack(res).view(torch.int16).numpy()

    batch_c = rng.integers(0, 256, (3, 16, 16, 3), dtype=np.uint8)    # C3-like: flip, [-1,1]
    flip = np.array([1, 0, 1])
    res = []
    for b in range(3):
        im = Image.fromarray(batch_c[b])
        if flip[b]:
            im = F.hflip(im)
        res.append(F.normalize(F.to_tensor(im), [.5, .5, .5], [.5, .5, .5]))
    out["c3_in"], out["c3_flip"], out["c3_out"] = batch_c, flip, torch.stack(res).numpy()

    batch_d = rng.integers(0, 256, (2, 28, 28, 1), dtype=np.uint8)    # gan.py: ToTensor then 1 - X, grayscale
    res = [1.0 - F.to_tensor(Image.fromarray(batch_d[b, :, :, 0])) for b in range(2)]
    out["gan_in"], out["gan_out"] = batch_d, torch.stack(res).numpy()

    batch_e = rng.integers(0, 256, (2, 60, 90, 3), dtype=np.uint8)    # C4-like: Resize(32)+CenterCrop(32)+Normalize
    res = [F.normalize(F.to_tensor(F.center_crop(F.resize(Image.fromarray(batch_e[b]), 32), 32)), *IMAGENET)
           for b in range(2)]
    out["c4_in"], out["c4_out"] = batch_e, torch.stack(res).numpy()
    batch_f = rng.integers(0, 256, (3, 40, 48, 3), dtype=np.uint8)    # RandomResizedCrop(24) with explicit boxes + flip
    boxes = np.array([[3, 5, 30, 36], [0, 0, 40, 48], [10, 20, 13, 17]]); flip = np.array([1, 0, 1])
    res = []
    for b in range(3):
        im = F.resized_crop(Image.fromarray(batch_f[b]), *[int(v) for v in boxes[b]], [24, 24])
        if flip[b]:
            im = F.hflip(im)
        res.append(F.normalize(F.to_tensor(im), IMAGENET[0], IMAGENET[1]))
    out["rrc_in"], out["rrc_boxes"], out["rrc_flip"], out["rrc_out"] = batch_f, boxes, flip, torch.stack(res).numpy()
    np.savez_compressed(HERE / "transform_torchvision.npz", **out)


def resize_pillow():
    rng = np.random.default_rng(13)
    out = {"versions": VERSIONS}
    cases = [((37, 53, 3), (20, 31)), ((64, 64, 3), (32, 32)), ((30, 20, 3), (45, 30)), ((50, 41), (17, 41)),
             ((33, 47, 3), (33, 20)), ((128, 96, 3), (7, 5))]
    for i, (shape, (oh, ow)) in enumerate(cases):
        img = rng.integers(0, 256, shape, dtype=np.uint8)
        res = np.asarray(Image.fromarray(img).resize((ow, oh), Image.BILINEAR))
        out[f"r{i}_in"], out[f"r{i}_out"] = img, res
    out["n_cases"] = np.array(len(cases))
    np.savez_compressed(HERE / "resize_pillow.npz", **out)


def params_torchvision():
    out = {"versions": VERSIONS}
    for name, (seed, B, h, w, th, tw, pad) in {"c2": (42, 64, 256, 256, 224, 224, 0), "c1": (42, 64, 32, 32, 32, 32, 4),
                                              "eq": (7, 16, 64, 64, 64, 64, 0), "rect": (123, 33, 48, 64, 48, 50, 0)}.items():
        torch.manual_seed(seed)
        img = Image.new("L", (w + 2 * pad, h + 2 * pad))
        rows = []
        for _ in range(B):
            i, j, _, _ = T.RandomCrop.get_params(img, (th, tw))
            rows.append((i, j, int(bool(torch.rand(1) < 0.5))))
        out[f"{name}_cfg"] = np.array([seed, B, h, w, th, tw, pad])
        out[f"{name}_params"] = np.array(rows, np.int64)
    # T.RandomResizedCrop.get_params + flip draw (examples/img_stt/online/online.py:173)
    for name, (seed, B, h, w, scale, ratio) in {"rrc_c2": (42, 96, 256, 256, (0.08, 1.0), (3 / 4, 4 / 3)),
                                                "rrc_wide": (5, 64, 48, 300, (0.3, 1.0), (0.9, 1.1)),
                                                "rrc_tall": (9, 64, 300, 40, (0.08, 1.0), (3 / 4, 4 / 3))}.items():
        torch.manual_seed(seed)
        img = Image.new("L", (w, h))
        rows = []
        for _ in range(B):
            i, j, hh, ww = T.RandomResizedCrop.get_params(img, list(scale), list(ratio))
            rows.append((i, j, hh, ww, int(bool(torch.rand(1) < 0.5))))
        out[f"{name}_cfg"] = np.array([seed, B, h, w, *scale, *ratio], np.float64)
        out[f"{name}_params"] = np.array(rows, np.int64)
    np.savez_compressed(HERE / "params_torchvision.npz", **out)


def augment_torchvision():
    from torchvision.transforms import autoaugment as AA
    from torchvision.transforms import InterpolationMode
    rng = np.random.default_rng(17)
    out = {"versions": VERSIONS, "cpu_capability": np.array(str(torch.backends.cpu.get_cpu_capability()))}
    # (a) the reference's own Compose (examples/img_cls/resnet/resnet.py:94-103), seed 42, 24 samples in order
    batch = rng.integers(0, 256, (24, 32, 32, 3), dtype=np.uint8)
    batch[3] = (batch[3] // 4) + 90          # low contrast
    batch[5, :, :, 1] = 17                   # one constant band (AutoContrast / Equalize identity case)
    normalize = T.Normalize(mean=CIFAR[0], std=CIFAR[1], inplace=True)
    compose = T.Compose([T.RandomCrop(32, 4, padding_mode="reflect"), T.RandomHorizontalFlip(), T.RandomRotation(15),
                         T.RandAugment(), T.ToTensor(), normalize])
    torch.manual_seed(42)
    res = torch.stack([compose(Image.fromarray(batch[b])) for b in range(batch.shape[0])])
    out["cifar_input"], out["cifar_seed"], out["cifar_output"] = batch, np.array(42), res.numpy()
    # (b) every operation alone, both signs, a few magnitudes, RGB and L
    names, mags, idx, outs = [], [], [], []
    imgs = [rng.integers(0, 256, (20, 28, 3), dtype=np.uint8), rng.integers(30, 200, (19, 19, 1), dtype=np.uint8)]
    space = T.RandAugment()._augmentation_space(31, (20, 28))
    for k, img in enumerate(imgs):
        pil = Image.fromarray(img if img.shape[2] == 3 else img[:, :, 0])
        for name, (table, signed) in space.items():
            for bin_ in (3, 9, 30):
                for sign in ((1.0, -1.0) if signed else (1.0,)):
                    m = float(table[bin_].item()) * sign if table.ndim > 0 else 0.0
                    r = np.asarray(AA._apply_op(pil, name, m, InterpolationMode.NEAREST, None)).reshape(img.shape)
                    names.append(name); mags.append(m); idx.append(k); outs.append(r.ravel())
        for angle in (-15.0, 7.25, 90.0, 180.0, 270.0, 359.5):
            r = np.asarray(F.rotate(pil, angle)).reshape(img.shape)
            names.append("RandomRotation"); mags.append(angle); idx.append(k); outs.append(r.ravel())
    out["op_image0"], out["op_image1"] = imgs
    out["op_names"], out["op_magnitudes"], out["op_image"] = np.array(names), np.array(mags, np.float64), np.array(idx)
    out["op_offsets"] = np.cumsum([0] + [len(o) for o in outs])
    out["op_outputs"] = np.concatenate(outs)
    np.savez_compressed(HERE / "augment_torchvision.npz", **out)


if __name__ == "__main__":
    augment_torchvision()
    metrics_reference()
    transform_torchvision()
    resize_pillow()
    params_torchvision()
    for f in sorted(HERE.glob("*.npz")):
        print(f.name, f.stat().st_size)
