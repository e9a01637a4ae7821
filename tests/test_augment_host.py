"""RandomRotation / RandAugment, host side (no GPU): the numpy restatement of Pillow's arithmetic
against torchvision (golden fixture + live), the replay of torchvision's draws, and the integer
encoding of the operations."""
import numpy as np
import pytest
import torch

from oracle import augment_oracle as AO
from oracle import transform_oracle as TO
from torchbooster_b200 import _torchrng
from torchbooster_b200 import augment as A

CIFAR = ((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))


def test_restatement_equals_golden_operations(golden):
    g = golden("augment_torchvision.npz")
    imgs = [g["op_image0"], g["op_image1"]]
    off = g["op_offsets"]
    seen = set()
    for i, (name, m, k) in enumerate(zip(g["op_names"], g["op_magnitudes"], g["op_image"])):
        want = g["op_outputs"][off[i]:off[i + 1]].reshape(imgs[k].shape)
        got = AO.restated_op(imgs[k], str(name), float(m))
        assert np.array_equal(got, want), (name, m, k)
        seen.add(str(name))
    assert seen == set(AO.RAND_AUGMENT_OPS) | {"RandomRotation"}


@pytest.mark.parametrize("seed", range(3))
def test_restatement_equals_torchvision_live(seed):
    rng = np.random.default_rng(100 + seed)
    names = list(AO.RAND_AUGMENT_OPS) + ["RandomRotation", "Invert"]
    for t in range(120):
        C = 1 if t % 5 == 0 else 3
        H, W = (int(v) for v in rng.integers(1, 40, 2))
        img = rng.integers(0, 256, (H, W, C), dtype=np.uint8)
        if t % 3 == 1:
            img = (img // 3 + 60).astype(np.uint8)
        mags = A.rand_augment_magnitudes(H, W, int(rng.integers(0, 31)))
        name = names[t % len(names)]
        m = float(rng.uniform(-180, 180)) if name == "RandomRotation" else mags.get(name, 0.0) * (1 if rng.integers(0, 2) else -1)
        if name in ("Posterize", "Solarize"):
            m = abs(m)
        assert np.array_equal(AO.restated_op(img, name, m), AO.apply_ops(img, [(name, m)])), (name, m, H, W, C)


@pytest.mark.parametrize("seed,B,h,w,crop,pad,p,deg,ra", [
    (42, 96, 32, 32, (32, 32), 4, 0.5, 15, (2, 9, 31)),             # resnet.py:96-100
    (7, 64, 40, 36, (32, 28), 0, 0.3, (-30, 10), (3, 5, 31)),
    (9, 50, 32, 32, (32, 32), 0, 0.5, None, (2, 9, 31)),            # no rotation, nothing drawn for the crop
    (3, 50, 32, 32, (32, 32), 4, 0.5, 180, None),                   # rotation only
    (11, 40, 28, 28, (24, 24), 2, 0.5, 10.5, (1, 30, 31)),
])
def test_replay_equals_torchvisions_draws(seed, B, h, w, crop, pad, p, deg, ra):
    top, left, flip, ops = AO.draw_augment(seed, B, h, w, crop[0], crop[1], pad, p, deg, *(ra or (2, 9, 31)),
                                           rand_augment=ra is not None)
    ap = A.AugmentParams(seed, deg, ra)
    first = ap.draw(B // 3, h, w, crop, (pad, pad), p)              # two calls: the word stream carries over
    second = ap.draw(B - B // 3, h, w, crop, (pad, pad), p)
    for k, want in enumerate((top, left, flip)):
        assert np.array_equal(np.concatenate([first[k], second[k]]), want)
    assert first[3] + second[3] == ops


@pytest.mark.parametrize("seed,B,h,w,crop,pad,p,deg,ra,C", [
    (42, 300, 32, 32, (32, 32), 4, 0.5, 15, (2, 9, 31), 3), (7, 200, 40, 36, (32, 28), 0, 0.3, (-30, 10), (3, 5, 31), 1),
    (3, 100, 32, 32, (32, 32), 4, None, 180, None, 3), (5, 100, 24, 24, None, 0, 0.5, None, (2, 30, 31), 3),
    (6, 64, 20, 20, (20, 20), 0, 0.5, (90, 90), (1, 9, 31), 3), (8, 64, 20, 30, (20, 30), 0, 0.5, (270, 270), None, 3)])
def test_native_walk_equals_python_statement(seed, B, h, w, crop, pad, p, deg, ra, C):
    """tb_augment_params_walk (C++) == AugmentParams.draw + encode_ops (the readable statement),
    records included: libm cos / sin, the 15-decimal rounding and the fixed-point conversion agree."""
    a, b = A.AugmentParams(seed, deg, ra), A.AugmentParams(seed, deg, ra)
    for part in (B // 2, B - B // 2):
        top, left, flip, ops = a.draw(part, h, w, crop, (pad, pad), p)
        t2, l2, f2, rec = b.draw_records(part, h, w, C, crop, (pad, pad), p)
        oh, ow = crop if crop is not None else (h + 2 * pad, w + 2 * pad)
        assert np.array_equal(top, t2) and np.array_equal(left, l2) and np.array_equal(flip, f2)
        assert np.array_equal(rec, A.encode_ops(ops, oh, ow, C))


def test_native_rotation_records_equal_python():
    from torchbooster_b200 import _native as N
    rng = np.random.default_rng(0)
    angles = np.concatenate([rng.uniform(-720, 720, 60000), np.float32(rng.uniform(-15, 15, 60000)).astype(np.float64),
                             rng.uniform(-1e-6, 1e-6, 5000), 90.0 + rng.uniform(-1e-9, 1e-9, 5000),
                             [0.0, 90.0, 180.0, 270.0, 360.0, -90.0, -180.0, 1e-12, 89.99999999999999, 45.0, 135.0, 225.0]])
    for (H, W) in ((32, 32), (17, 40)):
        rec = np.zeros((angles.shape[0], 8), np.int32)
        N.check(N.lib().tb_augment_rotation_records(angles.ctypes.data, angles.shape[0], H, W, rec.ctypes.data))
        want = np.array([A.encode_op("RandomRotation", float(v), H, W, 3) for v in angles], np.int64).astype(np.int32)
        assert np.array_equal(rec, want)


def test_uniform_replay_both_kernels():
    """torch's uniform_ contracts x * (b - a) + a into one FMA in its AVX2 / AVX512 kernels and not in
    the DEFAULT one; the replay follows the running process and is exact for it."""
    words = np.random.RandomState(5).randint(0, 2 ** 32, 4000, dtype=np.uint64)
    for a, b in ((-15.0, 15.0), (0.08, 1.0), (-180.0, 180.0), (0.1, 0.3)):
        torch.manual_seed(5)
        want = np.array([torch.empty(1).uniform_(a, b).item() for _ in range(400)], np.float32)
        assert np.array_equal(_torchrng.uniform(words[:400], a, b), want)
        fused, plain = _torchrng.uniform(words, a, b, fused=True), _torchrng.uniform(words, a, b, fused=False)
        assert np.abs(fused.astype(np.float64) - plain).max() <= 2.0 ** -22 * (abs(a) + abs(b))    # the product's rounding
        x = (words & np.uint64(0xFFFFFF)).astype(np.float64) * 2.0 ** -24
        exact = x * float(np.float32(np.float32(b) - np.float32(a))) + float(np.float32(a))
        assert np.abs(fused - exact).max() <= np.abs(plain - exact).max()
        assert (fused != exact.astype(np.float32)).mean() < 1e-3         # only the float64 detour's double rounding


def _apply_affine_record(img, rec):
    H, W, _ = img.shape
    a0, a1, a2, a3, a4, a5 = (int(v) for v in rec[1:7])
    y, x = np.mgrid[0:H, 0:W].astype(np.int64)
    xin, yin = (a2 + x * a0 + y * a1) >> 16, (a5 + x * a3 + y * a4) >> 16
    ok = (xin >= 0) & (xin < W) & (yin >= 0) & (yin < H)
    out = np.zeros_like(img)
    out[ok] = img[yin[ok], xin[ok]]
    return out


def test_affine_records():
    rng = np.random.default_rng(3)
    for H, W in ((9, 9), (7, 12), (32, 32), (1, 5)):
        img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        for name, m in [("RandomRotation", 90.0), ("RandomRotation", 180.0), ("RandomRotation", 270.0), ("Rotate", -9.0),
                        ("RandomRotation", 450.0), ("ShearX", 0.27), ("ShearY", -0.09), ("TranslateX", -4.35), ("TranslateY", 3.0),
                        ("RandomRotation", 14.999999046325684)]:
            rec = A.encode_op(name, m, H, W, 3)
            assert rec[0] == A.AFFINE and len(rec) == A.OP_WORDS
            assert all(-2 ** 31 <= v < 2 ** 31 for v in rec)
            assert np.array_equal(_apply_affine_record(img, rec), AO.apply_ops(img, [(name, m)])), (name, m, H, W)
    assert A.encode_op("RandomRotation", 360.0, 8, 8, 3)[0] == A.IDENTITY
    assert A.encode_op("Color", 0.3, 8, 8, 1)[0] == A.IDENTITY
    assert A.encode_op("Solarize", 178.5, 8, 8, 3)[:2] == [A.SOLARIZE, 179]
    assert A.encode_op("Posterize", 5, 8, 8, 3)[:2] == [A.POSTERIZE, 0xF8]
    with pytest.raises(ValueError, match="not recognized"):
        A.encode_op("Cutout", 1.0, 8, 8, 3)
    recs = A.encode_ops([[("Rotate", 9.0)], [("Invert", 0.0), ("Equalize", 0.0)], []], 8, 8, 3)
    assert recs.shape == (3, 2, 8) and recs.dtype == np.int32 and recs[0, 1, 0] == A.IDENTITY and recs[2].sum() == 0


def test_reference_cifar_compose_golden(golden):
    """resnet.py:96-103 as one T.Compose under manual_seed(42) (the fixture) == replayed draws +
    per-sample oracle chain."""
    g = golden("augment_torchvision.npz")
    if (str(g["cpu_capability"]).upper() == "DEFAULT") != (not _torchrng.fused_multiply_add()):
        pytest.skip("fixture was drawn by the other variant of torch's uniform_ kernel")
    batch = g["cifar_input"]
    top, left, flip, ops = A.AugmentParams(int(g["cifar_seed"]), 15, (2, 9, 31)).draw(batch.shape[0], 32, 32, (32, 32), (4, 4), 0.5)
    got = TO.transform_batch(batch, crop_top=top, crop_left=left, out_size=(32, 32), flip=flip, ops=ops, padding=4,
                             padding_mode="reflect", mean=CIFAR[0], std=CIFAR[1])
    assert torch.equal(got, torch.from_numpy(g["cifar_output"]))
    restated = np.stack([AO.restated_ops(AO.apply_ops(batch[b], []), []) for b in range(2)])     # smoke of the empty list
    assert np.array_equal(restated, batch[:2])


IMAGENET = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))


@pytest.mark.parametrize("pre,flip_p,rot,ra", [(180, None, None, (2, 9, 31)),      # online.py:173
                                                (None, 0.5, None, (2, 9, 31)), (30, 0.5, 15, None), (None, None, 10, (1, 20, 31))])
def test_chain_replay_equals_the_compose(pre, flip_p, rot, ra):
    """ChainParams against the literal T.Compose around T.RandomResizedCrop under torch.manual_seed:
    same draws <=> same images (both sides end in torchvision's functional ops)."""
    import torchvision.transforms as T
    from PIL import Image
    rng = np.random.default_rng(21)
    batch = rng.integers(0, 256, (10, 60, 80, 3), dtype=np.uint8)
    stages = ([T.RandomRotation(pre)] if pre is not None else []) + [T.RandomResizedCrop(32)] + \
        ([T.RandomHorizontalFlip(flip_p)] if flip_p is not None else []) + ([T.RandomRotation(rot)] if rot is not None else []) + \
        ([T.RandAugment(*ra)] if ra is not None else []) + [T.ToTensor(), T.Normalize(IMAGENET[0], IMAGENET[1])]
    compose = T.Compose(stages)
    state = torch.get_rng_state()
    try:
        torch.manual_seed(314)
        want = torch.stack([compose(Image.fromarray(batch[b])) for b in range(10)])
    finally:
        torch.set_rng_state(state)
    cp = A.ChainParams(314, pre_rotation=pre, rotation=rot, rand_augment=ra)
    box, flip, ops, pre_ops = cp.draw(6, 60, 80, 32, 32, flip_p)
    box2, flip2, ops2, pre2 = cp.draw(4, 60, 80, 32, 32, flip_p)
    join = lambda a, b: None if a is None else a + b
    got = TO.transform_batch(batch, boxes=np.concatenate([box, box2]), out_size=(32, 32), flip=np.concatenate([flip, flip2]),
                             ops=join(ops, ops2), pre_ops=join(pre_ops, pre2), mean=IMAGENET[0], std=IMAGENET[1])
    assert torch.equal(got, want)
