"""Parity of the fused RandomRotation / RandAugment kernel against torchvision (PIL backend),
through the C ABI: every operation reproduces Pillow's bytes, so after the exact table the
outputs must be BIT-EXACT."""
import numpy as np
import pytest
import torch

from oracle import augment_oracle as AO
from oracle import transform_oracle as TO
from torchbooster_b200 import _native as N
from torchbooster_b200 import _torchrng
from torchbooster_b200 import augment as A
from torchbooster_b200.transforms import BatchTransform, TransformSpec

pytestmark = pytest.mark.gpu

CIFAR = ((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))


def same(a: torch.Tensor, b: torch.Tensor) -> bool:
    a, b = a.cpu(), b.cpu()
    if a.dtype != b.dtype or a.shape != b.shape:
        return False
    view = {torch.float32: torch.int32, torch.bfloat16: torch.int16, torch.float16: torch.int16}[a.dtype]
    return torch.equal(a.contiguous().view(view), b.contiguous().view(view))


def to_bytes(out: torch.Tensor) -> np.ndarray:
    """ToTensor-only output [B,C,H,W] -> uint8 [B,H,W,C] (v / 255 is injective on bytes)."""
    return (out.float().cpu() * 255.0).round().to(torch.uint8).permute(0, 2, 3, 1).numpy()


def test_reference_cifar_compose_golden(golden):
    """The reference's training transform (resnet.py:96-103), run as its own T.Compose under
    torch.manual_seed(42) when the fixture was made, against ONE launch with replayed draws."""
    g = golden("augment_torchvision.npz")
    if (str(g["cpu_capability"]).upper() == "DEFAULT") != (not _torchrng.fused_multiply_add()):
        pytest.skip("fixture was drawn by the other variant of torch's uniform_ kernel")
    spec = TransformSpec(crop=32, padding=4, padding_mode="reflect", flip_p=0.5, rotation=15, rand_augment=True,
                         mean=CIFAR[0], std=CIFAR[1])
    bt = BatchTransform(spec, device="cuda", seed=int(g["cifar_seed"]))
    bt.lut(3)                                       # the table is built once per spec, not per batch
    before = N.lib().tb_launch_count()
    out = bt(torch.from_numpy(g["cifar_input"]).cuda())
    torch.cuda.synchronize()
    assert N.lib().tb_launch_count() - before == 1
    assert same(out, torch.from_numpy(g["cifar_output"]))


def test_every_operation_alone_golden(golden):
    g = golden("augment_torchvision.npz")
    imgs = [g["op_image0"], g["op_image1"]]
    off = g["op_offsets"]
    for k, img in enumerate(imgs):
        H, W, C = img.shape
        rows = [i for i in range(len(g["op_names"])) if g["op_image"][i] == k]
        ops = [[(str(g["op_names"][i]), float(g["op_magnitudes"][i]))] for i in rows]
        batch = torch.from_numpy(np.repeat(img[None], len(rows), 0)).cuda()
        bt = BatchTransform(TransformSpec(rand_augment=(1, 9, 31)), device="cuda")
        out = to_bytes(bt(batch, params=(None, None, None, ops)))
        for j, i in enumerate(rows):
            want = g["op_outputs"][off[i]:off[i + 1]].reshape(img.shape)
            assert np.array_equal(out[j], want), (ops[j], k)


@pytest.mark.parametrize("seed", range(6))
def test_fuzz_chains_vs_torchvision(seed):
    """Random sizes (odd plane sizes take the scalar store path), channels, paddings, dtypes and
    operation lists of 0-5 operations over the whole vocabulary, with every magnitude bin."""
    rng = np.random.default_rng(500 + seed)
    names = list(AO.RAND_AUGMENT_OPS) + ["RandomRotation", "Invert"]
    for trial in range(6):
        C = 1 if rng.integers(0, 4) == 0 else 3
        H, W = (int(v) for v in rng.integers(4, 44, 2))
        pad = int(rng.integers(0, 4))
        mode = ["constant", "reflect", "edge", "symmetric"][int(rng.integers(0, 4))]
        th, tw = int(rng.integers(3, H + 2 * pad + 1)), int(rng.integers(3, W + 2 * pad + 1))
        dtype = [torch.float32, torch.bfloat16, torch.float16][int(rng.integers(0, 3))]
        B = int(rng.integers(1, 40))
        batch = rng.integers(0, 256, (B, H, W, C), dtype=np.uint8)
        if trial % 2:
            batch = (batch // int(rng.integers(2, 6)) + int(rng.integers(0, 100))).astype(np.uint8)
        top = rng.integers(0, H + 2 * pad - th + 1, B).astype(np.int32)
        left = rng.integers(0, W + 2 * pad - tw + 1, B).astype(np.int32)
        flip = rng.integers(0, 2, B).astype(np.uint8)
        ops = []
        for b in range(B):
            mags = A.rand_augment_magnitudes(th, tw, int(rng.integers(0, 31)))
            mine = []
            for _ in range(int(rng.integers(0, 6))):
                name = names[int(rng.integers(0, len(names)))]
                m = float(np.float32(rng.uniform(-180, 180))) if name == "RandomRotation" else mags.get(name, 0.0)
                if name not in ("Posterize", "Solarize", "RandomRotation") and rng.integers(0, 2):
                    m = -m
                mine.append((name, m))
            ops.append(mine)
        spec = TransformSpec(crop=(th, tw), padding=pad, padding_mode=mode, flip_p=0.5, rotation=180, rand_augment=True,
                             mean=CIFAR[0][:C], std=CIFAR[1][:C], dtype=dtype)
        got = BatchTransform(spec, device="cuda")(torch.from_numpy(batch).cuda(), params=(top, left, flip, ops))
        want = TO.transform_batch(batch, crop_top=top, crop_left=left, out_size=(th, tw), flip=flip, ops=ops, padding=pad,
                                  padding_mode=mode, mean=CIFAR[0][:C], std=CIFAR[1][:C], dtype=dtype)
        assert same(got, want), (seed, trial, H, W, C, th, tw, pad, mode, dtype)


def test_full_cifar_batches_replayed_vs_torchvision():
    """BASELINE config C1's batch (256 x 32x32x3) through the whole training transform, three
    consecutive batches from one seed; the oracle draws with torchvision's own code."""
    rng = np.random.default_rng(9)
    spec = TransformSpec(crop=32, padding=4, padding_mode="reflect", flip_p=0.5, rotation=15, rand_augment=True,
                         mean=CIFAR[0], std=CIFAR[1], dtype=torch.bfloat16)
    bt = BatchTransform(spec, device="cuda", seed=1234)
    top, left, flip, ops = AO.draw_augment(1234, 3 * 256, 32, 32, 32, 32, 4, 0.5, 15)
    for step in range(3):
        batch = rng.integers(0, 256, (256, 32, 32, 3), dtype=np.uint8)
        got = bt(torch.from_numpy(batch).cuda())
        sl = slice(step * 256, (step + 1) * 256)
        want = TO.transform_batch(batch, crop_top=top[sl], crop_left=left[sl], out_size=(32, 32), flip=flip[sl], ops=ops[sl],
                                  padding=4, padding_mode="reflect", mean=CIFAR[0], std=CIFAR[1], dtype=torch.bfloat16)
        assert same(got, want), step


ALL_OPS = [("Rotate", 21.0), ("Sharpness", -0.63), ("Equalize", 0.0), ("ShearX", -0.21), ("Contrast", 0.45), ("AutoContrast", 0.0),
           ("TranslateY", 31.0), ("Color", -0.9), ("Solarize", 102.0), ("ShearY", 0.3), ("Brightness", 0.27), ("Posterize", 5.0),
           ("TranslateX", -17.0), ("Invert", 0.0), ("RandomRotation", -163.5), ("Identity", 0.0)]


@pytest.mark.parametrize("H,W,C,crop,dtype", [(256, 256, 3, 224, torch.bfloat16),       # ImageNet crop: cluster of 2
                                              (256, 256, 3, None, torch.float32),      # online.py's size: cluster of 2
                                              (300, 333, 1, (299, 331), torch.float32),  # grey, odd sizes: cluster of 1
                                              (512, 512, 3, None, torch.float16)])     # cluster of 8
def test_images_owned_by_a_cluster(H, W, C, crop, dtype):
    """Images that do not fit in one SM's shared memory twice are spread over a thread-block cluster
    (distributed shared memory); every operation, chained, against torchvision."""
    rng = np.random.default_rng(H + W)
    B = 5
    batch = rng.integers(0, 256, (B, H, W, C), dtype=np.uint8)
    batch[1] = (batch[1] // 3 + 50).astype(np.uint8)
    th, tw = (H, W) if crop is None else ((crop, crop) if np.isscalar(crop) else crop)
    top = rng.integers(0, H - th + 1, B).astype(np.int32)
    left = rng.integers(0, W - tw + 1, B).astype(np.int32)
    flip = rng.integers(0, 2, B).astype(np.uint8)
    ops = [[ALL_OPS[(3 * b + j) % len(ALL_OPS)] for j in range(4)] for b in range(B)]
    ops[4] = ALL_OPS[12:16] + ALL_OPS[0:3]
    spec = TransformSpec(crop=None if crop is None else (th, tw), flip_p=0.5, rand_augment=True, mean=CIFAR[0][:C], std=CIFAR[1][:C],
                         dtype=dtype)
    got = BatchTransform(spec, device="cuda")(torch.from_numpy(batch).cuda(), params=(top, left, flip, ops))
    want = TO.transform_batch(batch, crop_top=top, crop_left=left, out_size=(th, tw), flip=flip, ops=ops, mean=CIFAR[0][:C],
                              std=CIFAR[1][:C], dtype=dtype)
    assert same(got, want)


@pytest.mark.parametrize("cluster", [0, 2, 4, 8])
def test_forced_cluster_sizes_on_small_images(cluster):
    """The development knobs (flags bits 8-11: cluster size, bits 12-15: threads per CTA) run small
    images through the cluster path and every block size: odd chunk boundaries, every operation,
    scalar and vector store paths."""
    rng = np.random.default_rng(cluster)
    for (H, W, C) in ((32, 32, 3), (19, 23, 3), (28, 28, 1), (5, 7, 3)):
        B = 9
        batch = rng.integers(0, 256, (B, H, W, C), dtype=np.uint8)
        flip = rng.integers(0, 2, B).astype(np.uint8)
        ops = [[ALL_OPS[(5 * b + j) % len(ALL_OPS)] for j in range(5)] for b in range(B)]
        spec = TransformSpec(flip_p=0.5, rand_augment=True, mean=CIFAR[0][:C], std=CIFAR[1][:C])
        want = TO.transform_batch(batch, flip=flip, ops=ops, mean=CIFAR[0][:C], std=CIFAR[1][:C])
        for threads in (0, 2, 3, 4):                    # default, 256, 512, 1024
            bt = BatchTransform(spec, device="cuda")
            bt.flags = (cluster << 8) | (threads << 12)
            got = bt(torch.from_numpy(batch).cuda(), params=(None, None, flip, ops))
            assert same(got, want), (H, W, C, threads)


def test_image_sizes_around_the_shared_memory_opt_in():
    """Found by tools/soak.py: 2 x 20.5 KB of image + the kernel's 7 KB of static shared memory crosses the
    48 KB default although the dynamic part alone does not."""
    rng = np.random.default_rng(6)
    for (H, W) in ((58, 118), (92, 76), (64, 106), (63, 108), (80, 100)):
        batch = rng.integers(0, 256, (2, H, W, 3), dtype=np.uint8)
        ops = [[("Rotate", 12.0), ("Equalize", 0.0)], [("Sharpness", 0.5), ("AutoContrast", 0.0)]]
        spec = TransformSpec(rand_augment=True, mean=CIFAR[0], std=CIFAR[1])
        got = BatchTransform(spec, device="cuda")(torch.from_numpy(batch).cuda(), params=(None, None, None, ops))
        assert same(got, TO.transform_batch(batch, ops=ops, mean=CIFAR[0], std=CIFAR[1])), (H, W)


def test_mnist_rand_augment_replayed():
    """lenet.py:95: T.Compose([T.RandAugment(), T.ToTensor()]) on 28x28 grey images -- no crop, flip or
    rotation in the pipeline, so only RandAugment draws; two batches from one seed."""
    import torchvision.transforms as T
    from PIL import Image
    rng = np.random.default_rng(4)
    bt = BatchTransform(TransformSpec(rand_augment=True), device="cuda", seed=99)
    compose = T.Compose([T.RandAugment(), T.ToTensor()])
    torch.manual_seed(99)
    for _ in range(2):
        batch = rng.integers(0, 256, (128, 28, 28, 1), dtype=np.uint8)
        batch[batch < 140] = 0                                  # digit-like: mostly black
        want = torch.stack([compose(Image.fromarray(batch[b, :, :, 0])) for b in range(batch.shape[0])])
        assert same(bt(torch.from_numpy(batch).cuda()), want)


def test_larger_images_and_limits():
    rng = np.random.default_rng(2)
    batch = rng.integers(0, 256, (3, 150, 170, 3), dtype=np.uint8)          # 2 x 76.5 KB of shared memory
    ops = [[("Rotate", -30.0), ("Sharpness", 0.81), ("Equalize", 0.0)], [("ShearX", 0.3), ("Contrast", -0.9), ("AutoContrast", 0.0)],
           [("TranslateY", -50.0), ("Color", 0.9), ("Solarize", 10.0)]]
    spec = TransformSpec(rand_augment=True, mean=CIFAR[0], std=CIFAR[1])
    got = BatchTransform(spec, device="cuda")(torch.from_numpy(batch).cuda(), params=(None, None, None, ops))
    assert same(got, TO.transform_batch(batch, ops=ops, mean=CIFAR[0], std=CIFAR[1]))
    with pytest.raises(NotImplementedError, match="shared memory"):
        BatchTransform(spec, device="cuda")(torch.zeros((1, 600, 600, 3), dtype=torch.uint8, device="cuda"),
                                            params=(None, None, None, [[("Invert", 0.0)]]))
    with pytest.raises(ValueError, match="not in"):
        BatchTransform(TransformSpec(rand_augment=True), device="cuda")(
            torch.zeros((1, 8, 8, 4), dtype=torch.uint8, device="cuda"), params=(None, None, None, [[("Invert", 0.0)]]))
    with pytest.raises(ValueError, match="seed"):
        BatchTransform(spec, device="cuda")(torch.zeros((1, 8, 8, 3), dtype=torch.uint8, device="cuda"))
    assert TransformSpec(resize=64, rand_augment=True).two_stage        # resize hands a uint8 image to the augment kernel
    with pytest.raises(NotImplementedError, match="ragged"):
        BatchTransform(TransformSpec(resize=(8, 8), rand_augment=True), device="cuda").ragged([torch.zeros((9, 9, 3), dtype=torch.uint8)])
    # an empty batch and an empty operation list are fine
    empty = BatchTransform(spec, device="cuda")(torch.zeros((0, 8, 8, 3), dtype=torch.uint8, device="cuda"), params=(None, None, None, []))
    assert tuple(empty.shape) == (0, 3, 8, 8)
    plain = BatchTransform(spec, device="cuda")(torch.from_numpy(batch[:1]).cuda(), params=(None, None, None, [[]]))
    assert same(plain, TO.transform_batch(batch[:1], mean=CIFAR[0], std=CIFAR[1]))


IMAGENET = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))


@pytest.mark.parametrize("pre,flip_p,rot,ra,dtype", [(180, None, None, True, torch.float32),        # online.py:173
                                                      (None, 0.5, None, True, torch.bfloat16), (45, 0.5, 15, None, torch.float32),
                                                      (180, None, None, None, torch.float16)])
def test_resized_crop_chains_vs_the_compose(pre, flip_p, rot, ra, dtype):
    """T.RandomRotation(180) -> T.RandomResizedCrop(size) -> T.RandAugment() -> ToTensor -> Normalize
    (examples/img_stt/online/online.py:173) and its variants: the literal Compose under
    torch.manual_seed against the chained launches with replayed draws."""
    import torchvision.transforms as T
    from PIL import Image
    rng = np.random.default_rng(5)
    B, H, W, size = 8, 120, 160, 96
    batch = rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    stages = ([T.RandomRotation(pre)] if pre is not None else []) + [T.RandomResizedCrop(size)] + \
        ([T.RandomHorizontalFlip(flip_p)] if flip_p is not None else []) + ([T.RandomRotation(rot)] if rot is not None else []) + \
        ([T.RandAugment()] if ra else []) + [T.ToTensor(), T.Normalize(IMAGENET[0], IMAGENET[1])]
    compose = T.Compose(stages)
    torch.manual_seed(2718)
    want = torch.stack([compose(Image.fromarray(batch[b])) for b in range(B)]).to(dtype)
    spec = TransformSpec(random_resized_crop=size, pre_rotation=pre, flip_p=flip_p, rotation=rot, rand_augment=ra,
                         mean=IMAGENET[0], std=IMAGENET[1], dtype=dtype)
    got = BatchTransform(spec, device="cuda", seed=2718)(torch.from_numpy(batch).cuda())
    assert same(got, want)


def test_resized_crop_chain_on_frames_beyond_4x():
    """Found by tools/soak.py: the chained pipelines also take the per-sample-records route when the frame is
    more than four times the output (236x188 -> 46) or its rows are not 4-byte aligned (W * 3 = 375)."""
    import torchvision.transforms as T
    from PIL import Image
    rng = np.random.default_rng(12)
    for (H, W, size, pre, ra) in ((236, 188, 46, 180, False), (216, 144, 13, 15, True), (90, 125, 40, None, True)):
        batch = rng.integers(0, 256, (3, H, W, 3), dtype=np.uint8)
        stages = ([T.RandomRotation(pre)] if pre is not None else []) + [T.RandomResizedCrop(size), T.RandomHorizontalFlip()] + \
            ([T.RandAugment()] if ra else []) + [T.ToTensor(), T.Normalize(IMAGENET[0], IMAGENET[1])]
        compose = T.Compose(stages)
        torch.manual_seed(77)
        want = torch.stack([compose(Image.fromarray(batch[b])) for b in range(3)])
        spec = TransformSpec(random_resized_crop=size, pre_rotation=pre, flip_p=0.5, rand_augment=ra, mean=IMAGENET[0], std=IMAGENET[1])
        got = BatchTransform(spec, device="cuda", seed=77)(torch.from_numpy(batch).cuda())
        assert same(got, want), (H, W, size, pre, ra)


def _resize_chain_cases():
    import torchvision.transforms as T
    return [
        # Resize -> CenterCrop -> RandomCrop(reflect padding): the padding mirrors the centre crop, not the resized image
        ((96, 128), dict(resize=64, center_crop=56, crop=56, padding=4, padding_mode="reflect", flip_p=0.5),
         lambda: [T.Resize(64), T.CenterCrop(56), T.RandomCrop(56, padding=4, padding_mode="reflect"), T.RandomHorizontalFlip()]),
        # Resize -> RandomCrop(symmetric padding), no centre crop
        ((90, 70), dict(resize=(64, 60), crop=48, padding=6, padding_mode="symmetric", flip_p=0.5),
         lambda: [T.Resize((64, 60)), T.RandomCrop(48, padding=6, padding_mode="symmetric"), T.RandomHorizontalFlip()]),
        # CenterCrop -> RandomCrop(zero padding), no resize: the padding must not show what the centre crop cut away
        ((60, 60), dict(center_crop=40, crop=40, padding=4, flip_p=None),
         lambda: [T.CenterCrop(40), T.RandomCrop(40, padding=4)]),
        # CenterCrop LARGER than the image (zero fill) then edge padding: the edge is the zero fill
        ((30, 44), dict(center_crop=(40, 50), crop=(36, 50), padding=3, padding_mode="edge", flip_p=0.5),
         lambda: [T.CenterCrop((40, 50)), T.RandomCrop((36, 50), padding=3, padding_mode="edge"), T.RandomHorizontalFlip()]),
        # the ImageNet evaluation resize followed by the training augmentations
        ((300, 400), dict(resize=256, center_crop=224, flip_p=0.5, rand_augment=True),
         lambda: [T.Resize(256), T.CenterCrop(224), T.RandomHorizontalFlip(), T.RandAugment()]),
        ((150, 130), dict(resize=72, center_crop=64, crop=64, padding=4, padding_mode="edge", flip_p=0.5, rotation=15, rand_augment=True),
         lambda: [T.Resize(72), T.CenterCrop(64), T.RandomCrop(64, padding=4, padding_mode="edge"), T.RandomHorizontalFlip(),
                  T.RandomRotation(15), T.RandAugment()]),
    ]


@pytest.mark.parametrize("case", range(6))
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_resize_then_padding_or_augment_vs_the_compose(case, dtype):
    """Resize / CenterCrop pipelines whose tail needs the resized image as a whole (padding that mirrors it, padding after a
    centre crop, rotation / RandAugment): two launches with a planar uint8 hand-off, against the literal Compose under
    torch.manual_seed with the draws replayed from the same seed."""
    import torchvision.transforms as T
    from PIL import Image
    (H, W), kw, stages = _resize_chain_cases()[case]
    rng = np.random.default_rng(40 + case)
    B = 6
    batch = rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    compose = T.Compose(stages() + [T.ToTensor(), T.Normalize(IMAGENET[0], IMAGENET[1])])
    spec = TransformSpec(mean=IMAGENET[0], std=IMAGENET[1], dtype=dtype, **kw)
    assert spec.two_stage
    bt = BatchTransform(spec, device="cuda", seed=31 + case)
    torch.manual_seed(31 + case)
    for step in range(2):           # two consecutive batches from one generator
        want = torch.stack([compose(Image.fromarray(batch[b])) for b in range(B)]).to(dtype)
        got = bt(torch.from_numpy(batch).cuda())
        assert same(got, want), (case, step)
    # explicit parameters and the oracle's own chain
    g = bt.geometry(H, W)
    th, tw = g["out"]
    ph, pw = g["cur"][0] + 2 * g["pad"][1], g["cur"][1] + 2 * g["pad"][0]
    top = rng.integers(0, ph - th + 1, B).astype(np.int32)
    left = rng.integers(0, pw - tw + 1, B).astype(np.int32)
    flip = rng.integers(0, 2, B).astype(np.uint8)
    ops = [[("Rotate", 21.0), ("Posterize", 5.0)] if b % 2 else [] for b in range(B)] if spec.augments else None
    params = (top, left, flip, ops) if spec.augments else (top, left, flip)
    got = BatchTransform(spec, device="cuda")(torch.from_numpy(batch).cuda(), params=params)
    pad = kw.get("padding", 0)
    want = TO.transform_batch(batch, resize=kw.get("resize"), center_crop=kw.get("center_crop"), padding=pad,
                              padding_mode=kw.get("padding_mode", "constant"), crop_top=top, crop_left=left, out_size=(th, tw),
                              flip=flip, ops=ops, mean=IMAGENET[0], std=IMAGENET[1], dtype=dtype)
    assert same(got, want), case
