"""Parity of the CUDA transform path against the oracle (torchvision PIL backend), through the
C ABI.  Integer decisions (crop, flip, padding) and the table values must be BIT-EXACT: the
fp32 table reproduces torchvision's ((x/255) - mean)/std exactly, so tolerance is 0 ulp (the
north star allows 1); bf16 must equal RNE(.to(bfloat16)) of that value."""
import numpy as np
import pytest
import torch

from oracle import params_oracle as PO
from oracle import transform_oracle as TO
from torchbooster_b200 import _native as N
from torchbooster_b200.transforms import BatchTransform, CropFlipParams, TransformSpec

pytestmark = pytest.mark.gpu

IMAGENET = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
CIFAR = ((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))


def same(a: torch.Tensor, b: torch.Tensor) -> bool:
    a, b = a.cpu(), b.cpu()
    if a.dtype != b.dtype or a.shape != b.shape:
        return False
    view = {torch.float32: torch.int32, torch.bfloat16: torch.int16, torch.float16: torch.int16}[a.dtype]
    return torch.equal(a.contiguous().view(view), b.contiguous().view(view))


def run_both_paths(spec, batch, params):
    """Staged fast path (when eligible) and the gather path must both equal the oracle."""
    outs = []
    for flags in (0, N.FLAG_FORCE_GENERIC, N.FLAG_FORCE_STREAM, N.FLAG_FORCE_STREAM | N.FLAG_NO_PDL):
        bt = BatchTransform(spec, device="cuda")
        bt.flags = flags
        outs.append(bt(torch.from_numpy(batch).cuda(), params=params))
    torch.cuda.synchronize()
    return outs


def test_lut_is_torchvision_exact():
    for dtype in (torch.float32, torch.bfloat16, torch.float16):
        for mean, std in (IMAGENET, CIFAR, ((.5, .5, .5), (.5, .5, .5))):
            bt = BatchTransform(TransformSpec(mean=mean, std=std, dtype=dtype), device="cuda")
            x = torch.arange(256, dtype=torch.uint8).view(1, 256, 1).repeat(3, 1, 1)
            t = x.to(torch.float32).div(255)
            t.sub_(torch.tensor(mean).view(-1, 1, 1)).div_(torch.tensor(std).view(-1, 1, 1))
            assert same(bt.lut(3), t.view(3, 256).to(dtype))
    bt = BatchTransform(TransformSpec(one_minus=True), device="cuda")
    want = 1.0 - torch.arange(256, dtype=torch.uint8).to(torch.float32).div(255)
    assert same(bt.lut(1), want.view(1, 256))


def test_golden_fixtures(golden):
    g = golden("transform_torchvision.npz")
    spec = TransformSpec(crop=32, padding=4, padding_mode="reflect", flip_p=.5, mean=CIFAR[0], std=CIFAR[1])
    for out in run_both_paths(spec, g["c1_in"], (g["c1_top"], g["c1_left"], g["c1_flip"])):
        assert same(out, torch.from_numpy(g["c1_out"]))
    spec = TransformSpec(crop=(40, 56), flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    want = torch.from_numpy(g["c2_out_bf16_bits"]).view(torch.bfloat16)
    for out in run_both_paths(spec, g["c2_in"], (g["c2_top"], g["c2_left"], g["c2_flip"])):
        assert same(out, want)
    spec = TransformSpec(flip_p=.5, mean=[.5] * 3, std=[.5] * 3)
    for out in run_both_paths(spec, g["c3_in"], (None, None, g["c3_flip"])):
        assert same(out, torch.from_numpy(g["c3_out"]))
    spec = TransformSpec(one_minus=True)
    for out in run_both_paths(spec, g["gan_in"], None):
        assert same(out, torch.from_numpy(g["gan_out"]))


GRID = [
    # B, H, W, C, crop, padding, mode, dtype
    (5, 32, 32, 3, (32, 32), 4, "reflect", torch.float32),       # C1 CIFAR
    (3, 256, 256, 3, (224, 224), 0, "constant", torch.bfloat16),  # C2 ImageNet
    (6, 64, 64, 3, None, 0, "constant", torch.float32),           # C3 whole image, flip only
    (2, 40, 48, 3, (24, 40), 0, "constant", torch.float16),
    (3, 33, 37, 3, (20, 31), 0, "constant", torch.float32),       # odd W: pitch not 16-byte aligned
    (3, 40, 48, 3, (33, 24), 5, "symmetric", torch.bfloat16),
    (3, 40, 48, 3, (40, 48), 6, "edge", torch.float32),
    (3, 40, 48, 3, (44, 56), 7, "constant", torch.bfloat16),
    (1, 17, 19, 1, (9, 11), 2, "reflect", torch.float32),         # B=1, grayscale
    (2, 24, 16, 4, (16, 8), 0, "constant", torch.float32),        # RGBA -- gather path (PIL handles 4 channels)
    (2, 300, 512, 3, (300, 512), 0, "constant", torch.bfloat16),  # wide rows, whole image
    (4, 128, 1024, 3, (100, 1000), 0, "constant", torch.float32), # crop wider than one ring budget row
]


@pytest.mark.parametrize("B,H,W,C,crop,padding,mode,dtype", GRID)
def test_grid_vs_oracle(B, H, W, C, crop, padding, mode, dtype):
    rng = np.random.default_rng(B * 1000 + H + W)
    batch = rng.integers(0, 256, (B, H, W, C), dtype=np.uint8)
    th, tw = crop if crop else (H, W)
    top = rng.integers(0, H + 2 * padding - th + 1, B).astype(np.int32)
    left = rng.integers(0, W + 2 * padding - tw + 1, B).astype(np.int32)
    flip = rng.integers(0, 2, B).astype(np.uint8)
    # exercise all four borders
    top[0], left[0] = 0, 0
    top[-1], left[-1] = H + 2 * padding - th, W + 2 * padding - tw
    mean, std = (IMAGENET if C == 3 else ((.5,) * C, (.25,) * C))
    spec = TransformSpec(crop=crop, padding=padding, padding_mode=mode, flip_p=.5, mean=mean, std=std, dtype=dtype)
    want = TO.transform_batch(batch, crop_top=top, crop_left=left, out_size=(th, tw), flip=flip, padding=padding,
                              padding_mode=mode, mean=mean, std=std, dtype=dtype)
    for out in run_both_paths(spec, batch, (top, left, flip)):
        assert same(out, want)


def test_strided_source_and_out_argument():
    """A batch that is a strided view (crop of a larger buffer) and a caller-provided output."""
    rng = np.random.default_rng(9)
    big = torch.from_numpy(rng.integers(0, 256, (4, 80, 96, 3), dtype=np.uint8)).cuda()
    view = big[1:4, 8:72, 16:80, :]          # [3, 64, 64, 3], row pitch 288, batch stride 23040
    top = np.array([0, 5, 16], np.int32); left = np.array([16, 3, 0], np.int32); flip = np.array([1, 0, 1], np.uint8)
    spec = TransformSpec(crop=48, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    want = TO.transform_batch(view.cpu().numpy(), crop_top=top, crop_left=left, out_size=(48, 48), flip=flip,
                              mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    bt = BatchTransform(spec, device="cuda")
    out = torch.empty((3, 3, 48, 48), dtype=torch.bfloat16, device="cuda")
    res = bt(view, params=(top, left, flip), out=out)
    assert res.data_ptr() == out.data_ptr() and same(out, want)


def test_replayed_rng_equals_torchvision_draws():
    """Seeded BatchTransform draws what T.RandomCrop/T.RandomHorizontalFlip draw under manual_seed."""
    B, H, W = 48, 72, 80
    rng = np.random.default_rng(21)
    batch = rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    top, left, flip = PO.draw_crop_flip(1234, B, H, W, 64, 64, 0)
    want = TO.transform_batch(batch, crop_top=top, crop_left=left, out_size=(64, 64), flip=flip,
                              mean=IMAGENET[0], std=IMAGENET[1])
    bt = BatchTransform(TransformSpec(crop=64, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1]), device="cuda", seed=1234)
    assert same(bt(torch.from_numpy(batch).cuda()), want)


def test_host_batch_is_copied():
    rng = np.random.default_rng(2)
    batch = torch.from_numpy(rng.integers(0, 256, (4, 32, 32, 3), dtype=np.uint8)).pin_memory()
    bt = BatchTransform(TransformSpec(mean=[.5] * 3, std=[.5] * 3), device="cuda")
    want = TO.transform_batch(batch.numpy(), mean=[.5] * 3, std=[.5] * 3)
    assert same(bt(batch), want)


def test_full_size_c2_properties():
    """BASELINE config 2 at full size (512 x 256x256 -> 224, bf16): checked through
    size-independent properties -- (1) flipping twice is the identity on the crop, (2) the
    unflipped output equals a table lookup of the strided crop view, (3) fast == gather path."""
    B = 512
    gen = torch.Generator(device="cuda").manual_seed(5)
    batch = torch.randint(0, 256, (B, 256, 256, 3), dtype=torch.uint8, device="cuda", generator=gen)
    top, left, flip = CropFlipParams(42).draw(B, 256, 256, (224, 224), (0, 0), 0.5)
    spec = TransformSpec(crop=224, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    bt = BatchTransform(spec, device="cuda")
    out = bt(batch, params=(top, left, flip))
    bt_g = BatchTransform(spec, device="cuda"); bt_g.flags = N.FLAG_FORCE_GENERIC
    assert same(out, bt_g(batch, params=(top, left, flip)))
    unflipped = bt(batch, params=(top, left, np.zeros(B, np.uint8)))
    f = torch.from_numpy(flip.astype(bool)).cuda()
    assert same(torch.where(f.view(B, 1, 1, 1), out.flip(-1), out), unflipped)
    lut = bt.lut(3)
    for b in (0, 17, 511):
        crop = batch[b, top[b]:top[b] + 224, left[b]:left[b] + 224, :].permute(2, 0, 1).long()
        want = torch.stack([lut[c][crop[c]] for c in range(3)])
        assert same(unflipped[b], want)


def test_argument_errors():
    bt = BatchTransform(TransformSpec(), device="cuda")
    with pytest.raises(TypeError):
        bt(torch.zeros(2, 8, 8, 3, device="cuda"))
    L = N.lib()
    x = torch.zeros(1, 8, 8, 3, dtype=torch.uint8, device="cuda")
    o = torch.zeros(1, 3, 8, 8, device="cuda")
    lut = bt.lut(3)
    rc = L.tb_transform_u8_nhwc_to_nchw(x.data_ptr(), 1, 8, 8, 5, 192, 24, o.data_ptr(), 0, 8, 8, None, None, None,
                                        0, 0, 0, lut.data_ptr(), 0, None)
    assert rc == N.TB_EINVAL and "C=5" in N.last_error()
    rc = L.tb_transform_u8_nhwc_to_nchw(x.data_ptr(), 1, 8, 8, 3, 192, 24, o.data_ptr(), 0, 8, 8, None, None, None,
                                        8, 8, 2, lut.data_ptr(), 0, None)
    assert rc == N.TB_EINVAL and "reflect" in N.last_error()
    # empty batch is a no-op
    assert L.tb_transform_u8_nhwc_to_nchw(None, 0, 8, 8, 3, 192, 24, None, 0, 8, 8, None, None, None, 0, 0, 0, None,
                                          0, None) == 0


def test_pinned_batch_is_read_in_place():
    """Zero-copy: a pinned host batch goes to the kernel as is (host pointer); same bits as the
    copied path, for the streaming kernel and for the gather kernel."""
    rng = np.random.default_rng(12)
    B = 6
    batch = torch.from_numpy(rng.integers(0, 256, (B, 64, 80, 3), dtype=np.uint8)).pin_memory()
    top = rng.integers(0, 64 - 48 + 1, B).astype(np.int32); left = rng.integers(0, 80 - 64 + 1, B).astype(np.int32)
    flip = rng.integers(0, 2, B).astype(np.uint8)
    want = TO.transform_batch(batch.numpy(), crop_top=top, crop_left=left, out_size=(48, 64), flip=flip,
                              mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    spec = TransformSpec(crop=(48, 64), flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    for flags in (0, N.FLAG_FORCE_GENERIC):
        for zero_copy in (True, False):
            bt = BatchTransform(spec, device="cuda", zero_copy=zero_copy)
            bt.flags = flags
            assert same(bt(batch, params=(top, left, flip)), want)


def _cpu_lut(mean, std, dtype):
    x = torch.arange(256, dtype=torch.uint8).to(torch.float32).div(255)
    t = (x[None, :] - torch.tensor(mean, dtype=torch.float32)[:, None]) / torch.tensor(std, dtype=torch.float32)[:, None]
    return t.to(dtype)


@pytest.mark.parametrize("seed", range(6))
def test_fuzz_geometry_against_index_map(seed):
    """Random shapes / paddings / modes / dtypes against the numpy index-map restatement (itself
    pinned to torchvision in tests/test_oracle.py); C == 3 cases run the streaming kernel, the
    gather kernel and the forced-streaming route."""
    rng = np.random.default_rng(1000 + seed)
    for _ in range(10):
        C = int(rng.choice([1, 3, 3, 3, 4]))
        H, W = int(rng.integers(4, 90)), int(rng.integers(4, 120))
        if rng.random() < 0.4:
            W = int(rng.choice([16, 32, 64, 48, 80, 96]))        # 16-byte aligned pitches reach the streaming kernel
        mode = str(rng.choice(["constant", "reflect", "edge", "symmetric"]))
        pad = int(rng.integers(0, min(H, W, 9)))
        dtype = [torch.float32, torch.bfloat16, torch.float16][int(rng.integers(0, 3))]
        px = 4 if dtype == torch.float32 else 8
        th = int(rng.integers(1, H + 2 * pad + 1))
        tw = int(rng.integers(1, W + 2 * pad + 1))
        if rng.random() < 0.6 and (W + 2 * pad) >= px:
            tw = max(px, tw // px * px)
        B = int(rng.integers(1, 6))
        batch = rng.integers(0, 256, (B, H, W, C), dtype=np.uint8)
        top = rng.integers(0, H + 2 * pad - th + 1, B).astype(np.int32)
        left = rng.integers(0, W + 2 * pad - tw + 1, B).astype(np.int32)
        flip = rng.integers(0, 2, B).astype(np.uint8)
        mean = tuple(float(v) for v in rng.uniform(0.2, 0.6, C))
        std = tuple(float(v) for v in rng.uniform(0.1, 0.4, C))
        lut = _cpu_lut(mean, std, dtype)
        lut_np = lut.view(torch.int16 if dtype != torch.float32 else torch.int32).numpy()
        want = np.stack([TO.index_map_sample(batch[b], lut_np, top=int(top[b]), left=int(left[b]), out_h=th, out_w=tw,
                                             flip=bool(flip[b]), pad_y=pad, pad_x=pad, pad_mode=mode) for b in range(B)])
        spec = TransformSpec(crop=(th, tw), padding=pad, padding_mode=mode, flip_p=.5, mean=mean, std=std, dtype=dtype)
        for flags in (0, N.FLAG_FORCE_GENERIC, N.FLAG_FORCE_STREAM):
            bt = BatchTransform(spec, device="cuda")
            bt.flags = flags
            got = bt(torch.from_numpy(batch).cuda(), params=(top, left, flip)).cpu()
            got_bits = got.view(torch.int16 if dtype != torch.float32 else torch.int32).numpy()
            assert np.array_equal(got_bits, want), (seed, C, H, W, th, tw, pad, mode, dtype, flags)
