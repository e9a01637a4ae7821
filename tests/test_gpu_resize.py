"""Parity of the CUDA resize path against Pillow (through torchvision's PIL backend): bit-exact."""
import numpy as np
import pytest
import torch

from oracle import transform_oracle as TO
from torchbooster_b200.transforms import BatchTransform, TransformSpec

pytestmark = pytest.mark.gpu
IMAGENET = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))


def same(a, b):
    a, b = a.cpu().contiguous(), b.cpu().contiguous()
    view = {torch.float32: torch.int32, torch.bfloat16: torch.int16, torch.float16: torch.int16}[a.dtype]
    return a.dtype == b.dtype and a.shape == b.shape and torch.equal(a.view(view), b.view(view))


def test_golden_c4(golden):
    g = golden("transform_torchvision.npz")
    bt = BatchTransform(TransformSpec(resize=32, center_crop=32, mean=IMAGENET[0], std=IMAGENET[1]), device="cuda")
    assert same(bt(torch.from_numpy(g["c4_in"]).cuda()), torch.from_numpy(g["c4_out"]))


def test_golden_pillow_resize(golden):
    """Raw resize results: table = identity-like ToTensor, compare bytes via x/255 table inverse."""
    g = golden("resize_pillow.npz")
    for i in range(int(g["n_cases"])):
        src, want = g[f"r{i}_in"], g[f"r{i}_out"]
        if src.ndim == 2:
            src, want = src[:, :, None], want[:, :, None]
        bt = BatchTransform(TransformSpec(resize=want.shape[:2]), device="cuda")
        out = bt(torch.from_numpy(src[None]).cuda())
        ref = torch.from_numpy(want).permute(2, 0, 1)[None].float().div(255)
        assert same(out, ref), f"case {i}"


BOTH_KERNELS = (0, 16, 32)      # default (ring kernel where rows are 16-byte aligned, else sliding-window), TB_FLAG_TWO_PHASE, TB_FLAG_SLIDE

CASES = [
    # B, H, W, C, resize, center_crop, dtype
    (2, 512, 512, 3, 256, 256, torch.float32),       # C4
    (2, 375, 500, 3, 256, 256, torch.float32),       # adain.py on a COCO-shaped image
    (1, 480, 640, 3, 512, 512, torch.bfloat16),      # upscale
    (2, 300, 200, 3, (384, 256), None, torch.float32),
    (2, 64, 48, 1, 32, None, torch.float32),
    (1, 50, 50, 3, 20, 32, torch.float32),           # centre crop larger than the image: zero fill
    (1, 41, 57, 3, (41, 30), None, torch.float32),   # horizontal pass only
    (1, 41, 57, 3, (19, 57), None, torch.float32),   # vertical pass only
    (1, 1000, 900, 3, 64, 64, torch.float32),        # large reduction: 31-tap filters
    (1, 37, 61, 3, None, 24, torch.float16),         # centre crop only
    (2, 300, 260, 3, 96, 96, torch.float32),         # 2.7x - 3.1x reduction: eight-tap DP4A path
    (2, 256, 512, 3, (64, 128), None, torch.bfloat16),   # exactly 4x
    (1, 250, 128, 1, (63, 32), None, torch.float32),     # just under 4x, grayscale
]


@pytest.mark.parametrize("B,H,W,C,resize,cc,dtype", CASES)
def test_resize_grid_vs_oracle(B, H, W, C, resize, cc, dtype):
    rng = np.random.default_rng(H * 7 + W)
    batch = rng.integers(0, 256, (B, H, W, C), dtype=np.uint8)
    mean, std = (IMAGENET if C == 3 else ((.5,), (.5,)))
    want = TO.transform_batch(batch, resize=resize, center_crop=cc, mean=mean, std=std, dtype=dtype)
    for flags in BOTH_KERNELS:
        bt = BatchTransform(TransformSpec(resize=resize, center_crop=cc, mean=mean, std=std, dtype=dtype), device="cuda")
        bt.flags = flags
        assert same(bt(torch.from_numpy(batch).cuda()), want), flags


def test_resize_then_random_crop_flip():
    rng = np.random.default_rng(77)
    B = 3
    batch = rng.integers(0, 256, (B, 90, 120, 3), dtype=np.uint8)
    top = np.array([0, 4, 8], np.int32); left = np.array([21, 0, 7], np.int32); flip = np.array([1, 1, 0], np.uint8)
    want = TO.transform_batch(batch, resize=48, crop_top=top, crop_left=left, out_size=(40, 43), flip=flip,
                              mean=IMAGENET[0], std=IMAGENET[1])
    for flags in BOTH_KERNELS:
        bt = BatchTransform(TransformSpec(resize=48, crop=(40, 43), flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1]),
                            device="cuda")
        bt.flags = flags
        assert same(bt(torch.from_numpy(batch).cuda(), params=(top, left, flip)), want), flags


def test_resize_linearity_property_full_size():
    """Full C4 size (32 x 512x512): a constant image resizes to the same constant, and resizing
    is monotone: img_a <= img_b pointwise implies out_a <= out_b (non-negative taps)."""
    gen = torch.Generator(device="cuda").manual_seed(3)
    a = torch.randint(0, 200, (32, 512, 512, 3), dtype=torch.uint8, device="cuda", generator=gen)
    b = a + torch.randint(0, 56, a.shape, dtype=torch.uint8, device="cuda", generator=gen)
    bt = BatchTransform(TransformSpec(resize=256, center_crop=256), device="cuda")
    oa, ob = bt(a), bt(b)
    assert bool((oa <= ob).all())
    c = torch.full((2, 512, 512, 3), 77, dtype=torch.uint8, device="cuda")
    assert bool((bt(c) == torch.tensor(77.0).div(255)).all())


@pytest.mark.parametrize("seed", range(4))
def test_fuzz_resize_against_integer_oracle(seed):
    """Random source / target sizes (up- and down-scaling, one-pass cases, > 2x reductions), crops
    with constant padding and flips against the integer restatement of Pillow's resample (pinned
    to Pillow in tests/test_oracle.py) followed by the index map."""
    from oracle import resize_oracle as RO
    rng = np.random.default_rng(2000 + seed)
    for _ in range(8):
        C = int(rng.choice([1, 3, 3]))
        H, W = int(rng.integers(6, 110)), int(rng.integers(6, 130))
        if rng.random() < 0.5:
            W = int(rng.choice([32, 64, 96, 128]))            # word-readable rows: the DP4A path when scale <= 2
        rs_h = int(np.clip(int(H * rng.uniform(0.3, 1.6)), 1, 200))
        rs_w = int(np.clip(int(W * rng.uniform(0.3, 1.6)), 1, 200))
        if rng.random() < 0.15:
            rs_h = H
        if rng.random() < 0.15:
            rs_w = W
        pad = int(rng.integers(0, 4))
        th = int(rng.integers(1, rs_h + 2 * pad + 1)); tw = int(rng.integers(1, rs_w + 2 * pad + 1))
        B = int(rng.integers(1, 4))
        dtype = [torch.float32, torch.bfloat16][int(rng.integers(0, 2))]
        batch = rng.integers(0, 256, (B, H, W, C), dtype=np.uint8)
        top = rng.integers(0, rs_h + 2 * pad - th + 1, B).astype(np.int32)
        left = rng.integers(0, rs_w + 2 * pad - tw + 1, B).astype(np.int32)
        flip = rng.integers(0, 2, B).astype(np.uint8)
        mean = tuple(float(v) for v in rng.uniform(0.2, 0.6, C)); std = tuple(float(v) for v in rng.uniform(0.1, 0.4, C))
        x = torch.arange(256, dtype=torch.uint8).to(torch.float32).div(255)
        lut = ((x[None, :] - torch.tensor(mean)[:, None]) / torch.tensor(std)[:, None]).to(dtype)
        view = torch.int32 if dtype == torch.float32 else torch.int16
        want = []
        for b in range(B):
            img = batch[b] if C == 3 else batch[b, :, :, 0]
            rs = RO.resize_u8(img, rs_h, rs_w)
            rs = rs if C == 3 else rs[:, :, None]
            want.append(TO.index_map_sample(rs, lut.view(view).numpy(), top=int(top[b]), left=int(left[b]), out_h=th,
                                            out_w=tw, flip=bool(flip[b]), pad_y=pad, pad_x=pad, pad_mode="constant"))
        spec = TransformSpec(resize=(rs_h, rs_w), crop=(th, tw), padding=pad, flip_p=.5, mean=mean, std=std, dtype=dtype)
        for flags in (0, 1, 16, 32):
            bt = BatchTransform(spec, device="cuda")
            bt.flags = flags
            got = bt(torch.from_numpy(batch).cuda(), params=(top, left, flip)).cpu().view(view).numpy()
            assert np.array_equal(got, np.stack(want)), (seed, C, H, W, rs_h, rs_w, th, tw, pad, dtype, flags)


def test_resized_crop_golden(golden):
    g = golden("transform_torchvision.npz")
    spec = TransformSpec(random_resized_crop=24, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1])
    bt = BatchTransform(spec, device="cuda")
    out = bt(torch.from_numpy(g["rrc_in"]).cuda(), params=(g["rrc_boxes"], g["rrc_flip"]))
    assert same(out, torch.from_numpy(g["rrc_out"]))


@pytest.mark.parametrize("B,H,W,C,size,dtype", [(24, 256, 256, 3, 224, torch.bfloat16), (16, 64, 96, 3, 48, torch.float32),
                                                (8, 50, 40, 1, 32, torch.float32), (6, 300, 200, 3, (160, 100), torch.float16),
                                                (12, 375, 500, 3, 224, torch.bfloat16),      # ImageNet-sized source: up to 2.2x
                                                (8, 120, 100, 3, 32, torch.float32)])        # up to 3.75x: eight taps
def test_random_resized_crop_vs_torchvision(B, H, W, C, size, dtype):
    """Seeded RandomResizedCrop + flip: boxes drawn by the replayed torch generator, every box
    resized on the GPU with taps computed in-kernel; bit-exact against torchvision's own draws
    and F.resized_crop (PIL)."""
    from oracle import params_oracle as PO
    rng = np.random.default_rng(H + W)
    batch = rng.integers(0, 256, (B, H, W, C), dtype=np.uint8)
    boxes, flip = PO.draw_resized_crop_flip(77, B, H, W)
    mean, std = (IMAGENET if C == 3 else ((.5,), (.5,)))
    out_size = (size, size) if isinstance(size, int) else size
    want = TO.transform_batch(batch, boxes=boxes, out_size=out_size, flip=flip, mean=mean, std=std, dtype=dtype)
    for flags in BOTH_KERNELS:
        bt = BatchTransform(TransformSpec(random_resized_crop=size, flip_p=.5, mean=mean, std=std, dtype=dtype),
                            device="cuda", seed=77)
        bt.flags = flags
        assert same(bt(torch.from_numpy(batch).cuda()), want), flags


def test_resized_crop_edge_boxes_and_limits():
    rng = np.random.default_rng(4)
    H, W, S = 64, 64, 32
    batch = rng.integers(0, 256, (6, H, W, 3), dtype=np.uint8)
    # whole image, single pixel, one row, one column, identity-size box, bottom-right corner
    boxes = np.array([[0, 0, 64, 64], [10, 20, 1, 1], [5, 0, 1, 64], [0, 7, 64, 1], [16, 16, 32, 32], [63, 63, 1, 1]], np.int32)
    flip = np.array([0, 1, 0, 1, 1, 0], np.uint8)
    want = TO.transform_batch(batch, boxes=boxes, out_size=(S, S), flip=flip, mean=IMAGENET[0], std=IMAGENET[1])
    for flags in BOTH_KERNELS:
        bt = BatchTransform(TransformSpec(random_resized_crop=S, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1]), device="cuda")
        bt.flags = flags
        assert same(bt(torch.from_numpy(batch).cuda(), params=(boxes, flip)), want), flags
        # a box outside the image gives zeros (lut[0]) instead of reading out of bounds
        bad = boxes.copy(); bad[0] = [60, 60, 10, 10]
        out = bt(torch.from_numpy(batch).cuda(), params=(bad, flip))
        assert same(out[1:], want[1:]) and bool((out[0] == bt.lut(3)[:, 0].view(3, 1, 1)).all())
    # frames more than 4x the output (the uniform entry point refuses them) go through per-sample records;
    # boxes beyond 4x take the any-scale kernel; rows that are not 4-byte aligned are re-pitched
    for (Hb, Wb) in ((200, 200), (150, 201)):
        big = rng.integers(0, 256, (3, Hb, Wb, 3), dtype=np.uint8)
        bx = np.array([[0, 0, Hb, Wb], [20, 30, 100, 90], [7, 11, Hb - 10, 150]], np.int32)
        fl = np.array([1, 0, 1], np.uint8)
        want = TO.transform_batch(big, boxes=bx, out_size=(S, S), flip=fl, mean=IMAGENET[0], std=IMAGENET[1])
        got = bt(torch.from_numpy(big).cuda(), params=(bx, fl))
        assert same(got, want), (Hb, Wb, [int((got[b].cpu() != want[b]).sum()) for b in range(3)])
        outside = bx.copy(); outside[2] = [7, 11, Hb, 150]              # leaves the image: zeros, nothing read out of bounds
        got = bt(torch.from_numpy(big).cuda(), params=(outside, fl))
        assert same(got[:2], want[:2]) and bool((got[2] == bt.lut(3)[:, 0].view(3, 1, 1)).all())
    with pytest.raises(NotImplementedError, match="four times"):     # the C entry point itself still says so
        from torchbooster_b200 import _native as N
        z = torch.zeros((1, 200, 200, 3), dtype=torch.uint8, device="cuda")
        o = torch.empty((1, 3, S, S), device="cuda")
        bxd = torch.tensor([[0, 0, 200, 200]], dtype=torch.int32, device="cuda")
        N.check(N.lib().tb_resized_crop_bilinear_u8_nhwc_to_nchw(z.data_ptr(), 1, 200, 200, 3, z.stride(0), z.stride(1), bxd.data_ptr(),
                                                                o.data_ptr(), 0, S, S, None, bt.lut(3).data_ptr(), 0,
                                                                N.current_stream_ptr(z.device)))


def _ragged_images(rng, sizes, C=3):
    return [rng.integers(0, 256, (h, w, C), dtype=np.uint8) for h, w in sizes]


@pytest.mark.parametrize("resize,cc,dtype", [(64, 64, torch.float32),          # adain.py:170 at a small size
                                             (48, 40, torch.bfloat16), (40, 56, torch.float32)])   # crop smaller / larger than the resized image
def test_ragged_resize_center_crop_vs_torchvision(resize, cc, dtype):
    """T.Resize(int) + T.CenterCrop over images of DIFFERENT sizes in one launch: portrait, landscape,
    square, up- and down-scaling (1x .. 3.9x), identity-sized ones, odd widths (padded pitches)."""
    rng = np.random.default_rng(resize)
    big = int(3.9 * resize)
    sizes = [(97, 131), (131, 97), (64, 64), (50, 3 * resize), (big, 77), (resize, resize + 9), (40, 41), (big - 1, big), (33, 129)]
    imgs = _ragged_images(rng, sizes)
    spec = TransformSpec(resize=resize, center_crop=cc, mean=IMAGENET[0], std=IMAGENET[1], dtype=dtype)
    got = BatchTransform(spec, device="cuda").ragged(imgs)
    want = torch.stack([TO.transform_sample(im, resize=resize, center_crop=cc, mean=IMAGENET[0], std=IMAGENET[1], dtype=dtype)
                        for im in imgs])
    assert same(got, want)


def test_ragged_random_resized_crop_vs_torchvision():
    """T.RandomResizedCrop(96) + flip over differently sized images: the literal transform under
    torch.manual_seed against one launch with replayed draws; grey images too."""
    import torchvision.transforms as T
    from PIL import Image
    rng = np.random.default_rng(3)
    for C in (3, 1):
        sizes = [(int(h), int(w)) for h, w in zip(rng.integers(60, 300, 12), rng.integers(60, 300, 12))]
        imgs = _ragged_images(rng, sizes, C)
        mean, std = (IMAGENET if C == 3 else ((.5,), (.5,)))
        compose = T.Compose([T.RandomResizedCrop(96), T.RandomHorizontalFlip(), T.ToTensor(), T.Normalize(mean, std)])
        torch.manual_seed(1234)
        want = torch.stack([compose(Image.fromarray(im if C == 3 else im[:, :, 0])) for im in imgs])
        spec = TransformSpec(random_resized_crop=96, flip_p=.5, mean=mean, std=std)
        got = BatchTransform(spec, device="cuda", seed=1234).ragged(imgs)
        assert same(got, want), C


def test_ragged_limits():
    rng = np.random.default_rng(0)
    # regions beyond 4x take the any-scale kernel, one launch each, inside the same call
    big = _ragged_images(rng, [(64, 64), (200, 300), (90, 77), (500, 131), (33, 400)])
    spec = TransformSpec(resize=32, center_crop=32, mean=IMAGENET[0], std=IMAGENET[1])
    want = torch.stack([TO.transform_sample(im, resize=32, center_crop=32, mean=IMAGENET[0], std=IMAGENET[1]) for im in big])
    assert same(BatchTransform(spec, device="cuda").ragged(big), want)
    boxes = np.array([[0, 0, 64, 64], [10, 20, 180, 260], [5, 5, 80, 70], [100, 3, 390, 120], [1, 100, 30, 250]], np.int32)
    flip = np.array([0, 1, 1, 0, 1], np.uint8)
    spec = TransformSpec(random_resized_crop=24, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1])
    want = torch.stack([TO.transform_sample(im, resized_crop=tuple(int(v) for v in bx) + (24, 24), flip=bool(f), mean=IMAGENET[0],
                                            std=IMAGENET[1]) for im, bx, f in zip(big, boxes, flip)])
    assert same(BatchTransform(spec, device="cuda").ragged(big, params=(boxes, flip)), want)
    with pytest.raises(ValueError, match="different output sizes"):
        BatchTransform(TransformSpec(resize=32), device="cuda").ragged(_ragged_images(rng, [(64, 64), (64, 100)]))
    with pytest.raises(NotImplementedError):
        BatchTransform(TransformSpec(crop=16), device="cuda").ragged(_ragged_images(rng, [(64, 64)]))
    one = BatchTransform(TransformSpec(resize=(20, 30)), device="cuda").ragged(_ragged_images(rng, [(64, 64), (40, 100)]))
    assert tuple(one.shape) == (2, 3, 20, 30)


RING_CASES = [
    # B, H, W (W * 3 a multiple of 16), resized (h, w), crop (h, w), padding, dtype   -- all reach the ring kernel
    (3, 96, 128, (48, 64), (48, 64), 0, torch.float32),          # exact 2x: top byte plane only
    (2, 100, 160, (77, 123), (60, 100), 3, torch.float32),       # odd sizes, crop with constant padding
    (1, 64, 1024, (40, 600), (40, 600), 0, torch.bfloat16),      # three column tiles of 256, bf16
    (5, 120, 96, (30, 24), (28, 24), 0, torch.float32),          # exactly 4x: eight taps
    (2, 130, 112, (50, 40), (50, 33), 0, torch.float16),         # 2.6x / 2.8x: eight taps, narrow last tile
    (4, 48, 64, (70, 100), (64, 96), 0, torch.float32),          # upscaling
    (40, 32, 32, (24, 20), (24, 20), 0, torch.float32),          # many small images: chunks of a few rows
    (1, 700, 48, (400, 30), (400, 30), 0, torch.float32),        # tall: several 64-row chunks per image
]


@pytest.mark.parametrize("B,H,W,rs,crop,pad,dtype", RING_CASES)
def test_ring_kernel_vs_integer_oracle(B, H, W, rs, crop, pad, dtype):
    """The bulk-copy staged (ring) resize kernel, every group count, against the integer restatement of Pillow."""
    from oracle import resize_oracle as RO
    from torchbooster_b200 import _native as N
    rng = np.random.default_rng(B * 1000 + H + W)
    batch = rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    th, tw = crop
    top = rng.integers(0, rs[0] + 2 * pad - th + 1, B).astype(np.int32)
    left = rng.integers(0, rs[1] + 2 * pad - tw + 1, B).astype(np.int32)
    flip = rng.integers(0, 2, B).astype(np.uint8)
    mean, std = IMAGENET
    x = torch.arange(256, dtype=torch.uint8).to(torch.float32).div(255)
    lut = ((x[None, :] - torch.tensor(mean)[:, None]) / torch.tensor(std)[:, None]).to(dtype)
    view = torch.int32 if dtype == torch.float32 else torch.int16
    want = np.stack([TO.index_map_sample(RO.resize_u8(batch[b], rs[0], rs[1]), lut.view(view).numpy(), top=int(top[b]),
                                         left=int(left[b]), out_h=th, out_w=tw, flip=bool(flip[b]), pad_y=pad, pad_x=pad,
                                         pad_mode="constant") for b in range(B)])
    spec = TransformSpec(resize=rs, crop=crop, padding=pad, flip_p=.5, mean=mean, std=std, dtype=dtype)
    before = N.lib().tb_launch_count()
    for flags in (0, 1 << 8, 2 << 8, 3 << 8, N.FLAG_NO_PDL, N.FLAG_SLIDE):
        bt = BatchTransform(spec, device="cuda")
        bt.flags = flags
        got = bt(torch.from_numpy(batch).cuda(), params=(top, left, flip)).cpu().view(view).numpy()
        assert np.array_equal(got, want), flags
    assert N.lib().tb_launch_count() - before >= 6
