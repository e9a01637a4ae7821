"""Every BASELINE.json configuration AT ITS BENCHMARKED SIZE against the oracle.

The kernels pick their tile geometry from the batch size (rows / warps per tile in the crop kernel, strip height
and column tiles in the resize kernels, rows per warp in the scoring kernel), so the geometry ``bench.py`` times
must itself be compared with torchvision / the rank rule, not only small batches: C1 256 x 32^2, C2 512 x 256^2 ->
224 bf16, C3 1024 x 64^2, C4 32 x 512^2 -> 256^2, C5 125 000 x 1000 (one rank's share of the 1 M rows) in fp32 and
bf16.  Inputs are the synthetic ones of SURVEY.md section 8(d) (numpy default_rng(1234), MT19937 replay seed 42)."""
import numpy as np
import pytest
import torch

from oracle import metrics_oracle as MO
from oracle import transform_oracle as TO
from torchbooster_b200 import _native as N
from torchbooster_b200 import metrics as M
from torchbooster_b200.transforms import BatchTransform, CropFlipParams, TransformSpec

pytestmark = pytest.mark.gpu

IMAGENET = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
CIFAR = ((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))


def same(a: torch.Tensor, b: torch.Tensor) -> bool:
    a, b = a.cpu().contiguous(), b.cpu().contiguous()
    if a.dtype != b.dtype or a.shape != b.shape:
        return False
    view = {torch.float32: torch.int32, torch.bfloat16: torch.int16, torch.float16: torch.int16}[a.dtype]
    return torch.equal(a.view(view), b.view(view))


def images(shape):
    return np.random.default_rng(1234).integers(0, 256, shape, dtype=np.uint8)


def test_c1_cifar_reflect_crop_flip_batch_256():
    """resnet.py:94-98: RandomCrop(32, padding=4, reflect) + flip + CIFAR Normalize, fp32, batch 256."""
    batch = images((256, 32, 32, 3))
    top, left, flip = CropFlipParams(42).draw(256, 32, 32, (32, 32), (4, 4), 0.5)
    want = TO.transform_batch(batch, crop_top=top, crop_left=left, out_size=(32, 32), flip=flip, padding=4,
                              padding_mode="reflect", mean=CIFAR[0], std=CIFAR[1])
    spec = TransformSpec(crop=32, padding=4, padding_mode="reflect", flip_p=.5, mean=CIFAR[0], std=CIFAR[1])
    for flags in (0, N.FLAG_FORCE_GENERIC, N.FLAG_FORCE_STREAM):
        bt = BatchTransform(spec, device="cuda")
        bt.flags = flags
        assert same(bt(torch.from_numpy(batch).cuda(), params=(top, left, flip)), want), flags


def test_c2_imagenet_crop_flip_bf16_batch_512():
    """The headline configuration: 512 x 256^2 -> crop 224, flip, ImageNet Normalize, bf16.  Device batch through the
    streaming kernel (the benchmarked launch), the gather kernel, without PDL, and a pinned host batch read in place."""
    B = 512
    batch = images((B, 256, 256, 3))
    top, left, flip = CropFlipParams(42).draw(B, 256, 256, (224, 224), (0, 0), 0.5)
    want = TO.transform_batch(batch, crop_top=top, crop_left=left, out_size=(224, 224), flip=flip,
                              mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    spec = TransformSpec(crop=224, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    dev = torch.from_numpy(batch).cuda()
    for flags in (0, N.FLAG_FORCE_GENERIC, N.FLAG_FORCE_STREAM | N.FLAG_NO_PDL):
        bt = BatchTransform(spec, device="cuda")
        bt.flags = flags
        assert same(bt(dev, params=(top, left, flip)), want), flags
    pinned = torch.from_numpy(batch).pin_memory()
    assert same(BatchTransform(spec, device="cuda", zero_copy=True)(pinned, params=(top, left, flip)), want)
    # fp32 output of the same batch (two-table variant of the streaming kernel)
    want32 = TO.transform_batch(batch[:128], crop_top=top, crop_left=left, out_size=(224, 224), flip=flip,
                                mean=IMAGENET[0], std=IMAGENET[1])
    spec32 = TransformSpec(crop=224, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1])
    got32 = BatchTransform(spec32, device="cuda")(dev, params=(top, left, flip))
    assert same(got32[:128], want32)


def test_c3_gan_flip_to_minus_one_one_batch_1024():
    """gan.py:99,135 style: flip + Normalize(.5, .5) -> [-1, 1] fp32, 1024 x 64^2."""
    B = 1024
    batch = images((B, 64, 64, 3))
    _, _, flip = CropFlipParams(42).draw(B, 64, 64, (64, 64), (0, 0), 0.5)
    zeros = np.zeros(B, np.int32)
    want = TO.transform_batch(batch, crop_top=zeros, crop_left=zeros, out_size=(64, 64), flip=flip, mean=[.5] * 3, std=[.5] * 3)
    spec = TransformSpec(flip_p=.5, mean=[.5] * 3, std=[.5] * 3)
    for flags in (0, N.FLAG_FORCE_GENERIC):
        bt = BatchTransform(spec, device="cuda")
        bt.flags = flags
        assert same(bt(torch.from_numpy(batch).cuda(), params=(zeros, zeros, flip)), want), flags


@pytest.mark.parametrize("B", [32, 256])
def test_c4_resize_512_to_256_normalise(B):
    """adain.py:167-171: Resize(256) + CenterCrop(256) + ToTensor + Normalize on 512^2 sources, batch 32 (BASELINE)
    and 256 (the other geometry the resize kernels choose): every resize kernel against Pillow."""
    batch = images((B, 512, 512, 3))
    want = TO.transform_batch(batch, resize=256, center_crop=256, mean=IMAGENET[0], std=IMAGENET[1])
    dev = torch.from_numpy(batch).cuda()
    for flags in (0, N.FLAG_SLIDE, N.FLAG_TWO_PHASE, N.FLAG_FORCE_GENERIC):
        bt = BatchTransform(TransformSpec(resize=256, center_crop=256, mean=IMAGENET[0], std=IMAGENET[1]), device="cuda")
        bt.flags = flags
        got = bt(dev)
        assert same(got, want), flags


def _score_chunks(logits: torch.Tensor, labels: torch.Tensor, k: int, C: int, rows: int = 5000):
    """The rank-rule oracle over a large matrix, a few thousand rows at a time (float64 numpy)."""
    c1 = ck = 0
    cm = np.zeros((C, C), np.int64)
    lab = labels.cpu()
    for lo in range(0, logits.shape[0], rows):
        part = logits[lo:lo + rows].float().cpu()
        rank = MO.label_rank(part, lab[lo:lo + rows])
        c1 += int((rank < 1).sum())
        ck += int((rank < k).sum())
        cm += MO.confusion(part, lab[lo:lo + rows], C)
    return c1, ck, cm


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_c5_one_rank_share_125k_rows(dtype):
    """125 000 x 1000 logits (one of 8 ranks' share of the 1 M rows): top-1, top-5 and the confusion matrix against
    the rank rule / argmax / bincount of the oracle on the SAME values; the bf16 copy is full of exact ties."""
    n, C = 125_000, 1000
    gen = torch.Generator().manual_seed(1234)
    logits = torch.randn(n, C, generator=gen).to(dtype)
    labels = torch.randint(0, C, (n,), generator=gen)
    labels[::3] = logits[::3].float().argmax(-1)
    st = M.MetricState(C, k=5, device="cuda")
    st.update(logits.cuda(), labels.cuda())
    res = st.compute()
    c1, c5, cm = _score_chunks(logits, labels, 5, C)
    assert res["n"] == n and res["invalid_labels"] == 0
    assert res["correct_1"] == c1 and res["correct_5"] == c5
    assert np.array_equal(res["confusion"].cpu().numpy(), cm)
    assert c1 == int((logits.float().argmax(-1) == labels).sum())         # the reference's own count (metrics.py:27)
    assert float(M.accuracy(logits.cuda(), labels.cuda())) == float(MO.accuracy(logits.float(), labels))


def test_c5_latency_case_1000_rows():
    """The secondary C5 case of SURVEY.md 8(d): N = 1000 rows (10^6 logit values)."""
    gen = torch.Generator().manual_seed(99)
    logits = torch.randn(1000, 1000, generator=gen)
    labels = torch.randint(0, 1000, (1000,), generator=gen)
    st = M.MetricState(1000, k=5, device="cuda")
    st.update(logits.cuda(), labels.cuda())
    res = st.compute()
    assert res["correct_1"] == MO.topk_correct(logits, labels, 1) and res["correct_5"] == MO.topk_correct(logits, labels, 5)
    assert np.array_equal(res["confusion"].cpu().numpy(), MO.confusion(logits, labels, 1000))
