"""The reference's `utransform` (inverse Normalize + clamp / stretch + ToPILImage) on the GPU:
byte-identical to torchvision's chain on the CPU."""
import numpy as np
import pytest
import torch

from oracle import transform_oracle as TO
from torchbooster_b200.transforms import BatchTransform, ToImage, TransformSpec

pytestmark = pytest.mark.gpu

MEAN, STD = np.array((0.485, 0.456, 0.406)), np.array((0.229, 0.224, 0.225))        # adain.py:167


@pytest.mark.parametrize("mode", ["clamp", "stretch", None])
@pytest.mark.parametrize("shape", [(3, 64, 48), (3, 256, 256), (1, 33, 17), (3, 1, 1), (3, 517, 389)])
def test_utransform_equals_torchvision(mode, shape):
    g = torch.Generator().manual_seed(hash((mode, shape)) % 1000)
    x = torch.randn(shape, generator=g) * 1.3
    if mode is None:
        # without clamp / stretch ToPILImage is only defined on [0, 1]: a normalised real image
        x = (torch.rand(shape, generator=g) - torch.tensor(MEAN[:shape[0]], dtype=torch.float32)[:, None, None]) / \
            torch.tensor(STD[:shape[0]], dtype=torch.float32)[:, None, None]
        x = x.clamp(min=((0.001 - MEAN[:shape[0]]) / STD[:shape[0]]).max(), max=((0.998 - MEAN[:shape[0]]) / STD[:shape[0]]).min())
    m, s = MEAN[:shape[0]], STD[:shape[0]]
    ti = ToImage.inverse_of(m, s, mode)                                   # the reference's -mean / std, 1.0 / std
    got = ti(x.cuda())
    want = TO.to_image(x, mean=-m / s, std=1.0 / s, mode=mode)
    assert got.dtype == torch.uint8 and tuple(got.shape) == (shape[1], shape[2], shape[0])
    assert np.array_equal(got.cpu().numpy(), want)


def test_round_trip_and_batches():
    """ctransform then utransform returns the image (adain.py:170-171), batched; and a batch equals
    its images one by one for the clamp, the whole-tensor extrema for the stretch."""
    rng = np.random.default_rng(1)
    raw = rng.integers(0, 256, (5, 40, 56, 3), dtype=np.uint8)
    x = BatchTransform(TransformSpec(mean=MEAN, std=STD), device="cuda")(torch.from_numpy(raw).cuda())
    back = ToImage.inverse_of(MEAN, STD, "clamp")(x)
    want = np.stack([TO.to_image(x[b].cpu(), mean=-MEAN / STD, std=1.0 / STD, mode="clamp") for b in range(5)])
    assert np.array_equal(back.cpu().numpy(), want)
    assert np.abs(back.cpu().numpy().astype(int) - raw.astype(int)).max() <= 1          # float round trip: off by one at most
    st = ToImage.inverse_of(MEAN, STD, "stretch")(x)
    flat = x.cpu().permute(1, 0, 2, 3).reshape(3, 5 * 40, 56)                           # one tall image = make_grid(nrow=1) without padding
    want = TO.to_image(flat, mean=-MEAN / STD, std=1.0 / STD, mode="stretch").reshape(5, 40, 56, 3)
    assert np.array_equal(st.cpu().numpy(), want)
    assert ToImage.to_pil(back[0]).size == (56, 40)


def signed_zero_cases():
    """Tensors whose extrema involve -0.0 (its bit pattern is INT_MIN: an atomic on the signed view of a float
    must not treat it as 'non-negative')."""
    g = torch.Generator().manual_seed(11)
    # (a) the maximum is exactly -0.0, everything else negative: hi must be -0.0, not a value below it
    a = -torch.rand((3, 64, 96), generator=g) - 0.25
    a[:, 10, 32:64] = -0.0
    # (b) one whole warp's share is -0.0 while other warps hold the (negative) minimum and a positive maximum
    b = torch.randn((3, 128, 128), generator=g)
    b[:, 40:44, :] = -0.0
    # (c) only zeros of both signs and negatives; single channel
    c = -torch.rand((1, 50, 70), generator=g)
    c[0, ::3] = -0.0
    c[0, 1, 1] = 0.0
    return [a, b, c]


@pytest.mark.parametrize("case", [0, 1, 2])
def test_stretch_with_negative_zero_extrema(case):
    x = signed_zero_cases()[case]
    want = TO.to_image(x, mode="stretch")
    ti = ToImage(mode="stretch")
    for _ in range(8):                                       # the old fault depended on the order of the atomics
        assert np.array_equal(ti(x.cuda()).cpu().numpy(), want)
    n = x.shape[0]
    m, s = MEAN[:n], STD[:n]
    want = TO.to_image(x, mean=-m / s, std=1.0 / s, mode="stretch")
    assert np.array_equal(ToImage.inverse_of(m, s, "stretch")(x.cuda()).cpu().numpy(), want)


def test_constant_input_and_errors():
    x = torch.full((3, 8, 8), 0.25)
    got = ToImage(mode="stretch")(x.cuda())                  # 0 / 0 = NaN -> byte 0, as on the CPU
    assert np.array_equal(got.cpu().numpy(), TO.to_image(x, mode="stretch"))
    with pytest.raises(TypeError):
        ToImage()(torch.zeros((3, 4, 4), dtype=torch.float16, device="cuda"))
    with pytest.raises(ValueError, match="division by zero"):
        ToImage([0.0] * 3, [1.0, 0.0, 1.0])(torch.zeros((3, 4, 4), device="cuda"))
    with pytest.raises(ValueError, match="PIL mode"):
        ToImage()(torch.zeros((2, 4, 4), device="cuda"))
    with pytest.raises(ValueError):
        ToImage(mean=[0.5])
