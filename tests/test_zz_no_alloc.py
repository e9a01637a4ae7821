"""tb_to_image_stretch_u8_static: the min-max stretch with its extrema in a fixed device array of the
library -- same bytes as tb_channel_minmax + tb_to_image_u8, nothing allocated.  (Named to be
collected last: the other GPU tests run before the fixed array is first touched.)"""
import ctypes as C

import numpy as np
import pytest
import torch

from torchbooster_b200 import _native as N
from torchbooster_b200.transforms import ToImage

pytestmark = pytest.mark.gpu

MEAN, STD = np.array((0.485, 0.456, 0.406)), np.array((0.229, 0.224, 0.225))


def _static(x: torch.Tensor, mean32: np.ndarray, std32: np.ndarray, out: torch.Tensor) -> int:
    B, Cn, H, W = x.shape
    with torch.cuda.device(x.device):
        return N.lib().tb_to_image_stretch_u8_static(x.data_ptr(), B, Cn, H, W, mean32.ctypes.data, std32.ctypes.data,
                                                     out.data_ptr(), N.current_stream_ptr(x.device))


def test_two_calls_in_a_row_equal_the_two_call_form():
    ti = ToImage.inverse_of(MEAN, STD, "stretch")
    g = torch.Generator().manual_seed(3)
    a = (torch.randn((4, 3, 96, 80), generator=g) * 2.0).cuda()
    b = (torch.randn((2, 3, 33, 57), generator=g) * 0.1 + 5.0).cuda()      # different extrema: the array must be reset
    want_a, want_b = ti(a), ti(b)
    out_a = torch.empty((4, 96, 80, 3), dtype=torch.uint8, device="cuda")
    out_b = torch.empty((2, 33, 57, 3), dtype=torch.uint8, device="cuda")
    N.check(_static(a, ti.mean, ti.std, out_a))            # first call: looks the array's address up
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    N.check(_static(b, ti.mean, ti.std, out_b))            # back to back on one stream
    N.check(_static(a, ti.mean, ti.std, out_a))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before      # no device memory taken by the calls
    assert torch.equal(out_a, want_a) and torch.equal(out_b, want_b)


def test_grey_and_limits():
    ti = ToImage([-0.5], [2.0], "stretch")
    x = torch.rand((1, 1, 40, 24), device="cuda")
    out = torch.empty((1, 40, 24, 1), dtype=torch.uint8, device="cuda")
    N.check(_static(x, ti.mean, ti.std, out))
    assert torch.equal(out, ti(x))
    five = torch.zeros((1, 5, 4, 4), device="cuda")
    rc = _static(five, np.zeros(5, np.float32), np.ones(5, np.float32), torch.empty((1, 4, 4, 5), dtype=torch.uint8, device="cuda"))
    assert rc == N.TB_EINVAL and "beyond" in N.last_error()
    empty = torch.zeros((0, 3, 4, 4), device="cuda")
    assert _static(empty, np.zeros(3, np.float32), np.ones(3, np.float32), torch.empty((0, 4, 4, 3), dtype=torch.uint8, device="cuda")) == N.TB_OK


def test_negative_zero_extrema_through_the_fixed_array():
    """Same bytes as the two-call form when -0.0 is an extremum (the atomics branch on the sign bit)."""
    from test_gpu_image import signed_zero_cases
    for x in signed_zero_cases():
        Cn, H, W = x.shape
        ti = ToImage(mode="stretch")
        xb = x.cuda()[None]
        want = ti(xb[0])
        out = torch.empty((1, H, W, Cn), dtype=torch.uint8, device="cuda")
        mean32, std32 = np.zeros(Cn, np.float32), np.ones(Cn, np.float32)
        for _ in range(4):
            N.check(_static(xb, mean32, std32, out))
            assert torch.equal(out[0], want)


def test_a_refused_call_enqueues_nothing():
    """Every argument rule is checked before the first launch: the launch counter of the library stays put."""
    ok = torch.zeros((1, 3, 4, 4), device="cuda")
    out = torch.empty((1, 4, 4, 3), dtype=torch.uint8, device="cuda")
    one, zero = np.ones(3, np.float32), np.zeros(3, np.float32)
    two = torch.zeros((1, 2, 4, 4), device="cuda")
    lib = N.lib()
    before = lib.tb_launch_count()
    assert _static(two, zero[:2], one[:2], torch.empty((1, 4, 4, 2), dtype=torch.uint8, device="cuda")) == N.TB_EINVAL
    assert "tb_to_image_stretch_u8_static" in N.last_error() and "PIL mode" in N.last_error()
    assert _static(ok, zero, zero, out) == N.TB_EINVAL and "division by zero" in N.last_error()
    assert _static(ok, zero, -one, out) == N.TB_EINVAL and "std > 0" in N.last_error()
    B, Cn, H, W = ok.shape
    assert lib.tb_to_image_stretch_u8_static(ok.data_ptr(), B, Cn, H, W, zero.ctypes.data, None, out.data_ptr(),
                                             N.current_stream_ptr(ok.device)) == N.TB_EINVAL
    assert lib.tb_to_image_stretch_u8_static(None, B, Cn, H, W, zero.ctypes.data, one.ctypes.data, out.data_ptr(),
                                             N.current_stream_ptr(ok.device)) == N.TB_EINVAL
    assert lib.tb_launch_count() == before
