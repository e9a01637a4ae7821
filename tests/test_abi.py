"""The C-ABI boundary on a machine without a GPU: the library loads, exports every symbol the
header declares, its host-side functions are right, and argument errors are reported cleanly."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

from oracle import resize_oracle as RO
from torchbooster_b200 import _native as N

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "torchbooster_b200.h"


def declared_symbols():
    text = HEADER.read_text()
    return sorted(set(re.findall(r"^TB_API\s+[\w\s\*]+?\b(tb_\w+)\s*\(", text, flags=re.M)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 10
    handle = ctypes.CDLL(str(N.lib_path()))
    for name in names:
        assert hasattr(handle, name), f"{name} declared in {HEADER.name} but not exported"
    assert sorted(N.SIGNATURES) == names, "ctypes SIGNATURES and the header disagree"


def test_version_and_error_text():
    L = N.lib()
    assert L.tb_version() == 1
    assert L.tb_resize_ksize(0, 5) == N.TB_EINVAL
    assert "sizes" in N.last_error()


@pytest.mark.parametrize("n_in,n_out", [(512, 256), (375, 256), (500, 341), (480, 512), (200, 256), (1000, 64), (7, 7), (5, 1)])
def test_resize_taps_equal_oracle(n_in, n_out):
    L = N.lib()
    ks = L.tb_resize_ksize(n_in, n_out)
    ks_o, bounds_o, taps_o = RO.bilinear_taps(n_in, n_out)
    assert ks == ks_o
    bounds = np.empty((n_out, 2), np.int32)
    taps = np.empty((n_out, ks), np.int32)
    assert L.tb_resize_taps(n_in, n_out, bounds.ctypes.data, taps.ctypes.data) == 0
    assert np.array_equal(bounds, bounds_o) and np.array_equal(taps, taps_o)


def test_argument_validation_needs_no_gpu():
    L = N.lib()
    one = ctypes.c_void_p(16)
    # C out of range, unknown dtype, reflect padding too large, bad pitch
    assert L.tb_transform_u8_nhwc_to_nchw(one, 1, 8, 8, 5, 192, 24, one, 0, 8, 8, None, None, None, 0, 0, 0, one, 0, None) == N.TB_EINVAL
    assert L.tb_transform_u8_nhwc_to_nchw(one, 1, 8, 8, 3, 192, 24, one, 9, 8, 8, None, None, None, 0, 0, 0, one, 0, None) == N.TB_EINVAL
    assert L.tb_transform_u8_nhwc_to_nchw(one, 1, 8, 8, 3, 192, 24, one, 0, 8, 8, None, None, None, 8, 0, 2, one, 0, None) == N.TB_EINVAL
    assert L.tb_transform_u8_nhwc_to_nchw(one, 1, 8, 8, 3, 192, 23, one, 0, 8, 8, None, None, None, 0, 0, 0, one, 0, None) == N.TB_EINVAL
    assert "src_row_pitch" in N.last_error()
    assert L.tb_metrics_update(one, 0, 4, 0, 0, one, 1, one, None, None) == N.TB_EINVAL
    assert L.tb_metrics_update(one, 0, 4, 10, 10, one, 0, one, None, None) == N.TB_EINVAL
    assert L.tb_metrics_update(one, 0, 4, 10, 9, one, 1, one, None, None) == N.TB_EINVAL
    assert L.tb_resize_bilinear_u8_nhwc_to_nchw(one, 1, 8, 8, 4, 256, 32, 4, 4, None, None, 0, None, None, 0, one, 0, 4, 4,
                                                None, None, None, 0, 0, one, 0, None) == N.TB_EINVAL
    assert "RGBA" in N.last_error()
    assert L.tb_build_lut(1, None, None, 3, 0, one, None) == N.TB_EINVAL
    std0 = (ctypes.c_float * 3)(0.5, 0.0, 0.5)
    mean = (ctypes.c_float * 3)(0.5, 0.5, 0.5)
    assert L.tb_build_lut(1, mean, std0, 3, 0, one, None) == N.TB_EINVAL
    assert "division by zero" in N.last_error()      # torchvision's message for a zero std
    with pytest.raises(ValueError):
        N.check(N.TB_EINVAL)
    # empty inputs are no-ops that never touch the device
    assert L.tb_metrics_update(None, 0, 0, 10, 10, None, 1, one, None, None) == 0
    assert L.tb_running_sum_update(None, 0, 0, one, None) == 0


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(N, "_lib", None)
    monkeypatch.setattr(N, "_LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(N.NativeLibraryMissing, match="no CPU or PyTorch fallback"):
        N.lib()
