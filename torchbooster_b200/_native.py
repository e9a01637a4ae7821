"""ctypes binding of ``include/torchbooster_b200.h`` (the C-ABI boundary).

There is no CPU fallback: if the shared library is missing or does not load, every entry point
raises.  ctypes releases the GIL for the duration of each call.
"""
from __future__ import annotations

import ctypes as C
import threading
from pathlib import Path

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libtorchbooster_b200.so"
_lock = threading.Lock()
_lib = None

TB_OK, TB_EINVAL, TB_ENOTSUP, TB_ENOENT = 0, -22, -95, -2
TB_F32, TB_BF16, TB_F16, TB_U8 = 0, 1, 2, 3
PAD_MODES = {"none": 0, "constant": 1, "reflect": 2, "edge": 3, "symmetric": 4}
LUT_TO_TENSOR, LUT_NORMALIZE, LUT_ONE_MINUS, LUT_NORMALIZE_ONE_MINUS = 0, 1, 2, 3
FLAG_FORCE_GENERIC, FLAG_FORCE_STREAM, FLAG_NO_PDL, FLAG_SRC_PLANAR, FLAG_TWO_PHASE, FLAG_SLIDE = 1, 2, 4, 8, 16, 32

_i32, _i64, _vp = C.c_int, C.c_int64, C.c_void_p

# name -> (restype, argtypes); mirrors include/torchbooster_b200.h one to one
SIGNATURES = {
    "tb_version": (_i32, []),
    "tb_last_error": (C.c_char_p, []),
    "tb_launch_count": (_i64, []),
    "tb_build_lut": (_i32, [_i32, _vp, _vp, _i32, _i32, _vp, _vp]),
    "tb_transform_u8_nhwc_to_nchw": (_i32, [_vp, _i32, _i32, _i32, _i32, _i64, _i64, _vp, _i32, _i32, _i32,
                                            _vp, _vp, _vp, _i32, _i32, _i32, _vp, _i32, _vp]),
    "tb_transform_u8_records_to_nchw": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _i64, _vp, _i32, _i32, _i32,
                                               _vp, _vp, _vp, _i32, _i32, _i32, _vp, _i32, _vp]),
    "tb_resize_ksize": (_i32, [_i32, _i32]),
    "tb_resize_taps": (_i32, [_i32, _i32, _vp, _vp]),
    "tb_resize_bilinear_u8_nhwc_to_nchw": (_i32, [_vp, _i32, _i32, _i32, _i32, _i64, _i64, _i32, _i32,
                                                  _vp, _vp, _i32, _vp, _vp, _i32, _vp, _i32, _i32, _i32,
                                                  _vp, _vp, _vp, _i32, _i32, _vp, _i32, _vp]),
    "tb_resized_crop_bilinear_u8_nhwc_to_nchw": (_i32, [_vp, _i32, _i32, _i32, _i32, _i64, _i64, _vp, _vp, _i32, _i32, _i32,
                                                        _vp, _vp, _i32, _vp]),
    "tb_augment_u8_nhwc_to_nchw": (_i32, [_vp, _i32, _i32, _i32, _i32, _i64, _i64, _vp, _i32, _i32, _i32, _vp, _vp, _vp,
                                          _i32, _i32, _i32, _vp, _i32, _vp, _i32, _vp]),
    "tb_affine_nearest_u8_nhwc": (_i32, [_vp, _i32, _i32, _i32, _i32, _i64, _i64, _vp, _vp, _vp]),
    "tb_augment_rotation_records": (_i32, [_vp, _i64, _i32, _i32, _vp]),
    "tb_augment_params_walk": (_i64, [_vp, _i64, _i64, _i32, _i32, C.c_double, _i32, C.c_float, C.c_float, _i32, _i32, _i32,
                                      _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp]),
    "tb_channel_minmax": (_i32, [_vp, _i64, _i32, _i64, _vp, _vp]),
    "tb_to_image_u8": (_i32, [_vp, _i64, _i32, _i32, _i32, _vp, _vp, _i32, _vp, _vp, _vp]),
    "tb_to_image_stretch_u8_static": (_i32, [_vp, _i64, _i32, _i32, _i32, _vp, _vp, _vp, _vp]),
    "tb_resize_crop_ragged_u8_to_nchw": (_i32, [_vp, _vp, _vp, _i32, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _i32, _vp]),
    "tb_metrics_update": (_i32, [_vp, _i32, _i64, _i32, _i64, _vp, _i32, _vp, _vp, _vp]),
    "tb_running_sum_update": (_i32, [_vp, _i32, _i64, _vp, _vp]),
    "tb_metric_state_pack_i32": (_i32, [_vp, _i64, _vp, _vp]),
    "tb_metric_state_unpack_i32": (_i32, [_vp, _i64, _vp, _vp]),
    "tb_lmdb_open": (_i32, [C.c_char_p, C.POINTER(_vp)]),
    "tb_lmdb_close": (_i32, [_vp]),
    "tb_lmdb_entries": (_i64, [_vp]),
    "tb_lmdb_get": (_i32, [_vp, C.c_char_p, _i64, C.POINTER(_vp), C.POINTER(_i64)]),
    "tb_lmdb_gather": (_i32, [_vp, _vp, _i64, _i64, _vp, _i32]),
    "tb_lmdb_gather_ragged": (_i32, [_vp, _vp, _i64, _vp, _vp, _vp, _i64, _i32]),
    "tb_lmdb_cuda_register": (_i32, [_vp, C.POINTER(_vp)]),
    "tb_lmdb_cuda_unregister": (_i32, [_vp]),
    "tb_lmdb_record_offsets": (_i32, [_vp, _vp, _i64, _i64, _vp]),
    "tb_rrc_params_walk": (_i64, [_vp, _vp, _vp, _i64, _i64, _i32, _i32, C.c_double, C.c_double, C.c_double, _vp, _vp]),
    "tb_rrc_params_walk_ragged": (_i64, [_vp, _vp, _vp, _i64, _i64, _vp, _vp, C.c_double, C.c_double, C.c_double, _vp, _vp]),
}


class NativeLibraryMissing(ImportError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def lib():
    """The loaded library.  Raises ``NativeLibraryMissing`` -- never falls back."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not _LIB_PATH.exists():
                raise NativeLibraryMissing(
                    f"{_LIB_PATH} is missing: build it with `python -m torchbooster_b200.build` "
                    "(needs nvcc).  torchbooster_b200 has no CPU or PyTorch fallback.")
            try:
                handle = C.CDLL(str(_LIB_PATH))
            except OSError as e:  # pragma: no cover
                raise NativeLibraryMissing(f"cannot load {_LIB_PATH}: {e}") from e
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(handle, name)
                fn.restype, fn.argtypes = res, args
            _lib = handle
    return _lib


def last_error() -> str:
    msg = lib().tb_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc: int, what: str = "") -> None:
    """Translate a C-ABI return code: argument errors -> ValueError, the rest -> RuntimeError."""
    if rc == TB_OK:
        return
    msg = last_error() or what
    if rc == TB_EINVAL:
        raise ValueError(msg)
    if rc == TB_ENOTSUP:
        raise NotImplementedError(msg)
    raise RuntimeError(msg)


def dtype_code(dtype) -> int:
    import torch
    try:
        return {torch.float32: TB_F32, torch.bfloat16: TB_BF16, torch.float16: TB_F16}[dtype]
    except KeyError:
        raise ValueError(f"unsupported dtype {dtype}; use float32, bfloat16 or float16") from None


def require_cuda(t, name: str):
    if not t.is_cuda:
        raise RuntimeError(
            f"{name} must be a CUDA tensor: torchbooster_b200 runs this path on the GPU only "
            "(there is no CPU fallback)")
    return t


def current_stream_ptr(device) -> int:
    import torch
    return torch.cuda.current_stream(device).cuda_stream


def cuda_device(device=None, allow_cpu: bool = False):
    """Normalise to an indexed CUDA device (``cuda`` -> ``cuda:<current>``)."""
    import torch
    if device is None:
        return torch.device("cuda", torch.cuda.current_device())
    device = torch.device(device)
    if device.type == "cuda":
        return device if device.index is not None else torch.device("cuda", torch.cuda.current_device())
    if allow_cpu and device.type == "cpu":
        return device
    raise RuntimeError(f"{device} is not a CUDA device: torchbooster_b200 runs on the GPU only (no CPU fallback)")
