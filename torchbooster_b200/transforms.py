"""Batch-level replacement of the reference's per-sample torchvision pipelines.

The reference's examples build ``torchvision.transforms`` chains that run per sample on PIL
images inside DataLoader workers, collate to fp32 NCHW and copy to the GPU
(``examples/img_cls/resnet/resnet.py:94-107``, ``examples/img_stt/adain/adain.py:167-171``,
``torchbooster/config.py:371-379,174-175``).  ``BatchTransform`` takes the raw uint8 NHWC batch
instead and produces the same tensor with one CUDA launch through the C ABI
(``include/torchbooster_b200.h``).  PyTorch only owns the buffers.

Random parameters are *inputs*: pass ``params=(top, left, flip)`` to reproduce any given draw, or
let ``CropFlipParams`` replay torch's CPU generator so that the draws equal what
``T.RandomCrop`` / ``T.RandomHorizontalFlip`` would have produced under ``torch.manual_seed``.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple, Union

import numpy as np
import torch

from . import _native as N
from . import _torchrng
from . import augment as _augment

IntPair = Union[int, Tuple[int, int]]


def _pair(v) -> Tuple[int, int]:
    if isinstance(v, (tuple, list)):
        if len(v) == 1:
            return int(v[0]), int(v[0])
        return int(v[0]), int(v[1])
    return int(v), int(v)


@dataclass(frozen=True)
class TransformSpec:
    """What the pipeline does, in the reference's vocabulary.

    Order: ``Resize`` -> ``CenterCrop`` -> ``RandomCrop(size, padding, padding_mode)`` ->
    ``RandomHorizontalFlip(p)`` -> ``RandomRotation(rotation)`` -> ``RandAugment(*rand_augment)`` ->
    ``ToTensor`` -> ``Normalize(mean, std)`` -> ``1 - X`` -> cast.

    resize        ``T.Resize`` argument: int (short side) or (h, w); Pillow antialiased bilinear.
    center_crop   ``T.CenterCrop`` size (zero fill when larger than the image).
    crop          ``T.RandomCrop`` size; offsets come from ``params`` (0 when absent).
    random_resized_crop  ``T.RandomResizedCrop(size, scale=rrc_scale, ratio=rrc_ratio)``: every sample's own
                  source box (from ``params``) is resized to ``size`` (Pillow bilinear); exclusive of
                  resize / center_crop / crop / padding.
    padding       ``T.RandomCrop`` padding: int or (left/right, top/bottom).
    padding_mode  "constant" (fill 0) | "reflect" | "edge" | "symmetric".
    flip_p        ``T.RandomHorizontalFlip`` probability; ``None`` = transform not in the pipeline.
    pre_rotation  ``T.RandomRotation`` at the HEAD of a random_resized_crop pipeline (``online.py:173``:
                  ``RandomRotation(180), RandomResizedCrop(size), RandAugment()``): the whole source is rotated.
    rotation      ``T.RandomRotation`` degrees (number or (min, max)); nearest, no expand, fill 0.
    rand_augment  ``T.RandAugment`` as ``(num_ops, magnitude, num_magnitude_bins)``, ``True`` = its defaults
                  ``(2, 9, 31)`` (``resnet.py:99-100``).  Both apply to crop / flip pipelines (one launch) and
                  after a random_resized_crop or a Resize / CenterCrop (the resize hands a uint8 image to the
                  augment kernel); the per-sample operations come from ``params`` or the replayed generator.
    mean, std     ``T.Normalize``; ``None`` = ToTensor only.
    one_minus     the ``1.0 - X`` of ``examples/img_gen/gan/gan.py:99``.
    dtype         output dtype: float32 (the reference), bfloat16 or float16 (RNE of the fp32 value).
    """
    crop: Optional[IntPair] = None
    random_resized_crop: Optional[IntPair] = None
    rrc_scale: Tuple[float, float] = (0.08, 1.0)
    rrc_ratio: Tuple[float, float] = (3.0 / 4.0, 4.0 / 3.0)
    padding: IntPair = 0
    padding_mode: str = "constant"
    flip_p: Optional[float] = None
    pre_rotation: Union[None, float, Tuple[float, float]] = None
    rotation: Union[None, float, Tuple[float, float]] = None
    rand_augment: Union[None, bool, Tuple[int, int, int]] = None
    resize: Optional[IntPair] = None
    center_crop: Optional[IntPair] = None
    mean: Optional[Sequence[float]] = None
    std: Optional[Sequence[float]] = None
    one_minus: bool = False
    dtype: torch.dtype = torch.float32

    def __post_init__(self):
        if self.padding_mode not in ("constant", "reflect", "edge", "symmetric"):
            # message of torchvision/transforms/_functional_pil.py:176-177
            raise ValueError("Padding mode should be either constant, edge, reflect or symmetric")
        if (self.mean is None) != (self.std is None):
            raise ValueError("mean and std must be given together")
        if self.random_resized_crop is not None and (self.resize is not None or self.center_crop is not None or
                                                     self.crop is not None or _pair(self.padding) != (0, 0)):
            raise ValueError("random_resized_crop excludes resize, center_crop, crop and padding")
        if self.pre_rotation is not None and self.random_resized_crop is None:
            raise NotImplementedError("pre_rotation is the head of a random_resized_crop pipeline (use rotation otherwise)")
        if self.rand_augment not in (None, True, False) and len(tuple(self.rand_augment)) != 3:
            raise ValueError("rand_augment is True or (num_ops, magnitude, num_magnitude_bins)")
        N.dtype_code(self.dtype)

    @property
    def two_stage(self) -> bool:
        """Resize / CenterCrop pipelines whose tail reads the resized image as a whole: RandomCrop padding that mirrors it
        (or pads a centre crop: the padding must not show the pixels the centre crop cut away) and rotation / RandAugment.
        The resize kernel hands a planar uint8 image to the augment kernel (two launches)."""
        if self.resize is None and self.center_crop is None:
            return False
        padded = _pair(self.padding) != (0, 0)
        return self.augments or (padded and (self.padding_mode != "constant" or self.center_crop is not None))

    @property
    def augments(self) -> bool:
        return self.rotation is not None or bool(self.rand_augment)

    @property
    def rand_augment_args(self) -> Optional[Tuple[int, int, int]]:
        if not self.rand_augment:
            return None
        return (2, 9, 31) if self.rand_augment is True else tuple(int(v) for v in self.rand_augment)


def resized_output_size(h: int, w: int, size: IntPair) -> Tuple[int, int]:
    """Output size of ``T.Resize(size)``: an int sets the short side, the long side becomes
    ``int(size * long / short)`` (torchvision ``functional.py:353-384``)."""
    if isinstance(size, (tuple, list)) and len(size) == 2:
        return int(size[0]), int(size[1])
    s = int(size[0]) if isinstance(size, (tuple, list)) else int(size)
    short, long = (w, h) if w <= h else (h, w)
    new_short, new_long = s, int(s * long / short)
    return (new_long, new_short) if w <= h else (new_short, new_long)


class CropFlipParams:
    """Vectorised replay of the draws ``T.RandomCrop.get_params`` and ``T.RandomHorizontalFlip``
    make from torch's global CPU generator after ``torch.manual_seed(seed)``
    (``torchvision/transforms/transforms.py:634-655,712-722``; ``torchbooster/utils.py:59-61``
    seeds that generator).

    torch's CPU generator is mt19937 seeded like ``numpy.random.RandomState(seed)``; per sample
    it spends one 32-bit word per ``randint`` (``word % range``) and one per ``rand``
    (``(word & (2**24 - 1)) * 2**-24``), in the order top, left, flip.  Nothing is drawn for the
    crop when the padded image already has the crop size.
    """

    def __init__(self, seed: int) -> None:
        self._rs = np.random.RandomState(int(seed) & 0xFFFFFFFF)

    def draw(self, batch: int, h: int, w: int, crop: Optional[Tuple[int, int]], padding: Tuple[int, int] = (0, 0),
             flip_p: Optional[float] = None):
        """``h, w``: unpadded source size.  Returns numpy ``(top int32[B], left int32[B], flip uint8[B])``."""
        px, py = padding
        ph, pw = h + 2 * py, w + 2 * px
        need_crop = crop is not None and not (ph == crop[0] and pw == crop[1])
        if crop is not None and (ph < crop[0] or pw < crop[1]):
            raise ValueError(f"Required crop size {tuple(crop)} is larger than input image size {(ph, pw)}")
        per = (2 if need_crop else 0) + (1 if flip_p is not None else 0)
        top = np.zeros(batch, np.int32)
        left = np.zeros(batch, np.int32)
        flip = np.zeros(batch, np.uint8)
        if per == 0 or batch == 0:
            return top, left, flip
        words = self._rs.randint(0, 2 ** 32, size=(batch, per), dtype=np.uint64)
        col = 0
        if need_crop:
            top = (words[:, 0] % np.uint64(ph - crop[0] + 1)).astype(np.int32)
            left = (words[:, 1] % np.uint64(pw - crop[1] + 1)).astype(np.int32)
            col = 2
        if flip_p is not None:
            u = (words[:, col] & np.uint64((1 << 24) - 1)).astype(np.float32) * np.float32(2.0 ** -24)
            flip = (u < np.float32(flip_p)).astype(np.uint8)
        return top, left, flip


class ResizedCropParams:
    """Replay of ``T.RandomResizedCrop.get_params`` (+ ``T.RandomHorizontalFlip``) draws from torch's
    global CPU generator after ``torch.manual_seed(seed)``
    (``torchvision/transforms/transforms.py`` RandomResizedCrop.get_params: up to ten attempts of
    ``area * uniform(scale)``, ``exp(uniform(log ratio))``, then two ``randint``; central fallback).

    ``torch.empty(1).uniform_(a, b)`` spends one 32-bit word: ``x * fl32(b - a) + a`` with ``x = (w & (2**24 - 1)) * 2**-24``,
    fused or not as torch's CPU kernel on this host does it (``_torchrng.uniform``).
    The number of words per sample varies with the rejection loop, so words are drawn in blocks
    and the unused tail is kept for the next batch; the per-word candidates (area, aspect) are
    computed vectorised, the accept / reject walk is sequential.
    """

    def __init__(self, seed: int, scale=(0.08, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0)) -> None:
        self._rs = np.random.RandomState(int(seed) & 0xFFFFFFFF)
        self.scale = (float(scale[0]), float(scale[1]))
        self.ratio = (float(ratio[0]), float(ratio[1]))
        log_ratio = torch.log(torch.tensor(self.ratio))            # float32, as torchvision computes it
        self._lr = (np.float32(log_ratio[0].item()), np.float32(log_ratio[1].item()))
        self._words = np.zeros(0, np.uint64)

    def _take(self, n: int) -> np.ndarray:
        if self._words.shape[0] < n:
            more = self._rs.randint(0, 2 ** 32, size=max(n - self._words.shape[0], 4096), dtype=np.uint64)
            self._words = np.concatenate([self._words, more])
        return self._words

    @staticmethod
    def _uniform(words: np.ndarray, a: np.float32, b: np.float32) -> np.ndarray:
        return _torchrng.uniform(words, float(a), float(b))

    def draw(self, batch: int, height: int, width: int, flip_p: Optional[float] = None):
        """Returns numpy ``(box int32[B,4] = top, left, h, w ; flip uint8[B])``."""
        area = height * width
        s0, s1 = np.float32(self.scale[0]), np.float32(self.scale[1])
        box = np.zeros((batch, 4), np.int32)
        flip = np.zeros(batch, np.uint8)
        if batch == 0:
            return box, flip
        need = batch * 6 + 64
        walk = N.lib().tb_rrc_params_walk
        while True:
            words = np.ascontiguousarray(self._take(need))
            # per-word candidates with torch's own float32 arithmetic (uniform_, exp)
            ta_all = area * self._uniform(words, s0, s1).astype(np.float64)
            ar_all = torch.exp(torch.from_numpy(self._uniform(words, self._lr[0], self._lr[1]))).numpy().astype(np.float64)
            used = walk(words.ctypes.data, ta_all.ctypes.data, ar_all.ctypes.data, words.shape[0], batch, height, width,
                        min(self.ratio), max(self.ratio), -1.0 if flip_p is None else float(flip_p),
                        box.ctypes.data, flip.ctypes.data)
            if used >= 0:
                break
            need = words.shape[0] * 2 + 23 * batch        # a long rejection streak: draw more and walk again
            if need > 64 * (23 * batch + 64):
                N.check(int(used))
        self._words = words[used:]
        return box, flip

    def draw_ragged(self, heights, widths, flip_p: Optional[float] = None):
        """``ResizedCropParams.draw`` for images of different sizes: sample ``b`` draws from its own
        ``heights[b] x widths[b]``.  Returns numpy ``(box int32[B,4], flip uint8[B])``."""
        heights = np.ascontiguousarray(np.asarray(heights, np.int32))
        widths = np.ascontiguousarray(np.asarray(widths, np.int32))
        batch = int(heights.shape[0])
        s0, s1 = np.float32(self.scale[0]), np.float32(self.scale[1])
        box = np.zeros((batch, 4), np.int32)
        flip = np.zeros(batch, np.uint8)
        if batch == 0:
            return box, flip
        need = batch * 6 + 64
        walk = N.lib().tb_rrc_params_walk_ragged
        while True:
            words = np.ascontiguousarray(self._take(need))
            u_all = self._uniform(words, s0, s1).astype(np.float64)
            ar_all = torch.exp(torch.from_numpy(self._uniform(words, self._lr[0], self._lr[1]))).numpy().astype(np.float64)
            used = walk(words.ctypes.data, u_all.ctypes.data, ar_all.ctypes.data, words.shape[0], batch, heights.ctypes.data,
                        widths.ctypes.data, min(self.ratio), max(self.ratio), -1.0 if flip_p is None else float(flip_p),
                        box.ctypes.data, flip.ctypes.data)
            if used >= 0:
                break
            need = words.shape[0] * 2 + 23 * batch
            if need > 64 * (23 * batch + 64):
                N.check(int(used))
        self._words = words[used:]
        return box, flip


class BatchTransform:
    """``uint8 [B,H,W,C]`` -> ``[B,C,h,w]`` on the GPU, bit-identical to the reference pipeline.

    ``batch`` may be a CUDA tensor (any batch / row stride, pixels dense) or a host tensor.  A
    PINNED host batch of a crop/flip pipeline is read in place by the kernel (``zero_copy``): its
    bulk copies go over PCIe and fetch only the crop rows, on a handful of SMs, so there is no
    staging copy and ~20 % fewer PCIe bytes than shipping whole images.  Pageable host batches and
    resize pipelines are copied to the device on the current stream first.
    """

    def __init__(self, spec: TransformSpec, device: Union[str, torch.device, None] = None, seed: Optional[int] = None,
                 zero_copy: bool = True):
        self.spec = spec
        self.zero_copy = zero_copy
        self.device = N.cuda_device(device)
        self._luts = {}
        self._taps = {}
        if seed is None:
            self._rng = None
        elif spec.random_resized_crop is not None and (spec.augments or spec.pre_rotation is not None):
            self._rng = _augment.ChainParams(seed, spec.rrc_scale, spec.rrc_ratio, spec.pre_rotation, spec.rotation,
                                             spec.rand_augment_args)
        elif spec.random_resized_crop is not None:
            self._rng = ResizedCropParams(seed, spec.rrc_scale, spec.rrc_ratio)
        elif spec.augments:
            self._rng = _augment.AugmentParams(seed, spec.rotation, spec.rand_augment_args)
        else:
            self._rng = CropFlipParams(seed)
        self._param_host = None
        self.flags = 0

    # -- tables ----------------------------------------------------------------------------
    def lut(self, channels: int) -> torch.Tensor:
        """Per-channel 256-entry table of ToTensor/Normalize/1-X/cast, built on the device."""
        t = self._luts.get(channels)
        if t is not None:
            return t
        sp = self.spec
        if sp.mean is not None:
            mean = [float(m) for m in sp.mean]
            std = [float(s) for s in sp.std]
            if len(mean) == 1:
                mean = mean * channels
            if len(std) == 1:
                std = std * channels
            if len(mean) != channels or len(std) != channels:
                raise ValueError(f"mean/std have {len(mean)}/{len(std)} entries for {channels} channels")
            kind = N.LUT_NORMALIZE_ONE_MINUS if sp.one_minus else N.LUT_NORMALIZE
            m_arr = (C.c_float * channels)(*mean)
            s_arr = (C.c_float * channels)(*std)
        else:
            kind = N.LUT_ONE_MINUS if sp.one_minus else N.LUT_TO_TENSOR
            m_arr = s_arr = None
        t = torch.empty((channels, 256), dtype=sp.dtype, device=self.device)
        with torch.cuda.device(self.device):
            N.check(N.lib().tb_build_lut(kind, m_arr, s_arr, channels, N.dtype_code(sp.dtype), t.data_ptr(),
                                         N.current_stream_ptr(self.device)))
        self._luts[channels] = t
        return t

    def _resize_taps(self, in_size: int, out_size: int):
        key = (in_size, out_size)
        hit = self._taps.get(key)
        if hit is not None:
            return hit
        L = N.lib()
        ks = L.tb_resize_ksize(in_size, out_size)
        if ks < 0:
            N.check(ks)
        bounds = np.empty((out_size, 2), np.int32)
        taps = np.empty((out_size, ks), np.int32)
        N.check(L.tb_resize_taps(in_size, out_size, bounds.ctypes.data, taps.ctypes.data))
        hit = (ks, torch.from_numpy(bounds).to(self.device), torch.from_numpy(taps).to(self.device))
        self._taps[key] = hit
        return hit

    # -- parameters ------------------------------------------------------------------------
    def geometry(self, h: int, w: int):
        """Static geometry for a source of ``h x w``: resized size, static offsets, output size,
        padding and whether random offsets apply."""
        sp = self.spec
        if sp.random_resized_crop is not None:
            out = _pair(sp.random_resized_crop)
            return dict(rs=out, off=(0, 0), cur=out, pad=(0, 0), crop=None, out=out)
        rs_h, rs_w = (resized_output_size(h, w, sp.resize) if sp.resize is not None else (h, w))
        cur_h, cur_w, off_y, off_x = rs_h, rs_w, 0, 0
        if sp.center_crop is not None:
            ch, cw = _pair(sp.center_crop)
            # torchvision functional.py:556-594: zero-pad when larger, then int(round(./2.0))
            pad_l = (cw - cur_w) // 2 if cw > cur_w else 0
            pad_t = (ch - cur_h) // 2 if ch > cur_h else 0
            pad_r = (cw - cur_w + 1) // 2 if cw > cur_w else 0
            pad_b = (ch - cur_h + 1) // 2 if ch > cur_h else 0
            ph, pw = cur_h + pad_t + pad_b, cur_w + pad_l + pad_r
            top = int(round((ph - ch) / 2.0))
            left = int(round((pw - cw) / 2.0))
            off_y, off_x = top - pad_t, left - pad_l
            cur_h, cur_w = ch, cw
        px, py = _pair(sp.padding)
        crop = _pair(sp.crop) if sp.crop is not None else None
        out_h, out_w = crop if crop is not None else (cur_h + 2 * py, cur_w + 2 * px)
        return dict(rs=(rs_h, rs_w), off=(off_y, off_x), cur=(cur_h, cur_w), pad=(px, py), crop=crop,
                    out=(out_h, out_w))

    def draw_params(self, batch: int, h: int, w: int, channels: int = 3):
        """Next ``(top, left, flip)`` -- ``(box, flip)`` for random_resized_crop, ``(top, left, flip, ops)``
        with rotation / rand_augment -- from the replayed torch generator (needs ``seed=``)."""
        if self._rng is None:
            raise RuntimeError("BatchTransform was built without a seed; pass params= explicitly")
        if isinstance(self._rng, _augment.ChainParams):
            oh, ow = _pair(self.spec.random_resized_crop)
            return self._rng.draw(batch, h, w, oh, ow, self.spec.flip_p)
        if self.spec.random_resized_crop is not None:
            return self._rng.draw(batch, h, w, self.spec.flip_p)
        g = self.geometry(h, w)
        if self.spec.augments:
            return self._rng.draw_records(batch, g["cur"][0], g["cur"][1], channels, g["crop"], g["pad"], self.spec.flip_p)
        return self._rng.draw(batch, g["cur"][0], g["cur"][1], g["crop"], g["pad"], self.spec.flip_p)

    def _params_to_device(self, params, batch: int):
        top, left, flip = params
        if all(isinstance(t, torch.Tensor) and t.is_cuda for t in (top, left, flip) if t is not None):
            conv = lambda t, dt: None if t is None else t.to(device=self.device, dtype=dt).contiguous()
            return conv(top, torch.int32), conv(left, torch.int32), conv(flip, torch.uint8)
        # one pinned staging buffer, one async copy: [top i32 x B][left i32 x B][flip u8 x B]
        host = torch.empty(9 * batch, dtype=torch.uint8, pin_memory=True)
        hv = host.numpy()
        hv[:4 * batch].view(np.int32)[:] = 0 if top is None else np.asarray(top, dtype=np.int32)
        hv[4 * batch:8 * batch].view(np.int32)[:] = 0 if left is None else np.asarray(left, dtype=np.int32)
        hv[8 * batch:] = 0 if flip is None else np.asarray(flip, dtype=np.uint8)
        dev = host.to(self.device, non_blocking=True)
        self._param_host = host  # keep the pinned source alive until the copy has run
        return dev[:4 * batch].view(torch.int32), dev[4 * batch:8 * batch].view(torch.int32), dev[8 * batch:]

    def _ops_to_device(self, ops, batch: int, h: int, w: int, channels: int) -> torch.Tensor:
        """(name, magnitude) lists, or already encoded int32 ``[B, n_ops, 8]`` records -> device tensor."""
        if isinstance(ops, torch.Tensor) and ops.is_cuda:
            rec = ops.to(device=self.device, dtype=torch.int32).contiguous()
        else:
            arr = ops if isinstance(ops, (np.ndarray, torch.Tensor)) else _augment.encode_ops(ops, h, w, channels)
            arr = np.ascontiguousarray(np.asarray(arr, dtype=np.int32))
            host = torch.from_numpy(arr).pin_memory()
            self._ops_hosts = (getattr(self, "_ops_hosts", []) + [host])[-4:]    # pinned sources stay alive until the copies have run
            rec = host.to(self.device, non_blocking=True)
        if rec.dim() != 3 or rec.shape[0] != batch or rec.shape[2] != _augment.OP_WORDS:
            raise ValueError(f"ops must be [B={batch}, n_ops, {_augment.OP_WORDS}] int32 records; got {tuple(rec.shape)}")
        return rec

    def _boxes_to_device(self, params, batch: int):
        box, flip = params
        if isinstance(box, torch.Tensor) and box.is_cuda:
            dbox = box.to(device=self.device, dtype=torch.int32).contiguous()
            dflip = None if flip is None else (flip if isinstance(flip, torch.Tensor) else torch.as_tensor(flip)).to(
                device=self.device, dtype=torch.uint8).contiguous()
            return dbox, dflip
        host = torch.empty(17 * batch, dtype=torch.uint8, pin_memory=True)     # [box i32 x 4B][flip u8 x B]
        hv = host.numpy()
        hv[:16 * batch].view(np.int32)[:] = np.asarray(box, dtype=np.int32).reshape(-1)
        hv[16 * batch:] = 0 if flip is None else np.asarray(flip, dtype=np.uint8)
        dev = host.to(self.device, non_blocking=True)
        self._param_host = host
        return dev[:16 * batch].view(torch.int32).view(batch, 4), dev[16 * batch:]

    def from_records(self, base_ptr: int, offsets, shape: Tuple[int, int, int], params=None,
                     out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Transform records that lie at ``base_ptr + offsets[b]`` (device-visible memory, e.g. an
        LMDB map registered with ``LMDBReader.cuda_register``): the kernel reads them in place.
        Crop / flip pipelines only."""
        sp = self.spec
        if sp.resize is not None or sp.center_crop is not None or sp.random_resized_crop is not None or sp.augments:
            raise NotImplementedError("from_records supports crop / flip / normalise pipelines")
        H, W, Cn = (int(v) for v in shape)
        off = np.ascontiguousarray(np.asarray(offsets, dtype=np.int64)) if not isinstance(offsets, torch.Tensor) else None
        B = int(off.shape[0]) if off is not None else int(offsets.shape[0])
        flags = self.flags
        if off is not None:
            if B and (int((off % 16).max()) != 0 or base_ptr % 16):
                flags |= N.FLAG_FORCE_GENERIC          # bulk copies need 16-byte aligned records
            host = torch.from_numpy(off)
            host = host.pin_memory() if torch.cuda.is_available() else host
            self._offsets_host = host
            offsets = host.to(self.device, non_blocking=True)
        g = self.geometry(H, W)
        out_h, out_w = g["out"]
        if params is None and self._rng is not None and (g["crop"] is not None or sp.flip_p is not None):
            params = self.draw_params(B, H, W)
        top = left = flip = None
        if params is not None:
            top, left, flip = self._params_to_device(params, B)
        if out is None:
            out = torch.empty((B, Cn, out_h, out_w), dtype=sp.dtype, device=self.device)
        lut = self.lut(Cn)
        px, py = g["pad"]
        ptr = lambda t: None if t is None else t.data_ptr()
        mode = N.PAD_MODES[sp.padding_mode] if (px or py) else N.PAD_MODES["none"]
        with torch.cuda.device(self.device):
            rc = N.lib().tb_transform_u8_records_to_nchw(
                base_ptr, offsets.data_ptr(), B, H, W, Cn, W * Cn, out.data_ptr(), N.dtype_code(sp.dtype), out_h, out_w,
                ptr(top), ptr(left), ptr(flip), py, px, mode, lut.data_ptr(), flags, N.current_stream_ptr(self.device))
        N.check(rc)
        return out

    def _identity_bytes(self, channels: int) -> torch.Tensor:
        if getattr(self, "_bytes_lut", None) is None or self._bytes_lut.shape[0] != channels:
            self._bytes_lut = torch.arange(256, dtype=torch.uint8, device=self.device).repeat(channels, 1).contiguous()
        return self._bytes_lut

    @staticmethod
    def _needs_records(batch: torch.Tensor, out_h: int, out_w: int) -> bool:
        """Frames the uniform resized-crop entry point refuses: more than 4x the output, or rows / images not 4-byte aligned."""
        _, H, W, _ = batch.shape
        return bool(H > 4 * out_h or W > 4 * out_w or batch.stride(1) % 4 or batch.stride(0) % 4 or batch.data_ptr() % 4)

    def _resized_crop_records(self, batch: torch.Tensor, box, flip, out_h: int, out_w: int, out: Optional[torch.Tensor],
                              as_bytes: bool = False) -> torch.Tensor:
        """The resized crop of a uniform batch through per-sample records (boxes read on the host)."""
        B, H, W, Cn = batch.shape
        hbox = (box.cpu().numpy() if isinstance(box, torch.Tensor) else np.asarray(box)).astype(np.int32).reshape(B, 4)
        hflip = None
        if flip is not None:
            hflip = (flip.cpu().numpy() if isinstance(flip, torch.Tensor) else np.asarray(flip)).astype(np.uint8)
        if batch.stride(1) % 4 or batch.stride(0) % 4 or batch.data_ptr() % 4:
            mem = torch.zeros((B, H, (W * Cn + 3) // 4 * 4), dtype=torch.uint8, device=self.device)
            mem[:, :, :W * Cn] = batch.reshape(B, H, W * Cn)
        else:
            mem = batch
        geom = np.zeros((B, 10), np.int32)
        geom[:, 0:4] = hbox
        geom[:, 4], geom[:, 5], geom[:, 8], geom[:, 9] = out_h, out_w, H, W
        return self._launch_ragged(mem, np.arange(B, dtype=np.int64) * mem.stride(0), np.full(B, mem.stride(1), np.int32), geom,
                                   hflip, Cn, out_h, out_w, out, as_bytes=as_bytes)

    def _resized_crop_chain(self, batch: torch.Tensor, params, out: Optional[torch.Tensor]) -> torch.Tensor:
        """[RandomRotation on the source ->] RandomResizedCrop [-> flip] [-> rotation / RandAugment on the crop] ->
        table: up to three launches that hand uint8 images on (``online.py:173``)."""
        sp = self.spec
        B, H, W, Cn = batch.shape
        out_h, out_w = _pair(sp.random_resized_crop)
        box_flip, ops, pre = params[:2], params[2], params[3]
        if (ops is None) != (not sp.augments) or (pre is None) != (sp.pre_rotation is None):
            raise ValueError("params do not match the pipeline: ops / pre must be given exactly for the stages it has")
        if out is None:
            out = torch.empty((B, Cn, out_h, out_w), dtype=sp.dtype, device=self.device)
        elif tuple(out.shape) != (B, Cn, out_h, out_w) or out.dtype != sp.dtype or not out.is_contiguous() or out.device != self.device:
            raise ValueError(f"out must be a contiguous tensor of shape {(B, Cn, out_h, out_w)}, dtype {sp.dtype} on {self.device}")
        if B == 0:
            return out
        L = N.lib()
        lut = self.lut(Cn)
        ptr = lambda t: None if t is None else t.data_ptr()
        with torch.cuda.device(self.device):
            stream = N.current_stream_ptr(self.device)
            src = batch
            if pre is not None:                                             # stage A: rotate the whole source image
                rec = self._ops_to_device(pre, B, H, W, Cn)
                if rec.shape[1] != 1:
                    raise ValueError("pre holds exactly one operation per sample")
                src = torch.empty((B, H, W, Cn), dtype=torch.uint8, device=self.device)
                N.check(L.tb_affine_nearest_u8_nhwc(batch.data_ptr(), B, H, W, Cn, batch.stride(0), batch.stride(1),
                                                    rec.data_ptr(), src.data_ptr(), stream))
            records = self._needs_records(src, out_h, out_w)
            if ops is None:                                                 # stage B straight to the output
                if records:
                    return self._resized_crop_records(src, box_flip[0], box_flip[1], out_h, out_w, out)
                box, flip = self._boxes_to_device(box_flip, B)
                N.check(L.tb_resized_crop_bilinear_u8_nhwc_to_nchw(
                    src.data_ptr(), B, H, W, Cn, src.stride(0), src.stride(1), box.data_ptr(), out.data_ptr(),
                    N.dtype_code(sp.dtype), out_h, out_w, ptr(flip), lut.data_ptr(), self.flags, stream))
                return out
            if records:                                                     # stage B: bytes, planar
                crop = self._resized_crop_records(src, box_flip[0], box_flip[1], out_h, out_w, None, as_bytes=True)
            else:
                box, flip = self._boxes_to_device(box_flip, B)
                crop = torch.empty((B, Cn, out_h, out_w), dtype=torch.uint8, device=self.device)
                N.check(L.tb_resized_crop_bilinear_u8_nhwc_to_nchw(
                    src.data_ptr(), B, H, W, Cn, src.stride(0), src.stride(1), box.data_ptr(), crop.data_ptr(),
                    N.TB_U8, out_h, out_w, ptr(flip), self._identity_bytes(Cn).data_ptr(), self.flags, stream))
            rec = self._ops_to_device(ops, B, out_h, out_w, Cn)             # stage C: operations on the crop, then the table
            N.check(L.tb_augment_u8_nhwc_to_nchw(
                crop.data_ptr(), B, out_h, out_w, Cn, Cn * out_h * out_w, out_w, out.data_ptr(), N.dtype_code(sp.dtype),
                out_h, out_w, None, None, None, 0, 0, N.PAD_MODES["none"], rec.data_ptr(), int(rec.shape[1]),
                lut.data_ptr(), N.FLAG_SRC_PLANAR, stream))
        return out

    def _resize_chain(self, batch: torch.Tensor, g, params, out: Optional[torch.Tensor]) -> torch.Tensor:
        """Resize [-> CenterCrop] -> RandomCrop(padding, any mode) -> flip [-> rotation / RandAugment] -> table
        (``TransformSpec.two_stage``): the resize kernel writes the resized (and centre-cropped) image as planar uint8,
        the augment kernel pads / crops / flips it, runs the operations and applies the table."""
        sp = self.spec
        B, H, W, Cn = batch.shape
        out_h, out_w = g["out"]
        cur_h, cur_w = g["cur"]
        if sp.augments:
            if params is None or len(params) != 4:
                raise ValueError("rotation / rand_augment need params=(top, left, flip, ops) or a seed")
            ops = params[3]
            params = params[:3]
        else:
            ops = None
        if out is None:
            out = torch.empty((B, Cn, out_h, out_w), dtype=sp.dtype, device=self.device)
        elif tuple(out.shape) != (B, Cn, out_h, out_w) or out.dtype != sp.dtype or not out.is_contiguous() or out.device != self.device:
            raise ValueError(f"out must be a contiguous tensor of shape {(B, Cn, out_h, out_w)}, dtype {sp.dtype} on {self.device}")
        if B == 0:
            return out
        top = left = flip = None
        if params is not None:
            top, left, flip = self._params_to_device(params, B)
        rec = None if ops is None else self._ops_to_device(ops, B, out_h, out_w, Cn)
        lut = self.lut(Cn)
        px, py = g["pad"]
        ptr = lambda t: None if t is None else t.data_ptr()
        L = N.lib()
        with torch.cuda.device(self.device):
            stream = N.current_stream_ptr(self.device)
            rs_h, rs_w = g["rs"]
            ksx, bx, kx = self._resize_taps(W, rs_w) if rs_w != W else (0, None, None)
            ksy, by, ky = self._resize_taps(H, rs_h) if rs_h != H else (0, None, None)
            off_y, off_x = g["off"]
            mid = torch.empty((B, Cn, cur_h, cur_w), dtype=torch.uint8, device=self.device)
            N.check(L.tb_resize_bilinear_u8_nhwc_to_nchw(
                batch.data_ptr(), B, H, W, Cn, batch.stride(0), batch.stride(1), rs_h, rs_w,
                ptr(bx), ptr(kx), ksx, ptr(by), ptr(ky), ksy, mid.data_ptr(), N.TB_U8, cur_h, cur_w,
                None, None, None, off_y, off_x, self._identity_bytes(Cn).data_ptr(), self.flags, stream))
            mode = N.PAD_MODES[sp.padding_mode] if (px or py) else N.PAD_MODES["none"]
            N.check(L.tb_augment_u8_nhwc_to_nchw(
                mid.data_ptr(), B, cur_h, cur_w, Cn, Cn * cur_h * cur_w, cur_w, out.data_ptr(), N.dtype_code(sp.dtype),
                out_h, out_w, ptr(top), ptr(left), ptr(flip), py, px, mode, ptr(rec), 0 if rec is None else int(rec.shape[1]),
                lut.data_ptr(), N.FLAG_SRC_PLANAR, stream))
        return out

    # -- ragged batches --------------------------------------------------------------------
    def _ragged_geometry(self, sizes: np.ndarray, params):
        """Per-sample records ``[top, left, h, w, resized h, resized w, y, x, image h, image w]`` for images of the
        given sizes under this spec (resize + center_crop, or random_resized_crop), the flips, and the output size."""
        sp = self.spec
        rrc = sp.random_resized_crop is not None
        B = int(sizes.shape[0])
        geom = np.zeros((B, 10), np.int32)
        geom[:, 8:10] = sizes
        flip = None
        if rrc:
            out_h, out_w = _pair(sp.random_resized_crop)
            if params is None:
                if self._rng is None:
                    raise ValueError("random_resized_crop needs params=(box[B,4], flip) or a seed")
                params = self._rng.draw_ragged(sizes[:, 0], sizes[:, 1], sp.flip_p)
            box, flip = params
            geom[:, 0:4] = np.asarray(box, np.int32).reshape(B, 4)
            geom[:, 4], geom[:, 5] = out_h, out_w
        else:
            fixed = None
            for b, (h, w) in enumerate(sizes):
                g = self.geometry(int(h), int(w))
                geom[b, 0:4] = (0, 0, h, w)
                geom[b, 4:6] = g["rs"]
                geom[b, 6:8] = g["off"]
                if fixed is None:
                    fixed = g["out"]
                elif fixed != g["out"]:
                    raise ValueError(f"images resize to different output sizes ({fixed} and {g['out']}); add center_crop")
            out_h, out_w = fixed
        return geom, flip, out_h, out_w

    def _check_ragged_spec(self) -> None:
        sp = self.spec
        if sp.augments or sp.pre_rotation is not None or sp.crop is not None or _pair(sp.padding) != (0, 0) or \
                (sp.random_resized_crop is None and sp.resize is None):
            raise NotImplementedError("ragged batches: resize (+ center_crop) or random_resized_crop (+ flip) pipelines")

    def ragged_records(self, buffer: torch.Tensor, offsets, params=None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """The same over image records as ``write_ragged_image_lmdb`` stores them (16-byte ``TBI1`` header + rows
        padded to 4 bytes), lying in one host buffer at ``offsets`` -- what ``LMDBReader.get_ragged_batch``
        returns: the buffer goes to the device as it is and the kernel reads every image where it lies."""
        from .lmdb import RAGGED_MAGIC
        self._check_ragged_spec()
        off = np.ascontiguousarray(np.asarray(offsets, np.int64))
        B = int(off.shape[0])
        if B == 0:
            raise ValueError("ragged_records() needs at least one record")
        hv = buffer.numpy()
        if int((off % 4).max()):
            raise ValueError("record offsets must be multiples of 4 bytes")
        head = np.stack([hv[o:o + 16].view("<u4") for o in off]).astype(np.int64)
        if (head[:, 0] != RAGGED_MAGIC).any():
            raise ValueError("a record does not start with the TBI1 image header (write_ragged_image_lmdb)")
        sizes, chans = head[:, 1:3], head[:, 3]
        Cn = int(chans[0])
        if (chans != Cn).any():
            raise ValueError("records hold different channel counts")
        geom, flip, out_h, out_w = self._ragged_geometry(sizes, params)
        pitches = ((sizes[:, 1] * Cn + 3) // 4 * 4).astype(np.int32)
        if not buffer.is_pinned():
            buffer = buffer.pin_memory()
        self._ragged_host = buffer
        dsrc = buffer.to(self.device, non_blocking=True)
        return self._launch_ragged(dsrc, off + 16, pitches, geom, flip, Cn, out_h, out_w, out)

    def ragged(self, images, params=None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """A batch of images of DIFFERENT sizes (what the reference's datasets really hold: paintings and
        COCO photographs under ``T.Resize(size) + T.CenterCrop(size)``, ``adain.py:170``; ImageNet-style
        ``T.RandomResizedCrop``) -> one ``[B,C,h,w]`` tensor, one launch.

        ``images``: sequence of ``uint8 [H_b, W_b, C]`` host arrays / tensors.  Supported pipelines:
        ``resize`` (+ ``center_crop``) with a common output size, and ``random_resized_crop`` (+ flip);
        ``params=(boxes[B,4], flip)`` or a seed for the latter.  Every sample's taps are computed in the
        kernel for its own (size -> resized size); a region more than four times its resized size gets its
        own launch of the any-scale kernel inside the same call.
        """
        self._check_ragged_spec()
        arrs = [np.ascontiguousarray(im.numpy() if isinstance(im, torch.Tensor) else np.asarray(im)) for im in images]
        B = len(arrs)
        if B == 0:
            raise ValueError("ragged() needs at least one image")
        Cn = arrs[0].shape[2] if arrs[0].ndim == 3 else 0
        for a in arrs:
            if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] != Cn:
                raise TypeError("images must be uint8 [H,W,C] arrays with the same channel count")
        sizes = np.array([a.shape[:2] for a in arrs], np.int64)
        geom, flip, out_h, out_w = self._ragged_geometry(sizes, params)
        # ---- pack: rows padded to 4 bytes, images to 16 -----------------------------------
        pitches = ((sizes[:, 1] * Cn + 3) // 4 * 4).astype(np.int32)
        nbytes = (sizes[:, 0] * pitches + 15) // 16 * 16
        offsets = np.concatenate([[0], np.cumsum(nbytes)[:-1]]).astype(np.int64)
        host = torch.empty(int(nbytes.sum()), dtype=torch.uint8, pin_memory=True)
        hv = host.numpy()
        for a, off, pitch in zip(arrs, offsets, pitches):
            h, w = a.shape[:2]
            hv[off:off + h * pitch].reshape(h, pitch)[:, :w * Cn] = a.reshape(h, w * Cn)
        self._ragged_host = host                            # the pinned source stays alive until the copy has run
        dsrc = host.to(self.device, non_blocking=True)
        return self._launch_ragged(dsrc, offsets, pitches, geom, flip, Cn, out_h, out_w, out)

    def _launch_ragged(self, dsrc: torch.Tensor, offsets: np.ndarray, pitches: np.ndarray, geom: np.ndarray, flip, Cn: int,
                       out_h: int, out_w: int, out: Optional[torch.Tensor], as_bytes: bool = False) -> torch.Tensor:
        """Per-sample geometry records over device memory ``dsrc``: one launch of the eight-tap kernel for
        every region within 4x of its resized size, one launch of the any-scale kernel for each of the rest."""
        sp = self.spec
        B = int(geom.shape[0])
        geom = geom.copy()
        # regions more than four times their resized size do not fit the eight-tap kernel: the common launch
        # skips them (an empty region writes zeros) and each gets a launch of the any-scale kernel below
        too_big = (geom[:, 2] > 4 * geom[:, 4]) | (geom[:, 3] > 4 * geom[:, 5])
        # a region that leaves its image stays with the common launch, which writes zeros for it
        inside = (geom[:, 0] >= 0) & (geom[:, 1] >= 0) & (geom[:, 2] >= 1) & (geom[:, 3] >= 1) & \
                 (geom[:, 0] + geom[:, 2] <= geom[:, 8]) & (geom[:, 1] + geom[:, 3] <= geom[:, 9])
        too_big &= inside
        slow = [(int(b), geom[b].copy()) for b in np.nonzero(too_big)[0]]
        geom[too_big, 2:4] = 0
        ok = ~too_big
        taps = 4 if ((geom[ok, 2] <= 2 * geom[ok, 4]) & (geom[ok, 3] <= 2 * geom[ok, 5])).all() else 8
        meta = torch.empty(B * (8 + 4 + 40 + 1), dtype=torch.uint8, pin_memory=True)       # offsets | pitches | geometry | flip
        mv = meta.numpy()
        mv[:8 * B].view(np.int64)[:] = offsets
        mv[8 * B:12 * B].view(np.int32)[:] = pitches
        mv[12 * B:52 * B].view(np.int32)[:] = geom.reshape(-1)
        mv[52 * B:] = 0 if flip is None else np.asarray(flip, np.uint8)
        self._ragged_meta = meta                            # pinned source stays alive until the copy has run
        dmeta = meta.to(self.device, non_blocking=True)
        odt = torch.uint8 if as_bytes else sp.dtype         # as_bytes: planar uint8 out, the hand-off to the augment kernel
        if out is None:
            out = torch.empty((B, Cn, out_h, out_w), dtype=odt, device=self.device)
        elif tuple(out.shape) != (B, Cn, out_h, out_w) or out.dtype != odt or not out.is_contiguous() or out.device != self.device:
            raise ValueError(f"out must be a contiguous tensor of shape {(B, Cn, out_h, out_w)}, dtype {odt} on {self.device}")
        lut = self._identity_bytes(Cn) if as_bytes else self.lut(Cn)
        code = N.TB_U8 if as_bytes else N.dtype_code(sp.dtype)
        base = dmeta.data_ptr()
        ptr = lambda t: None if t is None else t.data_ptr()
        with torch.cuda.device(self.device):
            stream = N.current_stream_ptr(self.device)
            N.check(N.lib().tb_resize_crop_ragged_u8_to_nchw(
                dsrc.data_ptr(), base, base + 8 * B, B, Cn, base + 12 * B, taps, out.data_ptr(), code,
                out_h, out_w, None if flip is None else base + 52 * B, lut.data_ptr(), self.flags, stream))
            for b, g in slow:           # any-scale kernel, host-built taps (cached per size pair), one sample per launch
                top, left, in_h, in_w, rs_h, rs_w, oy, ox = (int(v) for v in g[:8])
                ksx, bx, kx = self._resize_taps(in_w, rs_w) if rs_w != in_w else (0, None, None)
                ksy, by, ky = self._resize_taps(in_h, rs_h) if rs_h != in_h else (0, None, None)
                N.check(N.lib().tb_resize_bilinear_u8_nhwc_to_nchw(
                    dsrc.data_ptr() + int(offsets[b]) + top * int(pitches[b]) + left * Cn, 1, in_h, in_w, Cn, 0, int(pitches[b]),
                    rs_h, rs_w, ptr(bx), ptr(kx), ksx, ptr(by), ptr(ky), ksy, out[b].data_ptr(), code,
                    out_h, out_w, None, None, None if flip is None else int(base + 52 * B + b), oy, ox, lut.data_ptr(), self.flags,
                    stream))
        return out

    # -- the call --------------------------------------------------------------------------
    def __call__(self, batch: torch.Tensor, params=None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        if batch.dtype != torch.uint8 or batch.dim() != 4:
            raise TypeError(f"batch must be uint8 [B,H,W,C]; got {batch.dtype} {tuple(batch.shape)}")
        sp = self.spec
        if not batch.is_cuda:
            in_place = (self.zero_copy and batch.is_pinned() and sp.resize is None and sp.center_crop is None and
                        sp.random_resized_crop is None and not sp.augments)
            if not in_place:
                batch = batch.to(self.device, non_blocking=True)
        elif batch.device != self.device:
            raise ValueError(f"batch is on {batch.device}, transform on {self.device}")
        B, H, W, Cn = batch.shape
        if batch.stride(3) != 1 or batch.stride(2) != Cn:
            batch = batch.contiguous()
        g = self.geometry(H, W)
        out_h, out_w = g["out"]
        rrc = sp.random_resized_crop is not None
        if params is None and self._rng is not None and (rrc or sp.augments or g["crop"] is not None or sp.flip_p is not None):
            params = self.draw_params(B, H, W, Cn)
        top = left = flip = box = ops = None
        if rrc and (sp.augments or sp.pre_rotation is not None):
            if params is None or len(params) != 4:
                raise ValueError("this pipeline needs params=(box[B,4], flip, ops, pre) or a seed")
            return self._resized_crop_chain(batch, params, out)
        if rrc and self._needs_records(batch, out_h, out_w):
            if params is None:
                raise ValueError("random_resized_crop needs params=(box[B,4], flip) or a seed")
            return self._resized_crop_records(batch, params[0], params[1], out_h, out_w, out)
        if rrc:
            if params is None:
                raise ValueError("random_resized_crop needs params=(box[B,4], flip) or a seed")
            box, flip = self._boxes_to_device(params, B)
        elif sp.two_stage:
            return self._resize_chain(batch, g, params, out)
        elif sp.augments:
            if params is None or len(params) != 4:
                raise ValueError("rotation / rand_augment need params=(top, left, flip, ops) or a seed")
            ops = self._ops_to_device(params[3], B, out_h, out_w, Cn)
            top, left, flip = self._params_to_device(params[:3], B)
        elif params is not None:
            top, left, flip = self._params_to_device(params, B)
        if out is None:
            out = torch.empty((B, Cn, out_h, out_w), dtype=sp.dtype, device=self.device)
        else:
            if tuple(out.shape) != (B, Cn, out_h, out_w) or out.dtype != sp.dtype or not out.is_contiguous() \
                    or out.device != self.device:
                raise ValueError("out must be a contiguous tensor of shape "
                                 f"{(B, Cn, out_h, out_w)}, dtype {sp.dtype} on {self.device}")
        lut = self.lut(Cn)
        px, py = g["pad"]
        ptr = lambda t: None if t is None else t.data_ptr()
        L = N.lib()
        with torch.cuda.device(self.device):
            stream = N.current_stream_ptr(self.device)
            if rrc:
                rc = L.tb_resized_crop_bilinear_u8_nhwc_to_nchw(
                    batch.data_ptr(), B, H, W, Cn, batch.stride(0), batch.stride(1), box.data_ptr(),
                    out.data_ptr(), N.dtype_code(sp.dtype), out_h, out_w, ptr(flip), lut.data_ptr(), self.flags, stream)
            elif ops is not None:
                mode = N.PAD_MODES[sp.padding_mode] if (px or py) else N.PAD_MODES["none"]
                rc = L.tb_augment_u8_nhwc_to_nchw(
                    batch.data_ptr(), B, H, W, Cn, batch.stride(0), batch.stride(1),
                    out.data_ptr(), N.dtype_code(sp.dtype), out_h, out_w,
                    ptr(top), ptr(left), ptr(flip), py, px, mode, ops.data_ptr(), int(ops.shape[1]),
                    lut.data_ptr(), self.flags, stream)
            elif sp.resize is None and g["off"] == (0, 0) and sp.center_crop is None:
                mode = N.PAD_MODES[sp.padding_mode] if (px or py) else N.PAD_MODES["none"]
                rc = L.tb_transform_u8_nhwc_to_nchw(
                    batch.data_ptr(), B, H, W, Cn, batch.stride(0), batch.stride(1),
                    out.data_ptr(), N.dtype_code(sp.dtype), out_h, out_w,
                    ptr(top), ptr(left), ptr(flip), py, px, mode, lut.data_ptr(), self.flags, stream)
            else:
                rs_h, rs_w = g["rs"]
                ksx, bx, kx = self._resize_taps(W, rs_w) if rs_w != W else (0, None, None)
                ksy, by, ky = self._resize_taps(H, rs_h) if rs_h != H else (0, None, None)
                off_y, off_x = g["off"]
                rc = L.tb_resize_bilinear_u8_nhwc_to_nchw(
                    batch.data_ptr(), B, H, W, Cn, batch.stride(0), batch.stride(1), rs_h, rs_w,
                    ptr(bx), ptr(kx), ksx, ptr(by), ptr(ky), ksy,
                    out.data_ptr(), N.dtype_code(sp.dtype), out_h, out_w,
                    ptr(top), ptr(left), ptr(flip), off_y - py, off_x - px, lut.data_ptr(), self.flags, stream)
        N.check(rc)
        return out


class ToImage:
    """The reference's ``utransform``: ``T.Compose([T.Normalize(mean, std), clamp | hdr, T.ToPILImage()])``
    (``examples/img_stt/adain/adain.py:168-171``, ``online/online.py:169-174``, ``offline/offline.py:108-111``)
    for a float32 CUDA tensor ``[C,H,W]`` or ``[B,C,H,W]`` -> uint8 ``[H,W,C]`` / ``[B,H,W,C]`` on the GPU,
    byte-identical to torchvision's chain.

    mode  ``None`` (Normalize + ToPILImage), ``"clamp"`` (``x.clamp(0, 1)``) or ``"stretch"``
          (``(x - x.min()) / (x.max() - x.min())`` over the whole tensor).
    mean, std are the ``T.Normalize`` arguments; ``ToImage.inverse_of(mean, std)`` builds the reference's
    ``mean=-mean / std, std=1.0 / std`` in float64 numpy as its examples do.
    """

    MODES = {None: 0, "plain": 0, "clamp": 1, "stretch": 2, "hdr": 2}

    def __init__(self, mean=None, std=None, mode: Optional[str] = "clamp"):
        if (mean is None) != (std is None):
            raise ValueError("mean and std must be given together")
        if mode not in self.MODES:
            raise ValueError(f"mode must be one of {sorted(k for k in self.MODES if k)} or None")
        self.mode = self.MODES[mode]
        # F.normalize: torch.as_tensor(mean, dtype=tensor.dtype) -- float32 of the given numbers
        self.mean = None if mean is None else np.atleast_1d(np.asarray(mean, dtype=np.float64)).astype(np.float32)
        self.std = None if std is None else np.atleast_1d(np.asarray(std, dtype=np.float64)).astype(np.float32)

    @classmethod
    def inverse_of(cls, mean, std, mode: Optional[str] = "clamp") -> "ToImage":
        mean, std = np.asarray(mean, dtype=np.float64), np.asarray(std, dtype=np.float64)
        return cls(-mean / std, 1.0 / std, mode)

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        N.require_cuda(x, "x")
        if x.dtype != torch.float32 or x.dim() not in (3, 4):
            raise TypeError(f"x must be float32 [C,H,W] or [B,C,H,W]; got {x.dtype} {tuple(x.shape)}")
        single = x.dim() == 3
        xb = (x.unsqueeze(0) if single else x).contiguous()
        B, Cn, H, W = xb.shape
        mean = std = None
        if self.mean is not None:
            mean = np.ascontiguousarray(np.broadcast_to(self.mean, (Cn,)) if self.mean.shape[0] == 1 else self.mean)
            std = np.ascontiguousarray(np.broadcast_to(self.std, (Cn,)) if self.std.shape[0] == 1 else self.std)
            if mean.shape[0] != Cn or std.shape[0] != Cn:
                raise ValueError(f"mean/std have {mean.shape[0]}/{std.shape[0]} entries for {Cn} channels")
        out = torch.empty((B, H, W, Cn), dtype=torch.uint8, device=x.device)
        L = N.lib()
        with torch.cuda.device(x.device):
            stream = N.current_stream_ptr(x.device)
            minmax = None
            if self.mode == 2:
                minmax = torch.tensor([[float("inf"), float("-inf")]] * Cn, dtype=torch.float32, device=x.device)
                N.check(L.tb_channel_minmax(xb.data_ptr(), B, Cn, H * W, minmax.data_ptr(), stream))
            N.check(L.tb_to_image_u8(xb.data_ptr(), B, Cn, H, W, None if mean is None else mean.ctypes.data,
                                     None if std is None else std.ctypes.data, self.mode,
                                     None if minmax is None else minmax.data_ptr(), out.data_ptr(), stream))
        return out[0] if single else out

    @staticmethod
    def to_pil(image: torch.Tensor):
        """uint8 ``[H,W,C]`` tensor -> ``PIL.Image`` (what ``T.ToPILImage`` returns)."""
        from PIL import Image
        arr = image.cpu().numpy()
        return Image.fromarray(arr[:, :, 0] if arr.shape[2] == 1 else arr)
