#!/usr/bin/env python
"""Kernel-level timing of every BASELINE.json config (C1..C5) through the C ABI: CUDA events,
inputs rotated over more than L2 of distinct buffers, algorithmic bytes from SURVEY.md 8(d).
One JSON line per case on stdout.  Development tool; the headline line comes from bench.py."""
import argparse
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from torchbooster_b200 import _native as N                      # noqa: E402
from torchbooster_b200.metrics import MetricState                # noqa: E402
from torchbooster_b200.transforms import BatchTransform, CropFlipParams, TransformSpec  # noqa: E402

IMAGENET = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
CIFAR = ((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))
PEAK = 6428.4
try:
    PEAK = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
except Exception:
    pass
L2 = 126 << 20
CASE = ""          # --case: run only cases whose name contains this


def time_loop(fn, iters, warm=5):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def transform_case(name, B, H, W, spec_kw, alg_bytes, iters, flags=0, params=True):
    if CASE and CASE not in name:
        return
    dev = torch.device("cuda", 0)
    spec = TransformSpec(**spec_kw)
    bt = BatchTransform(spec, device=dev)
    bt.flags = flags
    nbuf = min(24, max(3, -(-4 * L2 // (B * H * W * 3))))
    gen = torch.Generator(device=dev).manual_seed(1234)
    srcs = [torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, device=dev, generator=gen) for _ in range(nbuf)]
    g = bt.geometry(H, W)
    prm = None
    if params:
        px, py = g["pad"]
        t, l, f = CropFlipParams(42).draw(B, g["cur"][0], g["cur"][1], g["crop"], g["pad"], spec.flip_p)
        prm = tuple(torch.from_numpy(a).to(dev) for a in (t, l, f))
    out = [torch.empty((B, 3) + g["out"], dtype=spec.dtype, device=dev) for _ in range(2)]
    ms = time_loop(lambda i: bt(srcs[i % nbuf], params=prm, out=out[i & 1]), iters)
    gbs = alg_bytes * B / (ms * 1e-3) / 1e9
    print(json.dumps({"case": name, "B": B, "ms": ms, "images_per_s": B / (ms * 1e-3), "alg_GBps": gbs,
                      "frac_measured": gbs / PEAK, "frac_nominal": gbs / 8000, "nbuf": nbuf, "flags": flags}), flush=True)


def metrics_case(name, n, c, dtype, iters):
    if CASE and CASE not in name:
        return
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev).manual_seed(1234)
    es = 4 if dtype == torch.float32 else 2
    nbuf = min(8, max(2, -(-4 * L2 // (n * c * es))))
    logits = [torch.randn(n, c, device=dev, generator=gen).to(dtype) for _ in range(nbuf)]
    labels = [torch.randint(0, c, (n,), device=dev, generator=gen) for _ in range(nbuf)]
    st = MetricState(c, k=5, device=dev)
    ms = time_loop(lambda i: st.update(logits[i % nbuf], labels[i % nbuf]), iters)
    gbs = n * (c * es + 8) / (ms * 1e-3) / 1e9
    print(json.dumps({"case": name, "N": n, "C": c, "ms": ms, "samples_per_s": n / (ms * 1e-3), "alg_GBps": gbs,
                      "frac_measured": gbs / PEAK, "frac_nominal": gbs / 8000, "nbuf": nbuf}), flush=True)


def two_stream_case(name, B, iters, nstreams):
    """C2 transform launched round-robin on `nstreams` streams through the C ABI (no Python
    wrapper): does the tail of launch i overlap the prologue of launch i+1?"""
    if CASE and CASE not in name:
        return
    dev = torch.device("cuda", 0)
    spec = TransformSpec(crop=224, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    bt = BatchTransform(spec, device=dev)
    lut = bt.lut(3)
    L = N.lib()
    nbuf = 6
    gen = torch.Generator(device=dev).manual_seed(1234)
    srcs = [torch.randint(0, 256, (B, 256, 256, 3), dtype=torch.uint8, device=dev, generator=gen) for _ in range(nbuf)]
    t, l, f = CropFlipParams(42).draw(B, 256, 256, (224, 224), (0, 0), .5)
    t, l, f = (torch.from_numpy(a).to(dev) for a in (t, l, f))
    outs = [torch.empty((B, 3, 224, 224), dtype=torch.bfloat16, device=dev) for _ in range(max(2, nstreams))]
    streams = [torch.cuda.Stream(dev) for _ in range(nstreams)]
    main = torch.cuda.current_stream(dev)

    def run(n):
        for st in streams:
            st.wait_stream(main)
        for i in range(n):
            s, o = srcs[i % nbuf], outs[i % len(outs)]
            rc = L.tb_transform_u8_nhwc_to_nchw(s.data_ptr(), B, 256, 256, 3, s.stride(0), s.stride(1), o.data_ptr(), 1, 224, 224,
                                                t.data_ptr(), l.data_ptr(), f.data_ptr(), 0, 0, 0, lut.data_ptr(), 0,
                                                streams[i % nstreams].cuda_stream)
            assert rc == 0
        for st in streams:
            main.wait_stream(st)

    run(10)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    run(iters)
    e1.record(main)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    gbs = 224 * 224 * 9 * B / (ms * 1e-3) / 1e9
    print(json.dumps({"case": name, "B": B, "ms": ms, "images_per_s": B / (ms * 1e-3), "alg_GBps": gbs,
                      "frac_measured": gbs / PEAK, "streams": nstreams}), flush=True)


def rrc_case(name, B, iters, flags=0):
    """T.RandomResizedCrop(224) + flip + normalise -> bf16 on 256x256 sources (per-sample boxes)."""
    if CASE and CASE not in name:
        return
    from torchbooster_b200.transforms import ResizedCropParams
    dev = torch.device("cuda", 0)
    spec = TransformSpec(random_resized_crop=224, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    bt = BatchTransform(spec, device=dev)
    bt.flags = flags
    nbuf = 6
    gen = torch.Generator(device=dev).manual_seed(1234)
    srcs = [torch.randint(0, 256, (B, 256, 256, 3), dtype=torch.uint8, device=dev, generator=gen) for _ in range(nbuf)]
    box, flip = ResizedCropParams(42).draw(B, 256, 256, .5)
    prm = (torch.from_numpy(box).to(dev), torch.from_numpy(flip).to(dev))
    out = [torch.empty((B, 3, 224, 224), dtype=torch.bfloat16, device=dev) for _ in range(2)]
    ms = time_loop(lambda i: bt(srcs[i % nbuf], params=prm, out=out[i & 1]), iters)
    alg = float((box[:, 2] * box[:, 3]).sum()) * 3 + B * 224 * 224 * 3 * 2
    gbs = alg / (ms * 1e-3) / 1e9
    print(json.dumps({"case": name, "B": B, "ms": ms, "images_per_s": B / (ms * 1e-3), "alg_GBps": gbs,
                      "frac_measured": gbs / PEAK, "alg_bytes_per_image": alg / B}), flush=True)


def ragged_case(name, B, iters):
    """T.Resize(256) + T.CenterCrop(256) + normalise over COCO-like images of different sizes (short side
    360-640, aspect 0.6-1.7), packed back to back on the device: kernel time of the ragged launch."""
    if CASE and CASE not in name:
        return
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(0)
    short = rng.integers(360, 641, B)
    aspect = rng.uniform(1.0, 1.7, B)
    land = rng.integers(0, 2, B).astype(bool)
    hs = np.where(land, short, (short * aspect).astype(np.int64)).astype(np.int64)
    ws = np.where(land, (short * aspect).astype(np.int64), short).astype(np.int64)
    bt = BatchTransform(TransformSpec(resize=256, center_crop=256, mean=IMAGENET[0], std=IMAGENET[1]), device=dev)
    geom = np.zeros((B, 10), np.int32)
    for b in range(B):
        g = bt.geometry(int(hs[b]), int(ws[b]))
        geom[b] = (0, 0, hs[b], ws[b], g["rs"][0], g["rs"][1], g["off"][0], g["off"][1], hs[b], ws[b])
    pitches = ((ws * 3 + 3) // 4 * 4).astype(np.int32)
    nbytes = (hs * pitches + 15) // 16 * 16
    offsets = np.concatenate([[0], np.cumsum(nbytes)[:-1]]).astype(np.int64)
    src = torch.randint(0, 256, (int(nbytes.sum()),), dtype=torch.uint8, device=dev)
    d_off, d_pitch, d_geom = (torch.from_numpy(a).to(dev) for a in (offsets, pitches, geom))
    out = torch.empty((B, 3, 256, 256), dtype=torch.float32, device=dev)
    lut = bt.lut(3)
    L = N.lib()
    stream = N.current_stream_ptr(dev)

    def run(i):
        N.check(L.tb_resize_crop_ragged_u8_to_nchw(src.data_ptr(), d_off.data_ptr(), d_pitch.data_ptr(), B, 3, d_geom.data_ptr(), 8,
                                                   out.data_ptr(), N.dtype_code(torch.float32), 256, 256, None, lut.data_ptr(), 0, stream))
    ms = time_loop(run, iters)
    alg = float((hs * ws * 3).sum()) + B * 256 * 256 * 12
    gbs = alg / (ms * 1e-3) / 1e9
    print(json.dumps({"case": name, "B": B, "ms": ms, "images_per_s": B / (ms * 1e-3), "alg_GBps": gbs, "frac_measured": gbs / PEAK,
                      "source_MB": float(nbytes.sum()) / 1e6, "note": "whole source images counted as algorithmic bytes"}), flush=True)


def augment_imagenet_case(name, B, iters, flags=0):
    """RandomCrop(224) of 256x256 + flip + RandAugment() + normalise -> bf16: each image is owned by a
    cluster of 2 CTAs (distributed shared memory)."""
    if CASE and CASE not in name:
        return
    dev = torch.device("cuda", 0)
    spec = TransformSpec(crop=224, flip_p=.5, rand_augment=True, mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    bt = BatchTransform(spec, device=dev, seed=42)
    bt.flags = flags
    nbuf = 6
    gen = torch.Generator(device=dev).manual_seed(1234)
    srcs = [torch.randint(0, 256, (B, 256, 256, 3), dtype=torch.uint8, device=dev, generator=gen) for _ in range(nbuf)]
    prm = tuple(torch.from_numpy(a).to(dev) for a in bt.draw_params(B, 256, 256))
    out = [torch.empty((B, 3, 224, 224), dtype=torch.bfloat16, device=dev) for _ in range(2)]
    ms = time_loop(lambda i: bt(srcs[i % nbuf], params=prm, out=out[i & 1]), iters)
    gbs = 224 * 224 * 9 * B / (ms * 1e-3) / 1e9
    print(json.dumps({"case": name, "B": B, "ms": ms, "images_per_s": B / (ms * 1e-3), "alg_GBps": gbs,
                      "frac_measured": gbs / PEAK}), flush=True)


def augment_case(name, B, iters, host_params=False, flags=0):
    """The reference's CIFAR training transform (resnet.py:96-103): RandomCrop(32, 4, reflect), flip,
    RandomRotation(15), RandAugment(), normalise -> fp32; records pre-encoded on the device
    (kernel time) or drawn + encoded + uploaded every step (host_params)."""
    if CASE and CASE not in name:
        return
    import time
    from torchbooster_b200 import augment as A
    dev = torch.device("cuda", 0)
    spec = TransformSpec(crop=32, padding=4, padding_mode="reflect", flip_p=.5, rotation=15, rand_augment=True,
                         mean=CIFAR[0], std=CIFAR[1])
    bt = BatchTransform(spec, device=dev, seed=42)
    bt.flags = flags
    nbuf = 24
    gen = torch.Generator(device=dev).manual_seed(1234)
    srcs = [torch.randint(0, 256, (B, 32, 32, 3), dtype=torch.uint8, device=dev, generator=gen) for _ in range(nbuf)]
    t, l, f, rec = bt.draw_params(B, 32, 32)
    prm = tuple(torch.from_numpy(a).to(dev) for a in (t, l, f, rec))
    out = [torch.empty((B, 3, 32, 32), dtype=torch.float32, device=dev) for _ in range(2)]
    if host_params:
        ms = time_loop(lambda i: bt(srcs[i % nbuf], out=out[i & 1]), iters)
        t0 = time.perf_counter()
        for _ in range(20):
            bt.draw_params(B, 32, 32)
        host_ms = (time.perf_counter() - t0) / 20 * 1e3
    else:
        ms, host_ms = time_loop(lambda i: bt(srcs[i % nbuf], params=prm, out=out[i & 1]), iters), None
    gbs = 15360 * B / (ms * 1e-3) / 1e9
    cpu = None
    if host_params and B <= 256:
        # the reference's own Compose (resnet.py:96-103), one process, on this box's CPU
        import torchvision.transforms as T
        from PIL import Image
        compose = T.Compose([T.RandomCrop(32, 4, padding_mode="reflect"), T.RandomHorizontalFlip(), T.RandomRotation(15),
                             T.RandAugment(), T.ToTensor(), T.Normalize(CIFAR[0], CIFAR[1], inplace=True)])
        pil = [Image.fromarray(a) for a in srcs[0].cpu().numpy()]
        torch.stack([compose(im) for im in pil[:64]])
        t0 = time.perf_counter()
        for _ in range(4):
            torch.stack([compose(im) for im in pil])
        cpu = 4 * len(pil) / (time.perf_counter() - t0)
    print(json.dumps({"case": name, "B": B, "ms": ms, "images_per_s": B / (ms * 1e-3), "alg_GBps": gbs,
                      "frac_measured": gbs / PEAK, "host_draw_encode_ms": host_ms, "cpu_torchvision_1proc_images_per_s": cpu}),
          flush=True)


def to_image_case(name, B, H, W, mode, iters):
    """utransform (adain.py:171 / online.py:174): fp32 NCHW -> uint8 NHWC; 12 B read (+ 12 B for the
    stretch's extrema pass) + 3 B written per pixel."""
    if CASE and CASE not in name:
        return
    from torchbooster_b200.transforms import ToImage
    dev = torch.device("cuda", 0)
    nbuf = max(3, -(-4 * L2 // (B * H * W * 12)))
    xs = [torch.randn((B, 3, H, W), device=dev) for _ in range(nbuf)]
    ti = ToImage.inverse_of(IMAGENET[0], IMAGENET[1], mode)
    ms = time_loop(lambda i: ti(xs[i % nbuf]), iters)
    alg = B * H * W * (15 + (12 if mode == "stretch" else 0))
    gbs = alg / (ms * 1e-3) / 1e9
    print(json.dumps({"case": name, "B": B, "ms": ms, "images_per_s": B / (ms * 1e-3), "alg_GBps": gbs,
                      "frac_measured": gbs / PEAK}), flush=True)


def loader_case(name, n_images, batch, threads, prefetch=2, in_place=False, cifar_aug=False):
    """LMDB (tmpfs) -> BatchLoader background gather into pinned slots -> zero-copy C2 transform
    (cifar_aug: 32x32 records through the reference's whole CIFAR training transform instead)."""
    if CASE and CASE not in name:
        return
    import os
    import shutil
    import tempfile
    import time
    from torchbooster_b200.lmdb import LMDBReader, write_image_lmdb
    from torchbooster_b200.loader import BatchLoader
    dev = torch.device("cuda", 0)
    root = tempfile.mkdtemp(dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        rng = np.random.default_rng(0)
        side = 32 if cifar_aug else 256
        imgs = rng.integers(0, 256, (n_images, side, side, 3), dtype=np.uint8)
        t0 = time.perf_counter()
        write_image_lmdb(os.path.join(root, "db"), imgs, labels=np.arange(n_images) % 1000)
        t_write = time.perf_counter() - t0
        del imgs
        reader = LMDBReader(os.path.join(root, "db"))
        spec = TransformSpec(crop=224, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
        if cifar_aug:
            spec = TransformSpec(crop=32, padding=4, padding_mode="reflect", flip_p=.5, rotation=15, rand_augment=True,
                                 mean=CIFAR[0], std=CIFAR[1])
        bt = BatchTransform(spec, device=dev, seed=1)
        labels = reader.read_labels()
        order = np.random.default_rng(1).permutation(n_images).tolist()
        best = None
        for _ in range(3):
            loader = BatchLoader(reader, (side, side, 3), batch, sampler=order, transform=bt, labels=labels, drop_last=True,
                                 prefetch=prefetch, threads=threads, in_place=in_place)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            cnt = 0
            for x, y in loader:
                cnt += x.shape[0]
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        print(json.dumps({"case": name, "images": cnt, "batch": batch, "threads": threads, "images_per_s": cnt / best,
                          "host_gather_GBps": cnt * side * side * 3 / best / 1e9, "lmdb_write_s": t_write}), flush=True)
    finally:
        shutil.rmtree(root, ignore_errors=True)


def calibrate():
    """Library reference points on the same buffers: a read-only reduction and a copy."""
    dev = torch.device("cuda", 0)
    bufs = [torch.randn(125_000, 1000, device=dev) for _ in range(2)]
    dst = torch.empty_like(bufs[0])
    nbytes = bufs[0].numel() * 4
    for name, fn, moved in (("calib_torch_sum_500MB", lambda i: bufs[i & 1].sum(), nbytes),
                            ("calib_torch_amax_rows_500MB", lambda i: bufs[i & 1].amax(dim=1), nbytes),
                            ("calib_torch_copy_500MB", lambda i: dst.copy_(bufs[i & 1]), 2 * nbytes)):
        ms = time_loop(fn, 50)
        print(json.dumps({"case": name, "ms": ms, "GBps": moved / (ms * 1e-3) / 1e9, "frac_measured": moved / (ms * 1e-3) / 1e9 / PEAK}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--case", default="")
    a = ap.parse_args()
    global CASE
    CASE = a.case
    want = lambda n: not a.only or any(n.startswith(x) for x in a.only.split(","))
    c2 = dict(crop=224, flip_p=.5, mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    if want("c2"):
        knob = lambda warps=0, rows=0, extra=0: extra | (warps << 16) | (rows << 24)
        cases = [("c2_default", 0),  ("c2_no_pdl", N.FLAG_NO_PDL), ("c2_w11_r8", knob(11, 8)), ("c2_w22_r4", knob(22, 4)),
                 ("c2_w10_r8", knob(10, 8)), ("c2_w8_r12", knob(8, 12)), ("c2_w14_r6", knob(14, 6)), ("c2_generic", 1)]
        if a.quick:
            cases = cases[:3]
        for label, flags in cases:
            transform_case(label, 512, 256, 256, c2, 224 * 224 * 9, 200, flags)
        transform_case("c2_default_B2048", 2048, 256, 256, c2, 224 * 224 * 9, 60, 0)
        transform_case("c2_fp32_out", 512, 256, 256, dict(c2, dtype=torch.float32), 224 * 224 * 15, 200)
        transform_case("c2_fp32_out_w10_r4", 512, 256, 256, dict(c2, dtype=torch.float32), 224 * 224 * 15, 200, knob(10, 4))
    if want("c1"):
        c1 = dict(crop=32, padding=4, padding_mode="reflect", flip_p=.5, mean=CIFAR[0], std=CIFAR[1])
        for bsz in (256, 2048):
            transform_case(f"c1_cifar_b{bsz}", bsz, 32, 32, c1, 15360, 300)
            transform_case(f"c1_cifar_b{bsz}_stream", bsz, 32, 32, c1, 15360, 300, flags=N.FLAG_FORCE_STREAM)
            transform_case(f"c1_cifar_b{bsz}_generic", bsz, 32, 32, c1, 15360, 300, flags=1)
    if want("c3"):
        c3 = dict(flip_p=.5, mean=[.5] * 3, std=[.5] * 3)
        transform_case("c3_gen_b1024", 1024, 64, 64, c3, 61440, 300)
        transform_case("c3_gen_b1024_generic", 1024, 64, 64, c3, 61440, 300, flags=1)
    if want("c4"):
        c4 = dict(resize=256, center_crop=256, mean=IMAGENET[0], std=IMAGENET[1])
        transform_case("c4_stt_b32", 32, 512, 512, c4, 1572864, 200, params=False)
        transform_case("c4_stt_b256", 256, 512, 512, c4, 1572864, 50, params=False)
        for bsz in (1, 4, 8, 16, 64, 128, 512):
            transform_case(f"c4_scan_b{bsz}", bsz, 512, 512, c4, 1572864, 200 if bsz < 100 else 50, params=False)
            transform_case(f"c4_scan_b{bsz}_slide", bsz, 512, 512, c4, 1572864, 200 if bsz < 100 else 50, flags=32, params=False)
        transform_case("c4_stt_b32_slide", 32, 512, 512, c4, 1572864, 200, flags=32, params=False)
        transform_case("c4_stt_b256_slide", 256, 512, 512, c4, 1572864, 50, flags=32, params=False)
        for ng in (1, 2, 3):
            transform_case(f"c4_stt_b32_ng{ng}", 32, 512, 512, c4, 1572864, 200, flags=ng << 8, params=False)
            transform_case(f"c4_stt_b256_ng{ng}", 256, 512, 512, c4, 1572864, 50, flags=ng << 8, params=False)
        transform_case("c4_stt_b256_bf16", 256, 512, 512, dict(c4, dtype=torch.bfloat16), 786432 + 393216, 50, params=False)
        transform_case("c4_stt_b32_two_phase", 32, 512, 512, c4, 1572864, 200, flags=16, params=False)
        transform_case("c4_stt_b256_two_phase", 256, 512, 512, c4, 1572864, 50, flags=16, params=False)
        if a.quick:
            for tr in (8, 16, 32):
                for tc in (2, 4, 8):
                    transform_case(f"c4_stt_b256_tr{tr}_tc{tc * 32}", 256, 512, 512, c4, 1572864, 50, flags=(tr << 24) | (tc << 16), params=False)
                    transform_case(f"c4_stt_b32_tr{tr}_tc{tc * 32}", 32, 512, 512, c4, 1572864, 200, flags=(tr << 24) | (tc << 16), params=False)

    if want("loader"):
        for th in (8, 16):
            loader_case(f"loader_c2_threads{th}", 4096, 512, th)
        loader_case("loader_c2_in_place", 4096, 512, 8, in_place=True)
        loader_case("loader_cifar_aug_b256", 50000, 256, 8, cifar_aug=True)
        loader_case("loader_cifar_aug_b2048", 50000, 2048, 8, cifar_aug=True)
    if want("ragged"):
        ragged_case("ragged_coco_like_b32", 32, 100)
        ragged_case("ragged_coco_like_b256", 256, 30)
    if want("rrc"):
        rrc_case("rrc_c2_b512", 512, 100)
        rrc_case("rrc_c2_b2048", 2048, 30)
        rrc_case("rrc_c2_b512_two_phase", 512, 100, flags=16)
        if a.quick:
            for tr in (16, 32, 64):
                rrc_case(f"rrc_c2_b512_tr{tr}", 512, 100, flags=tr << 24)
                rrc_case(f"rrc_c2_b2048_tr{tr}", 2048, 30, flags=tr << 24)
    if want("aug"):
        for bsz in (256, 2048, 16384):
            augment_case(f"aug_cifar_b{bsz}", bsz, 200 if bsz < 10000 else 50)
        if a.quick:
            for bsz in (256, 512, 2048):
                for th in (2, 3, 4):
                    augment_case(f"aug_cifar_b{bsz}_threads{64 << th}", bsz, 200, flags=th << 12)
        augment_case("aug_cifar_b256_again", 256, 400)      # the first case of a process runs before the clocks have ramped
        augment_case("aug_cifar_b512", 512, 400)
        augment_imagenet_case("aug_imagenet224_b512", 512, 30)
        for cl in (2, 4, 8):
            augment_imagenet_case(f"aug_imagenet224_b512_cluster{cl}", 512, 30, flags=cl << 8)
        augment_case("aug_cifar_b256_host_params", 256, 100, host_params=True)
        augment_case("aug_cifar_b2048_host_params", 2048, 30, host_params=True)
    if want("image"):
        to_image_case("to_image_clamp_b32_256", 32, 256, 256, "clamp", 200)
        to_image_case("to_image_clamp_b256_256", 256, 256, 256, "clamp", 100)
        to_image_case("to_image_stretch_b256_256", 256, 256, 256, "stretch", 100)
    if want("streams"):
        for ns in (1, 2, 3):
            two_stream_case(f"c2_abi_{ns}stream", 512, 400, ns)
        two_stream_case("c2_abi_1stream_B128", 128, 400, 1)
        two_stream_case("c2_abi_1stream_B256", 256, 400, 1)
        two_stream_case("c2_abi_1stream_B1024", 1024, 200, 1)
    if want("calib"):
        calibrate()
    if want("c5"):
        metrics_case("c5_fp32_125k", 125_000, 1000, torch.float32, 50)
        metrics_case("c5_bf16_125k", 125_000, 1000, torch.bfloat16, 50)
        metrics_case("c5_fp32_1000", 1000, 1000, torch.float32, 300)
        metrics_case("c5_fp32_512", 512, 1000, torch.float32, 300)


if __name__ == "__main__":
    main()
