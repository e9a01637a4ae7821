#!/usr/bin/env python
"""The data + metric side of the reference's CIFAR example (examples/img_cls/resnet/resnet.py:52-122)
on torchbooster_b200, end to end, with a stand-in model:

    LMDB records -> BatchLoader (background gather into pinned slots)
                 -> ONE launch per batch: RandomCrop(32, 4, reflect) + flip + RandomRotation(15) + RandAugment()
                    + ToTensor + Normalize                                  (resnet.py:96-103)
                 -> model -> MetricState.update (top-1 / top-5 / confusion, no host sync)
                 -> DeviceRunningAverage(loss)                               (resnet.py:55,66)
                 -> all_reduce once per epoch                                (one NCCL call)

    python examples/cifar_pipeline.py                      # one GPU
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 examples/cifar_pipeline.py
"""
from __future__ import annotations

import argparse
import os
import sys
import tempfile
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from torchbooster_b200 import distributed as dist                      # noqa: E402  drop-in for torchbooster.distributed
from torchbooster_b200.lmdb import LMDBReader, write_image_lmdb        # noqa: E402
from torchbooster_b200.loader import BatchLoader                       # noqa: E402
from torchbooster_b200.metrics import DeviceRunningAverage, MetricState  # noqa: E402
from torchbooster_b200.transforms import BatchTransform, TransformSpec  # noqa: E402

CIFAR = ((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))           # resnet.py:94


def main(argv=None) -> dict:
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=10_000)
    ap.add_argument("--batch", type=int, default=256)                  # resnet.yml
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--db", default=None, help="an existing LMDB of raw 32x32x3 records (default: a synthetic one)")
    a = ap.parse_args(argv)

    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=dev)

    tmp = None
    path = a.db
    if path is None:
        tmp = tempfile.TemporaryDirectory()
        path = os.path.join(tmp.name, "db")
        if dist.is_primary():
            rng = np.random.default_rng(0)
            write_image_lmdb(path, rng.integers(0, 256, (a.images, 32, 32, 3), dtype=np.uint8), labels=rng.integers(0, 10, a.images))
        if world > 1:                                                  # every rank reads rank 0's file (single node)
            box = [path]
            torch.distributed.broadcast_object_list(box, src=0)
            path = box[0]
        dist.synchronize()
    reader = LMDBReader(path)
    labels = reader.read_labels()

    spec = TransformSpec(crop=32, padding=4, padding_mode="reflect", flip_p=0.5, rotation=15, rand_augment=True,
                         mean=CIFAR[0], std=CIFAR[1], dtype=torch.bfloat16)
    transform = BatchTransform(spec, device=dev, seed=42 + rank)       # torchbooster.utils.seed(42) per rank
    gen = torch.Generator(device=dev).manual_seed(0)
    head = torch.randn(3 * 32 * 32, 10, device=dev, generator=gen).to(torch.bfloat16) * 0.02     # the stand-in "model"
    state, loss_avg = MetricState(10, k=5, device=dev), DeviceRunningAverage(device=dev)

    report = {}
    for epoch in range(a.epochs):
        sampler = dist.data_sampler(range(len(reader)), shuffle=True, distributed=world > 1)
        if world > 1:
            sampler.set_epoch(epoch)
        loader = BatchLoader(reader, (32, 32, 3), a.batch, sampler=sampler, transform=transform, labels=labels, drop_last=True)
        state.reset()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        seen = 0
        for x, y in loader:
            logits = x.flatten(1) @ head
            loss = torch.nn.functional.cross_entropy(logits.float(), y)
            state.update(logits, y)                                    # no host sync
            loss_avg.update(loss)                                      # no .item()
            seen += x.shape[0]
        state.all_reduce()                                             # the one exchange
        res = state.compute()                                          # the one host sync
        dt = time.perf_counter() - t0
        report = {"epoch": epoch, "images": int(res["n"]), "images_per_s": world * seen / dt, "top1": res["correct_1"] / max(res["n"], 1),
                  "top5": res["correct_5"] / max(res["n"], 1), "loss": float(loss_avg.value)}
        if dist.is_primary():
            print(report, flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()
    if tmp is not None and dist.is_primary():
        tmp.cleanup()
    return report


if __name__ == "__main__":
    main()
