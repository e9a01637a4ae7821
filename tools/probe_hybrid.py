#!/usr/bin/env python
"""Development probe: e2e C2 steps from pinned host memory when a fraction of every batch goes through the copy engine
(whole images, DMA) and the rest is read in place by the kernel (crop rows only).  Two streams, as in bench.py."""
import sys, time
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from torchbooster_b200.transforms import BatchTransform, CropFlipParams, TransformSpec

dev = torch.device("cuda", 0)
B, H, W, crop = 512, 256, 256, 224
spec = TransformSpec(crop=crop, flip_p=.5, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225), dtype=torch.bfloat16)
bt = BatchTransform(spec, device=dev)
g = torch.Generator().manual_seed(1)
host = [torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g).pin_memory() for _ in range(3)]
rng = CropFlipParams(42)
streams = [torch.cuda.Stream(dev), torch.cuda.Stream(dev)]
outs = [torch.empty((B, 3, crop, crop), dtype=torch.bfloat16, device=dev) for _ in range(2)]
done = [torch.cuda.Event(), torch.cuda.Event()]


def run(frac, K=40):
    nd = int(B * frac) // 8 * 8
    stage = [torch.empty((max(nd, 1), H, W, 3), dtype=torch.uint8, device=dev) for _ in range(2)]

    def loop(n):
        for i in range(n):
            s, j = i & 1, i % 3
            t, l, f = rng.draw(B, H, W, (crop, crop), (0, 0), 0.5)
            with torch.cuda.stream(streams[s]):
                if nd:
                    stage[s][:nd].copy_(host[j][:nd], non_blocking=True)
                    bt(stage[s][:nd], params=(t[:nd], l[:nd], f[:nd]), out=outs[s][:nd])
                if nd < B:
                    bt(host[j][nd:], params=(t[nd:], l[nd:], f[nd:]), out=outs[s][nd:])
                done[s].record()
            if i:
                done[1 - s].synchronize()
        done[(n - 1) & 1].synchronize()
    loop(6)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    loop(K)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"dma fraction {frac:.2f} ({nd} images): {dt / K * 1e3:.3f} ms per step, {B * K / dt / 1e3:.1f} k img/s", flush=True)


for frac in (0.0, 0.1, 0.15, 0.2, 0.3, 0.5, 1.0, 0.0, 0.15):
    run(frac)
