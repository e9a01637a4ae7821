#!/usr/bin/env python
"""Development probe: does a small kernel on a second stream run WHILE the persistent transform kernel runs, or only
at its boundaries?  Times 20 small launches on a side stream against a train of C2 transform launches."""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from torchbooster_b200 import _native as N
from torchbooster_b200.metrics import MetricState
from torchbooster_b200.transforms import BatchTransform, CropFlipParams, TransformSpec

dev = torch.device("cuda", 0)
B = 512
spec = TransformSpec(crop=224, flip_p=.5, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225), dtype=torch.bfloat16)
bt = BatchTransform(spec, device=dev)
srcs = [torch.randint(0, 256, (B, 256, 256, 3), dtype=torch.uint8, device=dev) for _ in range(6)]
t, l, f = CropFlipParams(42).draw(B, 256, 256, (224, 224), (0, 0), .5)
prm = tuple(torch.from_numpy(a).to(dev) for a in (t, l, f))
outs = [torch.empty((B, 3, 224, 224), dtype=torch.bfloat16, device=dev) for _ in range(2)]
logits = torch.randn(512, 1000, device=dev)
labels = torch.randint(0, 1000, (512,), device=dev)
st = MetricState(1000, k=5, device=dev)
small = torch.zeros(1 << 16, device=dev)
main, side = torch.cuda.current_stream(dev), torch.cuda.Stream(dev)


def train(n, flags=0):
    bt.flags = flags
    for i in range(n):
        bt(srcs[i % 6], params=prm, out=outs[i & 1])


def timed(fn_side, with_main, flags=0, n_side=20):
    torch.cuda.synchronize()
    e0, e1, m1 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record(main)
    side.wait_stream(main)
    if with_main:
        train(20, flags)
    m1.record(main)
    with torch.cuda.stream(side):
        for _ in range(n_side):
            fn_side()
        e1.record(side)
    torch.cuda.synchronize()
    return round(e0.elapsed_time(e1), 3), round(e0.elapsed_time(m1), 3)


train(3); st.update(logits, labels); small.add_(1.0)
for name, fn in (("metrics kernel", lambda: st.update(logits, labels)), ("aten add_ (64K floats)", lambda: small.add_(1.0))):
    print(name, "alone (side done, main done) ms:", timed(fn, False))
    print(name, "with 20 transform launches:", timed(fn, True))
    print(name, "with 20 transform launches, no PDL:", timed(fn, True, N.FLAG_NO_PDL))
    print(name, "with 20 generic-kernel transform launches:", timed(fn, True, N.FLAG_FORCE_GENERIC))
