#!/usr/bin/env python
"""One small invocation of every kernel (both transform paths, both resize paths, metrics for the
three dtypes, running sum) for compute-sanitizer: `compute-sanitizer --tool memcheck python tools/sanitize_cases.py`."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from torchbooster_b200 import _native as N                                   # noqa: E402
from torchbooster_b200.metrics import DeviceRunningAverage, MetricState     # noqa: E402
from torchbooster_b200.transforms import BatchTransform, TransformSpec      # noqa: E402

IM = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
rng = np.random.default_rng(0)
dev = "cuda"


def xf(B, H, W, C, dtype, flags=0, pinned=False, **kw):
    x = torch.from_numpy(rng.integers(0, 256, (B, H, W, C), dtype=np.uint8))
    x = x.pin_memory() if pinned else x.to(dev)
    spec = TransformSpec(mean=IM[0][:C], std=IM[1][:C], dtype=dtype, **kw)
    bt = BatchTransform(spec, device=dev, seed=1)
    bt.flags = flags
    return bt(x)


xf(7, 64, 80, 3, torch.bfloat16, crop=(48, 64), flip_p=.5)                      # stream kernel, row windows
xf(7, 64, 80, 3, torch.float32, crop=(48, 64), flip_p=.5)
xf(5, 32, 32, 3, torch.float32, flip_p=.5)                                      # stream kernel, contiguous tiles
xf(5, 32, 32, 3, torch.float32, crop=32, padding=4, padding_mode="reflect", flip_p=.5, flags=N.FLAG_FORCE_STREAM)
xf(5, 32, 32, 3, torch.float16, crop=32, padding=4, padding_mode="reflect", flip_p=.5)   # gather kernel
xf(3, 17, 19, 1, torch.float32, crop=(9, 11), padding=2, padding_mode="symmetric", flip_p=.5)
xf(3, 64, 80, 3, torch.bfloat16, crop=(48, 64), flip_p=.5, pinned=True)         # zero-copy
xf(2, 96, 128, 3, torch.float32, resize=64, center_crop=64)                     # DP4A resize
xf(2, 96, 128, 1, torch.bfloat16, resize=64)
xf(1, 300, 260, 3, torch.float32, resize=64, center_crop=72)                    # generic resize (scale > 2), zero fill
xf(1, 40, 56, 3, torch.float32, resize=(40, 30))                                # one pass only
for dt in (torch.float32, torch.bfloat16, torch.float16):
    for n, c in ((33, 1000), (17, 1003), (9, 5), (5, 4099)):
        lg = torch.randn(n, c).to(dt).to(dev)
        y = torch.randint(-1, c + 1, (n,)).to(dev)
        st = MetricState(c, k=5, device=dev)
        st.update(lg, y)
        st.update_loss(lg[:, 0])
        st.compute()
d = DeviceRunningAverage(dev)
d.update(torch.rand(1000, device=dev))
assert abs(d.value - 0.5) < 0.1
torch.cuda.synchronize()
print("sanitize cases done, launches:", N.lib().tb_launch_count())
