"""Development: per-CTA %globaltimer trace of the resize ring kernel (needs a library built with the trace patch:
build_ab/libT.so, see ROUND_NOTES.md 'Pitfalls').  Prints, per trace point, min / median / max over CTAs in microseconds
after the earliest CTA entry of the launch."""
import ctypes, sys
import numpy as np
import torch
from torchbooster_b200 import _native as N
from torchbooster_b200.transforms import BatchTransform, TransformSpec

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
lib = N.lib()
lib.tb_trace_read.argtypes = [ctypes.c_void_p, ctypes.c_int]
dev = torch.device("cuda", 0)
spec = TransformSpec(resize=256, center_crop=256, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225))
bt = BatchTransform(spec, device=dev)
x = [torch.randint(0, 256, (B, 512, 512, 3), dtype=torch.uint8, device=dev) for _ in range(8)]
outs = [torch.empty((B, 3, 256, 256), device=dev) for _ in range(8)]
for i in range(20):
    bt(x[i % 8], out=outs[i % 8])
torch.cuda.synchronize()
names = ["entry", "dep wait done", "table filled", "first copy issued", "first slot full", "first store", "grp0 done", "grpN done", "producer0 done", "chunk set up", "waiting for slot", "chunk bounds read", "group barrier passed", "row taps stored", "column taps read"]
for mode in ("alone", "back-to-back"):
    lib.tb_trace_read(None, 1)
    torch.cuda.synchronize()
    if mode == "alone":
        bt(x[0], out=outs[0])
    else:
        for i in range(6):
            bt(x[i % 8], out=outs[i % 8])       # the trace holds the last launch
    torch.cuda.synchronize()
    buf = np.zeros(256 * 16, dtype=np.uint64)
    lib.tb_trace_read(buf.ctypes.data, 0)
    t = buf.reshape(256, 16).astype(np.int64)
    t = t[t[:, 0] > 0]
    t0 = t[:, 0].min()
    print(f"== {mode}: {len(t)} CTAs, batch {B}")
    for k, nm in enumerate(names):
        v = t[:, k]
        v = v[v > 0]
        if len(v) == 0:
            continue
        r = (v - t0) / 1e3
        print(f"{nm:18s} min {r.min():7.2f}  med {np.median(r):7.2f}  max {r.max():7.2f}")
    if mode == "alone":
        per = (np.maximum(t[:, 6], t[:, 7]) - t[:, 0]) / 1e3
        print(f"CTA lifetime       min {per.min():7.2f}  med {np.median(per):7.2f}  max {per.max():7.2f}")
