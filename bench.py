#!/usr/bin/env python
"""Headline benchmark: images/sec transformed + scored (BASELINE.json), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (N=1 and per rank for N>1): BASELINE.json configs[1] -- an ImageNet-shaped synthetic
batch, 512 x 256x256x3 uint8 -> random crop 224, horizontal flip, ImageNet mean/std, bf16 NCHW --
plus the scoring of that batch: 512 x 1000 fp32 logits -> top-1 / top-5 counts and the 1000x1000
confusion matrix.  A step = one batch through both kernels.  `value` has the inputs resident in
HBM; `e2e` starts from pinned HOST buffers and goes through the package's public API
(BatchTransform / MetricState), host->device copies and a device->host read of the counters
inside the timed region.  With N>1 the batches are sharded per rank (weak scaling) and the only
cross-GPU traffic is one all_reduce of the packed metric buffer per evaluation interval (= the K
timed steps), inside the timed region.

`--impl reference` times the reference's own CPU path (oracle/: torchvision PIL transforms in
DataLoader workers + torchbooster.metrics.accuracy / topk / bincount) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

IMAGENET_MEAN, IMAGENET_STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
L2_BYTES = 126 * 1024 * 1024

CIFAR_MEAN, CIFAR_STD = (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)

WORKLOADS = {
    # the headline workload (BASELINE.json configs[1])
    "c2": dict(B=512, H=256, W=256, crop=224, dtype=torch.bfloat16, classes=1000, k=5,
               alg_bytes=224 * 224 * 3 + 224 * 224 * 3 * 2,
               desc="C2 ImageNet-shaped synthetic batch: 512x256x256x3 u8 -> random crop 224, flip, "
                    "normalize, bf16 NCHW; scored: 512x1000 fp32 logits -> top-1/top-5/confusion"),
}

# Every BASELINE.json configuration, timed per kernel launch inside the same run (the `configs` block of the line).
# alg_bytes = SURVEY.md 8(d)'s algorithmic bytes per image / per sample.
CONFIGS = {
    "c1": dict(kind="transform", B=256, H=32, W=32, alg_bytes=3072 + 12288, iters=200, kernel="tb_xf_generic_kernel",
               spec=dict(crop=32, padding=4, padding_mode="reflect", flip_p=0.5, mean=CIFAR_MEAN, std=CIFAR_STD),
               desc="C1 CIFAR-shaped: 256x32x32x3 u8, reflect-pad 4, random crop 32, flip, CIFAR normalize -> fp32 "
                    "(3.9 MB batch: launch-latency bound, reported as latency)"),
    "c2": dict(kind="transform", B=512, H=256, W=256, alg_bytes=224 * 224 * 9, iters=100, kernel="tb_xf_stream_kernel",
               spec=dict(crop=224, flip_p=0.5, mean=IMAGENET_MEAN, std=IMAGENET_STD, dtype=torch.bfloat16),
               desc="C2 ImageNet-shaped: 512x256x256x3 u8 -> random crop 224, flip, normalize -> bf16 (one stream, back to back)"),
    "c3": dict(kind="transform", B=1024, H=64, W=64, alg_bytes=12288 + 49152, iters=200, kernel="tb_xf_stream_kernel",
               spec=dict(flip_p=0.5, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5)),
               desc="C3 img_gen: 1024x64x64x3 u8, flip, scale to [-1,1] -> fp32"),
    "c4": dict(kind="transform", B=32, H=512, W=512, alg_bytes=786432 + 786432, iters=200, kernel="tb_resize_ring_kernel",
               spec=dict(resize=256, center_crop=256, mean=IMAGENET_MEAN, std=IMAGENET_STD), params=False,
               desc="C4 img_stt: 32x512x512x3 u8, Pillow-exact bilinear resize to 256, centre crop, normalize -> fp32"),
    "c4_b256": dict(kind="transform", B=256, H=512, W=512, alg_bytes=786432 + 786432, iters=50, kernel="tb_resize_ring_kernel",
                    spec=dict(resize=256, center_crop=256, mean=IMAGENET_MEAN, std=IMAGENET_STD), params=False,
                    desc="C4 at batch 256 (content + style batches of eight steps in one launch)"),
    "c4_b1": dict(kind="transform", B=1, H=512, W=512, alg_bytes=786432 + 786432, iters=200, kernel="tb_resize_ring_kernel",
                  spec=dict(resize=256, center_crop=256, mean=IMAGENET_MEAN, std=IMAGENET_STD), params=False, dev_only=True,
                  desc="C4 at batch 1: the fixed cost of one launch (development)"),
    "c5": dict(kind="metrics", N=125_000, C=1000, dtype=torch.float32, alg_bytes=4008, iters=50, kernel="tb_metrics_kernel",
               desc="C5: one rank's share (125 000 x 1000 fp32 logits) of the 1 M rows: top-1 + top-5 + confusion in one pass"),
    "c5_bf16": dict(kind="metrics", N=125_000, C=1000, dtype=torch.bfloat16, alg_bytes=2008, iters=50, kernel="tb_metrics_kernel",
                    desc="C5 with bf16 logits (exact ties are common)"),
}


# -------------------------------------------------------------------------------------------------
# clocks
# -------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock and clock-event reasons of one GPU while the timed region runs (NVML,
    every few ms; nvidia-smi's own sampling period is too coarse for a sub-second region)."""

    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20}
    NOTED = {"sw_power_cap": 0x4, "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10,
             "applications_clocks_setting": 0x2, "display_clock_setting": 0x100}

    def __init__(self, index: int, period_s: float = 0.001):
        self.index, self.period = index, period_s
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self.active = False          # only samples taken while this is set are kept
        self.in_region = 0
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None

    def __enter__(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            uuid = torch.cuda.get_device_properties(self.index).uuid
            try:
                self._h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
            except Exception:
                self._h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        except Exception as e:  # pragma: no cover - NVML missing
            self.error = repr(e)
        return self

    def _run(self):
        nv = self._nvml
        while not self._stop.is_set():
            if self.active:
                try:
                    mhz = nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
                    if self.active:
                        self.samples.append(mhz)
                        for name, bit in {**self.BAD, **self.NOTED}.items():
                            if mask & bit:
                                self.reasons.add(name)
                except Exception:
                    pass
            self._stop.wait(self.period)

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=1.0)
        return False

    def summary(self) -> dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(statistics.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples), "samples_in_timed_region": self.in_region}


# -------------------------------------------------------------------------------------------------
# CPU leg (the oracle, timed): cpu_baseline of our arm and the whole `--impl reference` arm
# -------------------------------------------------------------------------------------------------
class _SyntheticImages(torch.utils.data.Dataset):
    """BaseDataset-shaped dataset over in-memory synthetic bytes (no LMDB B-tree lookup: this
    favours the CPU baseline) feeding the reference-style torchvision pipeline."""

    def __init__(self, images: np.ndarray, transform):
        self.images, self.transform = images, transform

    def __len__(self):
        return self.images.shape[0]

    def __getitem__(self, idx):
        from PIL import Image
        return self.transform(Image.fromarray(self.images[idx])), idx % 1000


def _cpu_compose(name: str):
    """The reference-style torchvision pipeline of a configuration (per-sample, PIL backend)."""
    import torchvision.transforms as T
    if name == "c1":      # examples/img_cls/resnet/resnet.py:94-98
        return T.Compose([T.RandomCrop(32, padding=4, padding_mode="reflect"), T.RandomHorizontalFlip(), T.ToTensor(),
                          T.Normalize(CIFAR_MEAN, CIFAR_STD)])
    if name == "c2":
        return T.Compose([T.RandomCrop(224), T.RandomHorizontalFlip(), T.ToTensor(), T.Normalize(IMAGENET_MEAN, IMAGENET_STD)])
    if name == "c3":      # examples/img_gen/gan/gan.py:99,135 style
        return T.Compose([T.RandomHorizontalFlip(), T.ToTensor(), T.Normalize((0.5,) * 3, (0.5,) * 3)])
    if name in ("c4", "c4_b256"):   # examples/img_stt/adain/adain.py:167-171
        return T.Compose([T.Resize(256), T.CenterCrop(256), T.ToTensor(), T.Normalize(IMAGENET_MEAN, IMAGENET_STD)])
    raise KeyError(name)


class _StaggeredBatches(torch.utils.data.Sampler):
    """Batches of B indices (with replacement from the pool); the FIRST batch of every worker is shorter -- worker w
    gets (w + 1) / workers of a batch -- so that the workers, which all start at the same moment and take the same time
    per batch, end up evenly out of phase: afterwards whole batches arrive one every period / workers instead of in
    bursts of `workers`, and a window of any K batches measures the steady-state rate."""

    def __init__(self, pool: int, B: int, workers: int, batches: int, seed: int = 1234):
        self.pool, self.B, self.workers, self.batches, self.seed = pool, B, max(workers, 1), batches, seed

    def __len__(self):
        return self.batches

    def __iter__(self):
        rng = np.random.default_rng(self.seed)
        for i in range(self.batches):
            n = self.B if i >= self.workers else max(1, (i + 1) * self.B // self.workers)
            yield rng.integers(0, self.pool, n).tolist()


def cpu_path(name: str, B: int, H: int, W: int, steps: int, workers: int, max_seconds: float = 1e9, score=None,
             out_dtype=torch.float32, prefetch: int = 2, warm_seconds: float = 8.0):
    """Reference-style CPU path: per-sample torchvision PIL transforms in DataLoader workers, B images per worker batch
    (resnet.py:94-107 pattern), default collate to fp32 NCHW, then (score = (classes, k)) accuracy (the oracle's
    restatement of torchbooster/metrics.py:27) + torch.topk + torch.bincount on all host threads.
    Steady state: (1) DataLoader workers run in lockstep, so batches arrive in bursts of `workers`;
    `_StaggeredBatches` puts the workers evenly out of phase; (2) the first seconds of such a run are 2-3x faster than
    what the host sustains (measured on the B200 box: 14 k img/s over the first 20 batches of C2, 8 k over 60, 5.8 k
    from then on), so the clock starts only after `warm_seconds` AND workers * (prefetch + 1) batches.
    Returns (images_per_second, steps_done, seconds)."""
    from oracle import metrics_oracle as MO
    rng = np.random.default_rng(1234)
    pool = rng.integers(0, 256, (min(2048, max(B * 2, 256)), H, W, 3), dtype=np.uint8)
    warm_batches = workers * (prefetch + 1) if workers else 2
    ds = _SyntheticImages(pool, _cpu_compose(name))
    loader = torch.utils.data.DataLoader(ds, batch_sampler=_StaggeredBatches(pool.shape[0], B, workers, 10 ** 7),
                                         num_workers=workers, pin_memory=False, persistent_workers=False,
                                         prefetch_factor=prefetch if workers else None)
    if score:
        classes, k = score
        gen = torch.Generator().manual_seed(1234)
        logits = torch.randn(B, classes, generator=gen)
        cm = torch.zeros(classes ** 2, dtype=torch.int64)
    done, t0, n1, nk = 0, None, 0, 0
    start = time.perf_counter()
    it = iter(loader)
    i = 0
    while True:
        x, y = next(it)
        if t0 is not None:
            pass
        x = x.to(out_dtype) if out_dtype != torch.float32 else x
        if score and x.shape[0] == B:
            n1 += int(MO.accuracy(logits, y) * B)
            nk += int((logits.topk(k, -1).indices == y[:, None]).any(-1).sum())
            cm += torch.bincount(y * classes + logits.argmax(-1), minlength=classes ** 2)
        i += 1
        if t0 is None:
            if i >= warm_batches and time.perf_counter() - start >= warm_seconds:
                t0 = time.perf_counter()          # the clock starts after this (untimed) batch
            continue
        done += 1
        if done >= steps or time.perf_counter() - t0 > max_seconds:
            break
    dt = time.perf_counter() - t0
    del it, loader
    return done * B / dt, done, dt


def cpu_metrics_path(n: int, classes: int, dtype, k: int = 5, max_seconds: float = 4.0):
    """accuracy (torchbooster/metrics.py:27 restated) + torch.topk + torch.bincount over [n, classes] logits on every
    host thread.  Returns (samples_per_second, passes, seconds)."""
    from oracle import metrics_oracle as MO
    gen = torch.Generator().manual_seed(1234)
    rows = min(n, 25_000)                       # a bounded sample of the rows; the rate does not depend on n
    logits = torch.randn(rows, classes, generator=gen).to(dtype)
    labels = torch.randint(0, classes, (rows,), generator=gen)
    cm = torch.zeros(classes ** 2, dtype=torch.int64)

    def once():
        float(MO.accuracy(logits, labels))
        int((logits.topk(k, -1).indices == labels[:, None]).any(-1).sum())
        cm.add_(torch.bincount(labels * classes + logits.argmax(-1), minlength=classes ** 2))
    once()
    t0, done = time.perf_counter(), 0
    while True:
        once()
        done += 1
        if time.perf_counter() - t0 > max_seconds or done >= 20:
            break
    dt = time.perf_counter() - t0
    return done * rows / dt, done, dt


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_model() -> str:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    workers = min(host_threads(), 64)
    torch.set_num_threads(max(1, host_threads()))
    # A step is one batch of the workload; for very long runs the per-step sample shrinks so that
    # the whole run stays within a few minutes (about one million images in total).
    full_b = B = wl["B"]
    if args.steps * full_b > 1_000_000:
        B = max(16, (1_000_000 // args.steps) // 16 * 16)
    ips, done, dt = cpu_path("c2", B, wl["H"], wl["W"], args.steps, workers, score=(wl["classes"], wl["k"]),
                             out_dtype=wl["dtype"])
    line = {
        "impl": "reference", "metric": "images/sec transformed+scored", "value": ips, "unit": "images/s",
        "n_gpus": args.gpus, "steps": done, "warmup": args.warmup, "ms_per_step": dt / max(done, 1) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": wl["desc"],
                   "per_step": f"one batch of {B} images on the host CPU" +
                               ("" if B == full_b else f" (bounded sample of the {full_b}-image batch)"),
                   "timing": "the DataLoader workers are put evenly out of phase (a shorter first batch each) and the clock starts "
                             "after 8 s of untimed batches: the first seconds of such a run are 2-3x faster than the host sustains"},
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": workers, "kind": "port",
                         "sample": f"{done} batches of {B} images after 8 s of untimed ones (workers out of phase, sustained rate); torchvision "
                                   f"{__import__('torchvision').__version__} PIL transforms in {workers} DataLoader workers + "
                                   f"accuracy/topk/bincount; {cpu_model()}"},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# -------------------------------------------------------------------------------------------------
# our arm
# -------------------------------------------------------------------------------------------------
def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(workload: str):
    """Per-launch DRAM bytes of the dominant kernel from the committed ncu --set full capture."""
    p = ROOT / "profiles" / "traffic.json"
    if p.exists():
        try:
            return json.loads(p.read_text()).get(workload)
        except Exception:
            pass
    return None


def bind_to_gpu_node(index: int):
    """Pins this process to the CPUs NVML lists as local to its GPU, before any pinned host buffer is
    allocated: the buffers the kernel reads over PCIe then lie on the GPU's own NUMA node instead of
    wherever the launcher happened to start the rank.  Returns the number of CPUs in the mask."""
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = torch.cuda.get_device_properties(index).uuid
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(index)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return len(os.sched_getaffinity(0))
    except Exception:
        return None


def _graph_or_eager(fn, iters, dev):
    """`iters` calls of fn(i) captured into one CUDA graph (the host then issues ONE launch, so small kernels are timed
    on the device, not on the Python call rate); returns a callable that runs them all.  Eager loop if capture fails."""
    try:
        cap = torch.cuda.Stream(dev)
        cap.wait_stream(torch.cuda.current_stream(dev))
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=cap):
            for i in range(iters):
                fn(i)
        torch.cuda.current_stream(dev).wait_stream(cap)
        torch.cuda.synchronize(dev)
        return graph.replay, True
    except Exception as e:      # pragma: no cover
        print(f"bench.py: CUDA graph capture failed ({e!r}); launching eagerly", file=sys.stderr)
        torch.cuda.synchronize(dev)

        def run():
            for i in range(iters):
                fn(i)
        return run, False


def _timed(run, dev):
    run()                       # graph upload / warm-up
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run()
    e1.record()
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1)


def run_configs(dev, rank, world, with_cpu, transport, only=None):
    """The `configs` block: every BASELINE.json configuration through the public API, `iters` launches back to back on
    one stream (captured into a CUDA graph), CUDA events, inputs rotated over more than L2 of distinct buffers.  With
    several ranks every rank runs its own copy at the same time and the slowest rank's time is reported."""
    import torch.distributed as dist
    from torchbooster_b200.metrics import MetricState
    from torchbooster_b200.transforms import BatchTransform, CropFlipParams, TransformSpec
    peak, peak_src = measured_peak()
    out, ms_list = {}, []
    names = [n for n in CONFIGS if (only is None and not CONFIGS[n].get("dev_only")) or (only is not None and n in only)]
    for name in names:
        cfg = CONFIGS[name]
        iters = cfg["iters"]
        if cfg["kind"] == "transform":
            B, H, W = cfg["B"], cfg["H"], cfg["W"]
            spec = TransformSpec(**cfg["spec"])
            bt = BatchTransform(spec, device=dev)
            nbuf = min(24, max(3, -(-4 * L2_BYTES // (B * H * W * 3))))
            gen = torch.Generator(device=dev).manual_seed(1234 + rank)
            srcs = [torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, device=dev, generator=gen) for _ in range(nbuf)]
            g = bt.geometry(H, W)
            prm = None
            if cfg.get("params", True):
                t, l, f = CropFlipParams(42 + rank).draw(B, g["cur"][0], g["cur"][1], g["crop"], g["pad"], spec.flip_p)
                prm = tuple(torch.from_numpy(a).to(dev) for a in (t, l, f))
            outs = [torch.empty((B, 3) + tuple(g["out"]), dtype=spec.dtype, device=dev) for _ in range(2)]
            fn = lambda i: bt(srcs[i % nbuf], params=prm, out=outs[i & 1])
            items, unit = B, "images/s"
            l2 = f"inputs rotate over {nbuf} distinct batches ({nbuf * B * H * W * 3 >> 20} MiB)"
        else:
            n, C, dt = cfg["N"], cfg["C"], cfg["dtype"]
            es = 4 if dt == torch.float32 else 2
            nbuf = min(8, max(2, -(-4 * L2_BYTES // (n * C * es))))
            gen = torch.Generator(device=dev).manual_seed(1234 + rank)
            logits = [torch.randn(n, C, device=dev, generator=gen).to(dt) for _ in range(nbuf)]
            labels = [torch.randint(0, C, (n,), device=dev, generator=gen) for _ in range(nbuf)]
            st = MetricState(C, k=5, device=dev)
            fn = lambda i: st.update(logits[i % nbuf], labels[i % nbuf])
            items, unit = n, "samples/s"
            l2 = f"inputs rotate over {nbuf} distinct matrices ({nbuf * n * C * es >> 20} MiB)"
        for i in range(3):
            fn(i)
        torch.cuda.synchronize(dev)
        run, graphed = _graph_or_eager(fn, iters, dev)
        if world > 1:
            dist.barrier()
        ms = _timed(run, dev) / iters
        ms_list.append(ms)
        out[name] = {"workload": cfg["desc"], "ms_per_launch": ms, "launches_timed": iters, "items_per_launch": items, "unit": unit,
                     "launch": "CUDA graph" if graphed else "eager", "l2": l2,
                     "roofline": {"bound": "hbm", "kernel": cfg["kernel"], "peak": peak, "unit": "GB/s", "peak_source": peak_src,
                                  "algorithmic_bytes_per_launch": cfg["alg_bytes"] * items, "traffic": ncu_traffic(name)}}
        del fn, run
        if cfg["kind"] == "transform":
            del srcs, outs, bt
        else:
            del logits, labels, st
        torch.cuda.empty_cache()
    if world > 1:       # slowest rank per configuration
        t = torch.tensor(ms_list, dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_list = [float(v) for v in t.tolist()]
    for name, ms in zip(names, ms_list):
        r = out[name]
        r["ms_per_launch"] = ms
        r["items_per_s"] = world * r["items_per_launch"] / (ms * 1e-3)
        rf = r["roofline"]
        rf["achieved"] = rf["algorithmic_bytes_per_launch"] / (ms * 1e-3) / 1e9
        rf["frac"] = rf["achieved"] / peak
        rf["frac_of_nominal_8000"] = rf["achieved"] / 8000.0
    if world > 1 and "c5" in out:
        # the exchange of the C5 state: the packed int64 buffer (8 MB at C = 1000) and its int32 transport (4 MB)
        st = MetricState(1000, k=5, device=dev)
        res = {}
        for tr in ("int64", "int32"):
            for _ in range(3):
                st.all_reduce(transport=tr)
            torch.cuda.synchronize(dev)
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                st.all_reduce(transport=tr)
            e1.record()
            torch.cuda.synchronize(dev)
            t = torch.tensor([e0.elapsed_time(e1) / 10], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res[tr] = float(t.item())
        out["c5"]["all_reduce"] = {"ranks": world, "state": "int64 [4 + 1000*1000] (8 MB)", "ms_int64_transport": res["int64"],
                                   "ms_int32_transport": res["int32"], "backend": "nccl",
                                   "timing": "10 back-to-back exchanges, CUDA events, max over ranks"}
    if with_cpu and world == 1 and rank == 0:
        workers = min(host_threads(), 64)
        torch.set_num_threads(max(1, host_threads()))
        import torchvision
        for name in names:
            cfg = CONFIGS[name]
            if name == "c2":
                continue                    # the line's own cpu_baseline
            if name == "c4_b256":
                out[name]["cpu_baseline"] = out.get("c4", {}).get("cpu_baseline")
                continue
            if cfg["kind"] == "transform":
                ips, done, dt = cpu_path(name, cfg["B"], cfg["H"], cfg["W"], steps=100000, workers=workers, max_seconds=3.0,
                                         out_dtype=cfg["spec"].get("dtype", torch.float32), warm_seconds=3.0)
                out[name]["cpu_baseline"] = {"value": ips, "unit": "images/s", "cores": workers, "kind": "port",
                                             "sample": f"{done} batches of {cfg['B']} images in {dt:.1f} s after 3 s of untimed ones; torchvision "
                                                       f"{torchvision.__version__} PIL transforms in {workers} DataLoader workers"}
            else:
                sps, done, dt = cpu_metrics_path(cfg["N"], cfg["C"], cfg["dtype"])
                out[name]["cpu_baseline"] = {"value": sps, "unit": "samples/s", "cores": host_threads(), "kind": "port",
                                             "sample": f"{done} passes over 25 000 x {cfg['C']} logits in {dt:.1f} s; accuracy + torch.topk + "
                                                       f"torch.bincount on {host_threads()} host threads"}
    return out


def run_ours(args, wl):
    from torchbooster_b200 import _native as N
    from torchbooster_b200.metrics import MetricState
    from torchbooster_b200.transforms import BatchTransform, CropFlipParams, TransformSpec
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; torchbooster_b200 has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    affinity = None if args.no_affinity else bind_to_gpu_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = N.lib()
    B, H, W, crop, classes, k = wl["B"], wl["H"], wl["W"], wl["crop"], wl["classes"], wl["k"]
    K, Wm = args.steps, max(args.warmup, 3)

    spec = TransformSpec(crop=crop, flip_p=0.5, mean=IMAGENET_MEAN, std=IMAGENET_STD, dtype=wl["dtype"])
    bt = BatchTransform(spec, device=dev)
    if args.variant == "generic":
        bt.flags = N.FLAG_FORCE_GENERIC
    lut = bt.lut(3)
    state = MetricState(classes, k=k, device=dev)

    # ---- synthetic inputs, resident in HBM; rotated so that consecutive steps never hit L2 -------
    in_bytes = B * H * W * 3
    nbuf = min(16, max(4, -(-4 * L2_BYTES // in_bytes)))
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    srcs = [torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, device=dev, generator=gen) for _ in range(nbuf)]
    outs = [torch.empty((B, 3, crop, crop), dtype=wl["dtype"], device=dev) for _ in range(2)]
    logits = [torch.randn(B, classes, device=dev, generator=gen) for _ in range(nbuf)]
    labels = [torch.randint(0, classes, (B,), device=dev, generator=gen) for _ in range(nbuf)]
    rng = CropFlipParams(42 + rank)
    params = []
    for _ in range(nbuf):
        t, l, f = rng.draw(B, H, W, (crop, crop), (0, 0), 0.5)
        params.append((torch.from_numpy(t).to(dev), torch.from_numpy(l).to(dev), torch.from_numpy(f).to(dev)))
    stream = torch.cuda.current_stream(dev)
    # Scoring does not depend on the transform of the same step (in training the logits come from
    # the model), so it runs on a second stream and overlaps the transform kernel.
    side = torch.cuda.Stream(dev, priority=-1 if args.side_priority else 0)
    # Consecutive batches are independent and write alternating output buffers, so their transform
    # launches alternate between two streams: the prologue of launch i+1 (launch latency, table
    # fill, first loads) overlaps the tail of launch i instead of idling the GPU.
    xf_streams = [stream, torch.cuda.Stream(dev)]
    dcode = N.dtype_code(wl["dtype"])
    cm_ptr = state.buffer[state.HEADER:].data_ptr()

    # In the timed region the launches alternate between two streams and the scoring kernels run beside them: there the
    # launches go WITHOUT programmatic dependent launch.  A pre-launched dependent grid waits at the head of the block
    # scheduler's queue for shared memory no SM can give it yet, and the small kernels of the scoring stream queue up
    # behind it (measured: 20 scoring launches done after 0.84 ms with it, 0.51 ms without; step 42.1 -> 41.1 us).
    # On ONE stream, back to back, it is worth 1.3 us per launch and stays on (kernel_ms_back_to_back).
    region_flags = bt.flags | N.FLAG_NO_PDL

    def step(i, streams=None, flags=None):
        j = i % nbuf
        s, o = srcs[j], outs[i & 1]
        t, l, f = params[j]
        st = (streams or xf_streams)[i & 1]
        rc = L.tb_transform_u8_nhwc_to_nchw(s.data_ptr(), B, H, W, 3, s.stride(0), s.stride(1), o.data_ptr(), dcode,
                                            crop, crop, t.data_ptr(), l.data_ptr(), f.data_ptr(), 0, 0, 0,
                                            lut.data_ptr(), region_flags if flags is None else flags, st.cuda_stream)
        if rc:
            N.check(rc)

    def score(i, st=None):
        j = i % nbuf
        rc = L.tb_metrics_update(logits[j].data_ptr(), N.TB_F32, B, classes, classes, labels[j].data_ptr(), k,
                                 state.buffer.data_ptr(), cm_ptr, (st or side).cuda_stream)
        if rc:
            N.check(rc)
        state.note_rows(B)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def fork():
        side.wait_stream(stream)
        xf_streams[1].wait_stream(stream)

    def join():
        stream.wait_stream(side)
        stream.wait_stream(xf_streams[1])

    fork()
    for i in range(Wm):
        step(i)
        score(i)
    join()
    if world > 1:
        state.all_reduce(transport=args.transport)
    sync_all()
    # isolated per-launch duration of the transform kernel: launches serialised on ONE stream,
    # each bracketed by CUDA events (the overlapped timed region below cannot bracket a single
    # launch: its start event would also time the wait for the other stream's kernel)
    iso = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
    for i, (a, b) in enumerate(iso):
        a.record(stream)
        step(i, streams=[stream, stream], flags=bt.flags)
        b.record(stream)
    sync_all()
    kernel_ms_isolated = sum(a.elapsed_time(b) for a, b in iso[4:]) / len(iso[4:])
    # ... and launched back to back on one stream without events in between (programmatic dependent launch overlaps
    # the next launch's set-up with the previous kernel's tail)
    b2b0, b2b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(4):
        step(i, streams=[stream, stream], flags=bt.flags)
    b2b0.record(stream)
    for i in range(40):
        step(i, streams=[stream, stream], flags=bt.flags)
    b2b1.record(stream)
    sync_all()
    kernel_ms_back_to_back = b2b0.elapsed_time(b2b1) / 40
    state.reset()

    # ---- CUDA graphs of the K steps: the launch-bound inner loop is captured once, so the host issues two graph
    # launches instead of 2K kernel launches (with 8 ranks sharing one host the eager loop was CPU-bound).  Two graphs:
    # the transform launches (two alternating streams) and the scoring launches (their own stream), so that the metric
    # exchange can start as soon as the last batch is scored, under the last transform launches.
    graph_x = graph_s = None
    if not args.no_graph and K <= 4096:
        try:
            cap = torch.cuda.Stream(dev)
            cap2 = torch.cuda.Stream(dev)
            cap.wait_stream(stream)
            graph_x = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph_x, stream=cap):
                cap2.wait_stream(cap)
                for i in range(K):
                    step(i, streams=[cap, cap2])
                cap.wait_stream(cap2)
            graph_s = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph_s, stream=cap):
                for i in range(K):
                    score(i, st=cap)
            stream.wait_stream(cap)
            sync_all()
            graph_x.replay()           # one untimed replay (graph upload)
            with torch.cuda.stream(side):
                graph_s.replay()
            sync_all()
            state.reset()
        except Exception as e:      # capture not possible on this driver: eager launches
            print(f"bench.py: CUDA graph capture failed ({e!r}); launching eagerly", file=sys.stderr)
            graph_x = graph_s = None
            torch.cuda.synchronize(dev)
            state.reset()

    # ---- timed region: K steps, device-resident inputs ---------------------------------------------
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev_x = torch.cuda.Event(enable_timing=True)          # last transform done
    ev_s = torch.cuda.Event(enable_timing=True)          # last batch scored (the exchange starts)
    ev_r = torch.cuda.Event(enable_timing=True)          # exchange done
    launches0 = L.tb_launch_count()
    with ClockSampler(local) as clocks:
        # barrier + device sync immediately before the start event: the ranks enter the region together
        sync_all()
        ev0.record(stream)
        clocks.active = True
        fork()
        if graph_x is not None:
            graph_x.replay()
            with torch.cuda.stream(side):
                graph_s.replay()
            state.note_rows(B * K)
        else:
            for i in range(K):
                step(i)
                score(i)
            stream.wait_stream(xf_streams[1])
        ev_x.record(stream)
        with torch.cuda.stream(side):
            ev_s.record(side)
            if world > 1:
                state.all_reduce(transport=args.transport)        # the one cross-GPU exchange, once per evaluation interval
            ev_r.record(side)
        stream.wait_stream(side)
        ev1.record(stream)
        sync_all()
        clocks.active = False
        clocks.in_region = len(clocks.samples)
        launches = L.tb_launch_count() - launches0 + (2 * K if graph_x is not None else 0)    # captured launches are counted at capture
        counts = state.compute()
        if clocks.in_region < 8:
            # the timed region was too short for the sampler: keep sampling under the same load, untimed
            clocks.active = True
            t_end = time.perf_counter() + 0.25
            i = 0
            while time.perf_counter() < t_end:
                fork()
                for _ in range(64):
                    step(i)
                    i += 1
                join()
                torch.cuda.synchronize(dev)
            clocks.active = False
    ms_total = ev0.elapsed_time(ev1)
    ck = clocks.summary()

    # ---- the exchange is checked against values no collective touched: every rank scores its own logits with torch ops
    # (argmax, the rank rule, a weighted checksum of the confusion bins); the per-rank figures are all-gathered and summed
    cls = torch.arange(classes, device=dev)
    wgt = ((cls[:, None] * 31 + cls[None, :] * 17) % 1009).to(torch.int64)
    mine = torch.zeros(5, dtype=torch.int64, device=dev)          # n, correct_1, correct_k, trace, weighted checksum
    for j in range(nbuf):
        uses = len(range(j, K, nbuf))
        lg, lb = logits[j], labels[j]
        pred = lg.argmax(-1)
        ly = lg.gather(1, lb[:, None])
        rank_y = (lg > ly).sum(-1) + ((lg == ly) & (cls[None, :] < lb[:, None])).sum(-1)
        hit1 = (pred == lb).sum()
        mine += uses * torch.stack([torch.tensor(B, device=dev), hit1, (rank_y < k).sum(), hit1, wgt[lb, pred].sum()])
    if world > 1:
        allm = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allm, mine)
        expect = torch.stack(allm).sum(0)
    else:
        expect = mine
    cm = counts["confusion"]
    got = torch.stack([torch.tensor(counts["n"], device=dev), torch.tensor(counts["correct_1"], device=dev),
                       torch.tensor(counts[f"correct_{k}"], device=dev), cm.diagonal().sum(), (cm * wgt).sum()])
    exchange_checked = bool(torch.equal(got, expect)) and int(cm.sum()) == counts["n"] == world * B * K
    if not exchange_checked:
        raise SystemExit(f"bench.py: reduced metric state {got.tolist()} differs from the independently computed {expect.tolist()}")

    per_rank = {"ms_per_step_transform": [round(ev0.elapsed_time(ev_x) / K, 5)], "scoring_done_ms": [round(ev0.elapsed_time(ev_s), 3)]}
    if world > 1:
        mine_t = torch.tensor([ev0.elapsed_time(ev_x), ck["sm_mhz"] or 0.0, ev_s.elapsed_time(ev_r), ev0.elapsed_time(ev_s)],
                              dtype=torch.float64, device=dev)
        allr = [torch.zeros_like(mine_t) for _ in range(world)]
        dist.all_gather(allr, mine_t)
        per_rank = {"ms_per_step_transform": [round(float(a[0]) / K, 5) for a in allr],
                    "sm_mhz": [float(a[1]) for a in allr], "exchange_ms": [round(float(a[2]), 3) for a in allr],
                    "scoring_done_ms": [round(float(a[3]), 3) for a in allr]}
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * B * K / (ms_total * 1e-3)
    # One transform launch per step, launches back to back on two alternating streams: the
    # average launch duration over the timed region is the region's time per step (this charges
    # the kernel with every bubble and with the scoring kernel that shares the SMs).
    kernel_ms = ms_total / K

    peak, peak_src = measured_peak()
    gbs = lambda ms: wl["alg_bytes"] * B / (ms * 1e-3) / 1e9
    achieved = gbs(kernel_ms)
    roofline = {"bound": "hbm", "kernel": "tb_xf_generic_kernel" if bt.flags & N.FLAG_FORCE_GENERIC else "tb_xf_stream_kernel",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "frac_of_nominal_8000": achieved / 8000.0, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": wl["alg_bytes"] * B, "kernel_ms": kernel_ms,
                "kernel_ms_isolated": kernel_ms_isolated, "frac_isolated": gbs(kernel_ms_isolated) / peak,
                "kernel_ms_back_to_back": kernel_ms_back_to_back, "frac_back_to_back": gbs(kernel_ms_back_to_back) / peak,
                "method": "kernel_ms = CUDA-event time of the K-step timed region / K launches (launches overlap across two "
                          "streams, so single launches cannot be bracketed there); kernel_ms_isolated = mean CUDA-event bracket "
                          "of 16 serialised launches on one stream (each bracket also holds one launch latency); "
                          "kernel_ms_back_to_back = 40 launches on one stream between two events",
                "traffic": ncu_traffic(args.workload)}

    # ---- e2e: pinned host buffers -> public API -> counters read back, every step -------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, wl, dev, world, bt, spec, rank)

    # ---- every BASELINE configuration, per kernel launch ---------------------------------------------
    del srcs, outs, logits, labels
    torch.cuda.empty_cache()
    configs = None
    if not args.no_configs:
        configs = run_configs(dev, rank, world, with_cpu=not args.no_cpu, transport=args.transport)

    # ---- CPU baseline (rank 0, N=1) -----------------------------------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu:
        workers = min(host_threads(), 64)
        torch.set_num_threads(max(1, host_threads()))
        ips, done, dt = cpu_path("c2", B, H, W, steps=100000, workers=workers, max_seconds=12.0, score=(classes, k),
                                 out_dtype=wl["dtype"])
        import torchvision
        cpu = {"value": ips, "unit": "images/s", "cores": workers, "kind": "port",
               "sample": f"{done} batches of {B} images in {dt:.1f} s after 8 s of untimed ones (workers out of phase, sustained rate); torchvision "
                         f"{torchvision.__version__} PIL transforms in {workers} DataLoader workers + accuracy/topk/bincount on the host; {cpu_model()}"}

    if rank == 0:
        line = {
            "metric": "images/sec transformed+scored", "value": value, "unit": "images/s", "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": wl["desc"], "per_gpu_batch": B, "output_dtype": str(wl["dtype"]).replace("torch.", ""),
                       "l2": f"inputs rotate over {nbuf} distinct batches ({nbuf * in_bytes >> 20} MiB) > 126 MiB L2; "
                             "each output (147 MiB) exceeds L2",
                       "exchange": (f"one all_reduce(SUM) of the packed metric state per K steps ({args.transport} transport: "
                                    f"{'4' if args.transport == 'int32' else '8'} MB), issued on the scoring stream as soon as the last "
                                    "batch is scored, inside the timed region") if world > 1 else "none (1 GPU)",
                       "streams": "transform launches alternate between two streams (independent batches, alternating "
                                  "output buffers; launched without programmatic dependent launch, see bench.py), scoring on a "
                                  "third; all joined before the stop event",
                       "launch": (f"two CUDA graphs ({K} transform launches, {K} scoring launches), one replay each"
                                  if graph_x is not None else "eager launches"),
                       "variant": args.variant,
                       "cpu_affinity": (f"rank pinned to its GPU's {affinity} local CPUs (NVML) before pinned buffers are allocated"
                                        if affinity else "not set")},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": ck, "per_rank": per_rank, "exchange_checked": exchange_checked,
            "counts": {"n": counts["n"], "correct_1": counts["correct_1"], f"correct_{k}": counts[f"correct_{k}"]},
            "configs": configs,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def h2d_probe(dev, world, host, nbytes, iters=10, repeats=3):
    """What this box can move host -> device: every rank copies `nbytes` of (already pinned, already touched) host
    memory to its GPU `iters` times at the same moment (plain cudaMemcpyAsync through torch, the copy engine); best of
    `repeats` rounds.  Returns aggregate GB/s over all ranks."""
    import torch.distributed as dist
    src = host.view(-1)[:nbytes]
    dst = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    best = None
    for _ in range(repeats + 1):            # the first round is a warm-up
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(iters):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        best = dt if best is None else min(best, dt)
    return world * nbytes * iters / best / 1e9


def run_e2e(args, wl, dev, world, bt, spec, rank):
    """The calls a user makes, starting from HOST memory every step: the crop / flip draw for the step (replayed torch
    generator), BatchTransform(pinned uint8 batch, params) -- the kernel reads the pinned batch in place over PCIe (only
    the crop rows cross the bus; 8 CTAs) --, logits/labels copied host->device, MetricState.update, and a
    device->host read of the counters.  Steps alternate between two streams so that PCIe never
    idles; the host blocks on step i-1's counters right after enqueuing step i."""
    import torch.distributed as dist
    from torchbooster_b200.metrics import MetricState
    from torchbooster_b200.transforms import CropFlipParams

    B, H, W, crop, classes, k = wl["B"], wl["H"], wl["W"], wl["crop"], wl["classes"], wl["k"]
    K, Wm = args.steps, max(args.warmup, 3)
    nh = 3
    g = torch.Generator().manual_seed(99 + rank)
    host_imgs = [torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, generator=g).pin_memory() for _ in range(nh)]
    host_logits = [torch.randn(B, classes, generator=g).pin_memory() for _ in range(nh)]
    host_labels = [torch.randint(0, classes, (B,), generator=g).pin_memory() for _ in range(nh)]
    rng = CropFlipParams(42 + rank)
    state = MetricState(classes, k=k, device=dev)
    streams = [torch.cuda.Stream(dev), torch.cuda.Stream(dev)]
    outs = [torch.empty((B, 3, crop, crop), dtype=wl["dtype"], device=dev) for _ in range(2)]
    dev_logits = [torch.empty((B, classes), device=dev) for _ in range(2)]
    dev_labels = [torch.empty((B,), dtype=torch.int64, device=dev) for _ in range(2)]
    results = [torch.zeros(4, dtype=torch.int64).pin_memory() for _ in range(2)]
    done = [torch.cuda.Event(), torch.cuda.Event()]

    def loop(n):
        for i in range(n):
            s, j = i & 1, i % nh
            prm = rng.draw(B, H, W, (crop, crop), (0, 0), 0.5)               # this step's crop / flip decisions
            with torch.cuda.stream(streams[s]):
                dev_logits[s].copy_(host_logits[j], non_blocking=True)
                dev_labels[s].copy_(host_labels[j], non_blocking=True)
                bt(host_imgs[j], params=prm, out=outs[s])                 # pinned batch, read in place
                state.update(dev_logits[s], dev_labels[s])
                results[s].copy_(state.counters, non_blocking=True)       # device -> host read of the metric
                done[s].record()
            if i:
                done[1 - s].synchronize()                                 # block on the previous step's counters
        if n:
            done[(n - 1) & 1].synchronize()

    for st in streams:
        st.wait_stream(torch.cuda.current_stream(dev))
    loop(Wm)
    torch.cuda.synchronize(dev)
    state.reset()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    loop(K)
    if world > 1:
        torch.cuda.current_stream(dev).wait_stream(streams[0])
        torch.cuda.current_stream(dev).wait_stream(streams[1])
        state.all_reduce(transport=args.transport)
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    # (the counters are a running total shared by both streams: a read is a snapshot, the total is exact)
    assert int(state.counters[0].item()) == world * B * K, (results, state.counters)
    assert max(int(results[0][0]), int(results[1][0])) >= B * (K - 1)
    window = B * crop * (((crop * 3 + 15 + 15) // 16) * 16)       # bytes the bulk copies pull over PCIe per step
    h2d = window + B * classes * 4 + B * 8 + 9 * B
    # what the box can move: N ranks copying the same number of bytes host -> device at once (untimed probe)
    roof = h2d_probe(dev, world, host_imgs[0], h2d)
    moved = world * h2d * K / dt / 1e9
    return {"value": world * B * K / dt, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 32,
            "ms_per_step": dt / K * 1e3, "h2d_gbs": moved, "h2d_roofline_gbs": roof, "frac_of_h2d_roofline": moved / roof,
            "h2d_roofline_method": f"{world} rank(s) x 10 plain pinned->device copies of {h2d} bytes at the same moment, wall clock, max over ranks, best of 3 rounds",
            "api": "CropFlipParams.draw + BatchTransform.__call__(pinned uint8 batch, params): the kernel reads the pinned batch in "
                   "place (zero-copy, crop rows only, 8 CTAs) + H2D of logits/labels + MetricState.update + counters read "
                   "back every step; two streams, one step of lookahead"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--variant", default="stream", choices=["stream", "generic"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--side-priority", type=int, default=1, help="1: the scoring / exchange stream has high priority")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-configuration block (C1 .. C5)")
    ap.add_argument("--transport", default="int64", choices=["int32", "int64"],
                    help="wire format of the metric exchange (int32: half the bytes, exact while the ranks together have scored "
                         "< 2^31 rows; measured no faster than int64 at 2-8 ranks: the exchange is latency-bound)")
    ap.add_argument("--no-affinity", action="store_true", help="do not pin the rank to its GPU's local CPUs")
    ap.add_argument("--configs-only", default="", help="development: time only these configurations (comma separated) and print them")
    args = ap.parse_args()
    if args.configs_only:
        torch.cuda.set_device(0)
        res = run_configs(torch.device("cuda", 0), 0, 1, with_cpu=False, transport=args.transport, only=args.configs_only.split(","))
        print(json.dumps(res), flush=True)
        return
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
