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

WORKLOADS = {
    # name: (B, H, W, C, crop, flip_p, out dtype, description)
    "c2": dict(B=512, H=256, W=256, crop=224, dtype=torch.bfloat16, classes=1000, k=5,
               alg_bytes=224 * 224 * 3 + 224 * 224 * 3 * 2,
               desc="C2 ImageNet-shaped synthetic batch: 512x256x256x3 u8 -> random crop 224, flip, "
                    "normalize, bf16 NCHW; scored: 512x1000 fp32 logits -> top-1/top-5/confusion"),
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


def cpu_path(wl: dict, steps: int, warmup: int, workers: int, max_seconds: float = 1e9):
    """Reference-style CPU path: per-sample torchvision PIL transforms in DataLoader workers
    (resnet.py:94-107 pattern), default collate to fp32 NCHW, then accuracy (the oracle's
    restatement of torchbooster/metrics.py:27) + torch.topk + torch.bincount on all host threads.
    Returns (images_per_second, steps_done, seconds)."""
    import torchvision.transforms as T
    from oracle import metrics_oracle as MO
    B = wl["B"]
    rng = np.random.default_rng(1234)
    pool = rng.integers(0, 256, (min(2048, B * 2), wl["H"], wl["W"], 3), dtype=np.uint8)
    tf = T.Compose([T.RandomCrop(wl["crop"]), T.RandomHorizontalFlip(), T.ToTensor(),
                    T.Normalize(IMAGENET_MEAN, IMAGENET_STD)])
    total = (steps + warmup) * B
    ds = _SyntheticImages(pool, tf)
    sampler = torch.utils.data.RandomSampler(ds, replacement=True, num_samples=total)
    loader = torch.utils.data.DataLoader(ds, batch_size=B, sampler=sampler, num_workers=workers, pin_memory=False,
                                         drop_last=True, persistent_workers=False,
                                         prefetch_factor=2 if workers else None)
    gen = torch.Generator().manual_seed(1234)
    logits = torch.randn(B, wl["classes"], generator=gen)
    cm = torch.zeros(wl["classes"] ** 2, dtype=torch.int64)
    done, t0, n1, nk = 0, None, 0, 0
    for i, (x, y) in enumerate(loader):
        if i == warmup:
            t0 = time.perf_counter()
        x = x.to(torch.bfloat16) if wl["dtype"] == torch.bfloat16 else x
        n1 += int(MO.accuracy(logits, y) * B)
        nk += int((logits.topk(wl["k"], -1).indices == y[:, None]).any(-1).sum())
        cm += torch.bincount(y * wl["classes"] + logits.argmax(-1), minlength=wl["classes"] ** 2)
        if i >= warmup:
            done += 1
            if time.perf_counter() - t0 > max_seconds:
                break
    dt = time.perf_counter() - t0
    del loader
    return done * B / dt, done, dt


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
    wl = dict(wl)
    full_b = wl["B"]
    if args.steps * full_b > 1_000_000:
        wl["B"] = max(16, (1_000_000 // args.steps) // 16 * 16)
    # untimed warm-up covers one full round of worker prefetch, so that timed steps are steady state
    ips, done, dt = cpu_path(wl, args.steps, max(args.warmup, workers + 2), workers)
    line = {
        "impl": "reference", "metric": "images/sec transformed+scored", "value": ips, "unit": "images/s",
        "n_gpus": args.gpus, "steps": done, "warmup": args.warmup, "ms_per_step": dt / max(done, 1) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": wl["desc"],
                   "per_step": f"one batch of {wl['B']} images on the host CPU" +
                               ("" if wl["B"] == full_b else f" (bounded sample of the {full_b}-image batch)")},
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": workers, "kind": "port",
                         "sample": f"{done} batches of {wl['B']} images; torchvision {__import__('torchvision').__version__} "
                                   f"PIL transforms in {workers} DataLoader workers + accuracy/topk/bincount; {cpu_model()}"},
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
    sp = stream.cuda_stream
    # Scoring does not depend on the transform of the same step (in training the logits come from
    # the model), so it runs on a second stream and overlaps the transform kernel.
    side = torch.cuda.Stream(dev)
    sp2 = side.cuda_stream
    # Consecutive batches are independent and write alternating output buffers, so their transform
    # launches alternate between two streams: the prologue of launch i+1 (launch latency, table
    # fill, first loads) overlaps the tail of launch i instead of idling the GPU (measured: 45.0 us
    # -> 39.8 us per 512-image batch, profiles/r1_configs.jsonl "c2_abi_2stream").
    xf_streams = [stream, torch.cuda.Stream(dev)]
    dcode = N.dtype_code(wl["dtype"])
    cm_ptr = state.buffer[state.HEADER:].data_ptr()

    def step(i):
        j = i % nbuf
        s, o = srcs[j], outs[i & 1]
        t, l, f = params[j]
        rc = L.tb_transform_u8_nhwc_to_nchw(s.data_ptr(), B, H, W, 3, s.stride(0), s.stride(1), o.data_ptr(), dcode,
                                            crop, crop, t.data_ptr(), l.data_ptr(), f.data_ptr(), 0, 0, 0,
                                            lut.data_ptr(), bt.flags, xf_streams[i & 1].cuda_stream)
        if rc:
            N.check(rc)

    def score(i):
        j = i % nbuf
        rc = L.tb_metrics_update(logits[j].data_ptr(), N.TB_F32, B, classes, classes, labels[j].data_ptr(), k,
                                 state.buffer.data_ptr(), cm_ptr, sp2)
        if rc:
            N.check(rc)

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
        state.all_reduce()
    sync_all()
    # isolated per-launch duration of the transform kernel: launches serialised on ONE stream,
    # each bracketed by CUDA events (the overlapped timed region below cannot bracket a single
    # launch: its start event would also time the wait for the other stream's kernel)
    iso = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
    saved = xf_streams[1]
    xf_streams[1] = stream
    for i, (a, b) in enumerate(iso):
        a.record(stream)
        step(i)
        b.record(stream)
    xf_streams[1] = saved
    sync_all()
    kernel_ms_isolated = sum(a.elapsed_time(b) for a, b in iso[4:]) / len(iso[4:])
    state.reset()

    # ---- CUDA graph of G steps: the launch-bound inner loop is captured once and replayed, so the
    # host issues one graph launch per G steps instead of 2G kernel launches (with 8 ranks sharing
    # one host the eager loop was CPU-bound: 52.9 us/step against 41.8 us at one rank).
    G = 2 * nbuf                       # every (input buffer, output buffer, stream) combination once
    graph = None
    if not args.no_graph and K >= G:
        try:
            cap = torch.cuda.Stream(dev)
            cap.wait_stream(stream)
            graph = torch.cuda.CUDAGraph()
            saved_streams = (xf_streams[0], sp2)
            with torch.cuda.graph(graph, stream=cap):
                xf_streams[0] = cap
                side.wait_stream(cap)
                xf_streams[1].wait_stream(cap)
                for i in range(G):
                    step(i)
                    score(i)
                cap.wait_stream(side)
                cap.wait_stream(xf_streams[1])
            xf_streams[0] = saved_streams[0]
            sync_all()
            state.reset()
        except Exception as e:      # capture not possible on this driver: eager launches
            print(f"bench.py: CUDA graph capture failed ({e!r}); launching eagerly", file=sys.stderr)
            graph = None
            xf_streams[0] = stream
            torch.cuda.synchronize(dev)
            state.reset()
    if graph is not None:           # one untimed replay (graph upload)
        graph.replay()
        sync_all()
        state.reset()

    # ---- timed region: K steps, device-resident inputs ---------------------------------------------
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev_mid = torch.cuda.Event(enable_timing=True)
    launches0 = L.tb_launch_count()
    with ClockSampler(local) as clocks:
        # barrier + device sync immediately before the start event: the ranks enter the region
        # together (the sampler's NVML start-up used to skew them by milliseconds, which the final
        # all_reduce then charged to every rank)
        sync_all()
        ev0.record(stream)
        clocks.active = True
        n_graph = K // G if graph is not None else 0
        for _ in range(n_graph):
            graph.replay()
        fork()
        for i in range(n_graph * G, K):
            step(i)
            score(i)
        join()
        ev_mid.record(stream)
        if world > 1:
            state.all_reduce()        # the one cross-GPU exchange, once per evaluation interval
        ev1.record(stream)
        sync_all()
        clocks.active = False
        clocks.in_region = len(clocks.samples)
        launches = L.tb_launch_count() - launches0 + n_graph * 2 * G     # captured launches are counted once, at capture
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
                    score(i)
                    i += 1
                join()
                torch.cuda.synchronize(dev)
            clocks.active = False     # captured launches are counted once, at capture
    ms_total = ev0.elapsed_time(ev1)
    ck = clocks.summary()
    per_rank = None
    if world > 1:
        mine = torch.tensor([ev0.elapsed_time(ev_mid), ck["sm_mhz"] or 0.0, ev_mid.elapsed_time(ev1)],
                            dtype=torch.float64, device=dev)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = {"ms_per_step_before_exchange": [round(float(a[0]) / K, 5) for a in allr],
                    "sm_mhz": [float(a[1]) for a in allr], "exchange_ms": [round(float(a[2]), 3) for a in allr]}
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * B * K / (ms_total * 1e-3)
    # One transform launch per step, launches back to back on two alternating streams: the
    # average launch duration over the timed region is the region's time per step (this charges
    # the kernel with every bubble and with the scoring kernel that shares the SMs).
    kernel_ms = ms_total / K
    assert counts["n"] == world * B * K, (counts["n"], world * B * K)

    peak, peak_src = measured_peak()
    achieved = wl["alg_bytes"] * B / (kernel_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "tb_xf_stream_kernel" if bt.flags == 0 else "tb_xf_generic_kernel",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "frac_of_nominal_8000": achieved / 8000.0, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": wl["alg_bytes"] * B, "kernel_ms": kernel_ms,
                "kernel_ms_isolated": kernel_ms_isolated,
                "frac_isolated": wl["alg_bytes"] * B / (kernel_ms_isolated * 1e-3) / 1e9 / peak,
                "method": "kernel_ms = CUDA-event time of the K-step timed region / K launches (launches overlap "
                          "across two streams, so single launches cannot be bracketed there); kernel_ms_isolated = "
                          "mean CUDA-event bracket of 16 serialised launches on one stream after warm-up",
                "traffic": ncu_traffic(args.workload)}

    # ---- e2e: pinned host buffers -> public API -> counters read back, every step -------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, wl, dev, world, bt, spec, rank)

    # ---- CPU baseline (rank 0, N=1) -----------------------------------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu:
        workers = min(host_threads(), 64)
        torch.set_num_threads(max(1, host_threads()))
        ips, done, dt = cpu_path(wl, steps=max(80, 12 * workers), warmup=workers + 2, workers=workers,
                                 max_seconds=20.0)
        import torchvision
        cpu = {"value": ips, "unit": "images/s", "cores": workers, "kind": "port",
               "sample": f"{done} batches of {B} images in {dt:.1f} s; torchvision {torchvision.__version__} PIL "
                         f"transforms in {workers} DataLoader workers + accuracy/topk/bincount on the host; {cpu_model()}"}

    if rank == 0:
        line = {
            "metric": "images/sec transformed+scored", "value": value, "unit": "images/s", "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": wl["desc"], "per_gpu_batch": B, "output_dtype": str(wl["dtype"]).replace("torch.", ""),
                       "l2": f"inputs rotate over {nbuf} distinct batches ({nbuf * in_bytes >> 20} MiB) > 126 MiB L2; "
                             "each output (147 MiB) exceeds L2",
                       "exchange": "one all_reduce(SUM) of the packed int64 metric buffer (8 MB) per K steps, inside the "
                                   "timed region" if world > 1 else "none (1 GPU)",
                       "streams": "transform launches alternate between two streams (independent batches, alternating "
                                  "output buffers), scoring on a third; all joined before the stop event",
                       "launch": (f"CUDA graph of {G} steps replayed {K // G} times + {K % G} eager steps"
                                  if graph is not None else "eager launches"),
                       "variant": args.variant,
                       "cpu_affinity": (f"rank pinned to its GPU's {affinity} local CPUs (NVML) before pinned buffers are allocated"
                                        if affinity else "not set")},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": ck, "per_rank": per_rank,
            "counts": {"n": counts["n"], "correct_1": counts["correct_1"], f"correct_{k}": counts[f"correct_{k}"]},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_e2e(args, wl, dev, world, bt, spec, rank):
    """The calls a user makes, starting from HOST memory every step: BatchTransform(pinned uint8
    batch, params) -- the kernel reads the pinned batch in place over PCIe (only the crop rows
    cross the bus; 8 CTAs) --, logits/labels copied host->device, MetricState.update, and a
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
    host_params = [rng.draw(B, H, W, (crop, crop), (0, 0), 0.5) for _ in range(nh)]
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
            with torch.cuda.stream(streams[s]):
                dev_logits[s].copy_(host_logits[j], non_blocking=True)
                dev_labels[s].copy_(host_labels[j], non_blocking=True)
                bt(host_imgs[j], params=host_params[j], out=outs[s])      # pinned batch, read in place
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
        state.all_reduce()
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t.item())
    # (the counters are a running total shared by both streams: a read is a snapshot, the total is exact)
    assert world > 1 or int(state.counters[0].item()) == B * K, (results, state.counters)
    assert max(int(results[0][0]), int(results[1][0])) >= B * (K - 1)
    window = B * crop * (((crop * 3 + 15 + 15) // 16) * 16)       # bytes the bulk copies pull over PCIe per step
    h2d = window + B * classes * 4 + B * 8 + 9 * B
    return {"value": world * B * K / dt, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 32,
            "ms_per_step": dt / K * 1e3,
            "api": "BatchTransform.__call__(pinned uint8 batch, params): the kernel reads the pinned batch in place "
                   "(zero-copy, crop rows only, 8 CTAs) + H2D of logits/labels + MetricState.update + counters read "
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
    ap.add_argument("--no-affinity", action="store_true", help="do not pin the rank to its GPU's local CPUs")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
