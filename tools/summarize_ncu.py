#!/usr/bin/env python
"""Turns the ncu outputs a gpurun call brought back (gpurun_out/) into the small text summaries
committed under profiles/.  Runs in the build container (ncu -i needs no GPU)."""
import csv
import io
import json
import subprocess
import sys
from collections import OrderedDict
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
OUT = ROOT / "profiles"
KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed.avg.per_cycle_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
    "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]


def raw_rows(rep):
    txt = subprocess.run(["ncu", "-i", str(rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    return hdr, units, rows[2:]


def summarize_report(rep, out, title):
    hdr, units, rows = raw_rows(rep)
    lines = [f"# {title}", f"# source: {rep.name} (ncu --set full --clock-control none --import-source on)", ""]
    traffic = OrderedDict()
    for r in rows:
        d = dict(zip(hdr, r))
        name = d.get("Kernel Name", "?")
        lines.append(f"kernel: {name}")
        for k in KEYS:
            if k in d:
                lines.append(f"  {k:75s} {d[k]:>16s} {units[hdr.index(k)]}")
        try:
            to_b = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            rd = float(d["dram__bytes_read.sum"]) * to_b[units[hdr.index("dram__bytes_read.sum")]]
            wr = float(d["dram__bytes_write.sum"]) * to_b[units[hdr.index("dram__bytes_write.sum")]]
            lines.append(f"  {'dram traffic (read + write) per launch':75s} {rd + wr:16.0f} byte")
            traffic.setdefault(name, []).append(rd + wr)
        except Exception:
            pass
        lines.append("")
    out.write_text("\n".join(lines))
    return traffic


def summarize_launches(csv_path, out):
    rows = [r for r in csv.reader(open(csv_path)) if len(r) > 10 and r[0].isdigit()]
    per = OrderedDict()
    total = 0.0
    for r in rows:
        name, ns = r[4], float(r[-1])
        per.setdefault(name, []).append(ns)
        total += ns
    lines = ["# ncu --metrics gpu__time_duration.sum --clock-control none  (python bench.py --steps 20 --warmup 5 --no-cpu --no-e2e --no-configs)",
             "# per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes", "",
             f"{'kernel':90s} {'launches':>8s} {'mean_us':>9s} {'total_us':>10s} {'share':>7s}"]
    for name, v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
        lines.append(f"{name[:90]:90s} {len(v):8d} {sum(v) / len(v) / 1e3:9.2f} {sum(v) / 1e3:10.1f} {sum(v) / total:7.1%}")
    out.write_text("\n".join(lines) + "\n")


SASS_OPS = ["UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "IDP", "REDUX", "HSET2", "HMNMX2", "VHMNMX", "FMNMX3", "LDGSTS", "UTCHMMA", "HMMA", "LDTM"]


def sass_census(out):
    """Opcode census of the built library: which Blackwell paths the kernels really use (cuobjdump -sass)."""
    lib = ROOT / "torchbooster_b200" / "lib" / "libtorchbooster_b200.so"
    txt = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
    per, cur = OrderedDict(), None
    for line in txt.splitlines():
        line = line.strip()
        if line.startswith("Function :"):
            cur = line.split(":", 1)[1].strip()
            per[cur] = OrderedDict((op, 0) for op in SASS_OPS)
        elif cur and line.startswith("/*") and "*/" in line:
            body = line.split("*/", 1)[1].strip()
            toks = body.split()
            if not toks:
                continue
            op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
            base = op.split(".")[0]
            if base in per[cur]:
                per[cur][base] += 1
    lines = ["# SASS opcode census of torchbooster_b200/lib/libtorchbooster_b200.so (cuobjdump -sass, sm_100a)",
             "# UBLKCP = cp.async.bulk (1-D bulk copy, TMA engine); UTMALDG / UTMASTG = cp.async.bulk.tensor; SYNCS = mbarrier;",
             "# IDP = DP4A; REDUX = redux.sync; HSET2 / HMNMX2 / VHMNMX = packed 16-bit compare / max; no tensor-core opcode is",
             "# expected (nothing on this path is a contraction).", "",
             f"{'function':100s} " + " ".join(f"{op:>7s}" for op in SASS_OPS)]
    tot = OrderedDict((op, 0) for op in SASS_OPS)
    for fn, c in per.items():
        if any(c.values()):
            lines.append(f"{fn[:100]:100s} " + " ".join(f"{c[op]:7d}" for op in SASS_OPS))
        for op in SASS_OPS:
            tot[op] += c[op]
    lines.append(f"{'TOTAL over ' + str(len(per)) + ' functions':100s} " + " ".join(f"{tot[op]:7d}" for op in SASS_OPS))
    out.write_text("\n".join(lines) + "\n")


CAPTURES = [
    # report stem, output name, title, traffic key
    ("prof_c1", "ncu_c1_crop_generic", "C1: tb_xf_generic_kernel, 256 x 32x32 reflect-pad crop / flip -> fp32", "c1"),
    ("prof_c2", "ncu_c2_transform", "C2: tb_xf_stream_kernel, 512 x 256x256 -> crop 224, flip, bf16 (one stream, back to back)", "c2"),
    ("prof_c3", "ncu_c3_transform", "C3: tb_xf_stream_kernel, 1024 x 64x64 flip -> fp32", "c3"),
    ("prof_c4", "ncu_c4_resize_b32", "C4: tb_resize_ring_kernel, 32 x 512x512 -> 256x256 fp32", "c4"),
    ("prof_c4_b256", "ncu_c4_resize_b256", "C4 at batch 256: tb_resize_ring_kernel", "c4_b256"),
    ("prof_c5", "ncu_c5_metrics_f32", "C5: tb_metrics_row_kernel, 125 000 x 1000 fp32", "c5"),
    ("prof_c5_bf16", "ncu_c5_metrics_bf16", "C5 bf16: tb_metrics_row_kernel, 125 000 x 1000 bf16", "c5_bf16"),
    ("prof_rrc", "ncu_resized_crop", "resized-crop kernel, 512 x (256x256 box -> 224x224, bf16)", "rrc_c2_b512"),
    ("prof_aug", "ncu_augment", "augment kernel, CIFAR training transform, batch 2048", "aug_cifar_b2048"),
]


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
    src = ROOT / "gpurun_out"
    OUT.mkdir(exist_ok=True)
    traffic = {}
    if (src / "launches.csv").exists():
        summarize_launches(src / "launches.csv", OUT / f"{tag}_launches.txt")
    for stem, name, title, key in CAPTURES:
        rep = src / f"{stem}.ncu-rep"
        if rep.exists():
            t = summarize_report(rep, OUT / f"{tag}_{name}.txt", title)
            for _, v in t.items():
                traffic[key] = sum(v) / len(v)
    if traffic:
        old = json.loads((OUT / "traffic.json").read_text()) if (OUT / "traffic.json").exists() else {}
        old.update(traffic)
        (OUT / "traffic.json").write_text(json.dumps(old, indent=1) + "\n")
    for f in ("bench_driver.json", "bench_ref.json"):
        if (src / f).exists() and (src / f).stat().st_size:
            (OUT / f"{tag}_{f}").write_text((src / f).read_text())
    sass_census(OUT / f"{tag}_sass_census.txt")


if __name__ == "__main__":
    main()
