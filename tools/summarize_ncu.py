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
    lines = ["# ncu --metrics gpu__time_duration.sum --clock-control none  (python bench.py --steps 40 --warmup 3 --no-cpu --no-e2e)",
             "# per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes", "",
             f"{'kernel':90s} {'launches':>8s} {'mean_us':>9s} {'total_us':>10s} {'share':>7s}"]
    for name, v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
        lines.append(f"{name[:90]:90s} {len(v):8d} {sum(v) / len(v) / 1e3:9.2f} {sum(v) / 1e3:10.1f} {sum(v) / total:7.1%}")
    out.write_text("\n".join(lines) + "\n")


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
    src = ROOT / "gpurun_out"
    OUT.mkdir(exist_ok=True)
    traffic = {}
    if (src / "launches.csv").exists():
        summarize_launches(src / "launches.csv", OUT / f"{tag}_launches.txt")
    if (src / "prof_xf.ncu-rep").exists():
        t = summarize_report(src / "prof_xf.ncu-rep", OUT / f"{tag}_ncu_transform.txt", "transform kernel, C2 workload")
        for name, v in t.items():
            if "tb_xf_stream" in name:
                traffic["c2"] = sum(v) / len(v)
    if (src / "prof_mx.ncu-rep").exists():
        t = summarize_report(src / "prof_mx.ncu-rep", OUT / f"{tag}_ncu_metrics.txt", "metrics kernel, C5 workload (125 000 x 1000 fp32)")
        for name, v in t.items():
            traffic["c5"] = sum(v) / len(v)
    if (src / "prof_rs.ncu-rep").exists():
        t = summarize_report(src / "prof_rs.ncu-rep", OUT / f"{tag}_ncu_resize.txt", "resize kernel, C4 workload (batch 256)")
        for name, v in t.items():
            traffic["c4_b256"] = sum(v) / len(v)
    if (src / "prof_rrc.ncu-rep").exists():
        t = summarize_report(src / "prof_rrc.ncu-rep", OUT / f"{tag}_ncu_resized_crop.txt",
                             "resized-crop kernel, 512 x (256x256 box -> 224x224, bf16), per-sample boxes, taps computed in-kernel")
        for name, v in t.items():
            traffic["rrc_c2_b512"] = sum(v) / len(v)
    if (src / "prof_aug.ncu-rep").exists():
        t = summarize_report(src / "prof_aug.ncu-rep", OUT / f"{tag}_ncu_augment.txt",
                             "augment kernel, CIFAR training transform (resnet.py:96-103), batch 2048")
        for name, v in t.items():
            traffic["aug_cifar_b2048"] = sum(v) / len(v)
    if traffic:
        old = json.loads((OUT / "traffic.json").read_text()) if (OUT / "traffic.json").exists() else {}
        old.update(traffic)
        (OUT / "traffic.json").write_text(json.dumps(old, indent=1) + "\n")
    for f in ("bench_full.json", "bench_ref.json"):
        if (src / f).exists():
            (OUT / f"{tag}_{f}").write_text((src / f).read_text())
    if (src / "configs.jsonl").exists():        # merge by case name: a partial run must not drop the other cases
        dst = OUT / f"{tag}_configs.jsonl"
        rows = OrderedDict()
        for path in (dst, src / "configs.jsonl"):
            if path.exists():
                for line in path.read_text().splitlines():
                    if line.strip():
                        rows[json.loads(line)["case"]] = line
        dst.write_text("\n".join(rows.values()) + "\n")


if __name__ == "__main__":
    main()
