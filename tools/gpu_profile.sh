#!/bin/bash
# Bench + ncu evidence for profiles/: launch list and full captures of the three kernels.
set -u
mkdir -p gpurun_out
BENCH="python bench.py --steps 40 --warmup 3 --no-cpu --no-e2e"
echo "== bench (full)"
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/bench_full.json; tail -3 gpurun_out/bench_full.err
echo "== reference arm"
timeout 600 python bench.py --impl reference --steps 60 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-300 gpurun_out/bench_ref.json
echo "== ncu launch list"
timeout 300 $BENCH > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launch.log 2>&1; echo "launch list rc=$?"
echo "== ncu full: transform"
timeout 300 $BENCH > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tb_xf_stream -s 5 -c 2 -f -o gpurun_out/prof_xf $BENCH > gpurun_out/ncu_xf.log 2>&1; echo "ncu xf rc=$?"
echo "== ncu full: metrics"
CFG="python tools/bench_configs.py --only c5 --case c5_fp32_125k"
timeout 300 $CFG > gpurun_out/plain3.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tb_metrics_kernel -s 10 -c 1 -f -o gpurun_out/prof_mx $CFG > gpurun_out/ncu_mx.log 2>&1; echo "ncu mx rc=$?"
echo "== ncu full: resize"
CFG="python tools/bench_configs.py --only c4 --case c4_stt_b256"
timeout 300 $CFG > gpurun_out/plain4.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tb_resize -s 8 -c 1 -f -o gpurun_out/prof_rs $CFG > gpurun_out/ncu_rs.log 2>&1; echo "ncu rs rc=$?"
echo "== all configs"
timeout 900 python tools/bench_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
cut -c1-200 gpurun_out/configs.jsonl
