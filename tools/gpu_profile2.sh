#!/bin/bash
set -u
mkdir -p gpurun_out
echo "== ncu full: metrics"
CFG="python tools/bench_configs.py --only c5 --case c5_fp32_125k"
timeout 300 $CFG > gpurun_out/plain3.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tb_metrics_kernel -s 10 -c 1 -f -o gpurun_out/prof_mx $CFG > gpurun_out/ncu_mx.log 2>&1; echo "ncu mx rc=$?"
echo "== ncu full: resize"
CFG="python tools/bench_configs.py --only c4 --case c4_stt_b256"
timeout 300 $CFG > gpurun_out/plain4.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tb_resize -s 8 -c 1 -f -o gpurun_out/prof_rs $CFG > gpurun_out/ncu_rs.log 2>&1; echo "ncu rs rc=$?"
tail -2 gpurun_out/ncu_mx.log gpurun_out/ncu_rs.log
