#!/bin/bash
set -u
mkdir -p gpurun_out
CFG="python tools/bench_configs.py --only c4 --case ${RS_CASE:-c4_stt_b256}"
timeout 300 $CFG > gpurun_out/plain4.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tb_resize -s 8 -c 1 -f -o gpurun_out/prof_rs $CFG > gpurun_out/ncu_rs.log 2>&1; echo "ncu rs rc=$?"
