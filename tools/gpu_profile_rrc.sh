#!/bin/bash
set -u
mkdir -p gpurun_out
CFG="python tools/bench_configs.py --only rrc --case rrc_c2_b512"
timeout 300 $CFG > gpurun_out/plain_rrc.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tb_resize -s 8 -c 1 -f -o gpurun_out/prof_rrc $CFG > gpurun_out/ncu_rrc.log 2>&1; echo "ncu rrc rc=$?"
