#!/bin/bash
set -u
mkdir -p gpurun_out
CFG="python tools/bench_configs.py --only aug --case aug_cifar_b2048"
timeout 300 $CFG > gpurun_out/plain_aug.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tb_augment -s 8 -c 1 -f -o gpurun_out/prof_aug $CFG > gpurun_out/ncu_aug.log 2>&1; echo "ncu aug rc=$?"
