#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python tools/bench_configs.py ${CONFIG_ARGS:-} > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
cat gpurun_out/configs.jsonl | cut -c1-220; tail -5 gpurun_out/configs.err
