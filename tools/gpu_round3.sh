#!/bin/bash
set -u
mkdir -p gpurun_out
echo "== pytest -m gpu"
timeout 1200 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
echo "== configs"
timeout 900 python tools/bench_configs.py ${CONFIG_ARGS:-} > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
cat gpurun_out/configs.jsonl | cut -c1-200; tail -5 gpurun_out/configs.err
echo "== bench"
timeout 600 python bench.py --steps 500 --warmup 10 --no-cpu > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json | cut -c1-1500; tail -3 gpurun_out/bench.err
