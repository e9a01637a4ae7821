#!/bin/bash
set -u
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi -L
echo "== multi-GPU pytest"
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_pipeline.py -q -m gpu > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_multi.log
for n in 1 $N; do
  echo "== bench N=$n"
  if [ "$n" = "1" ]; then
    timeout 600 python bench.py --gpus 1 --steps 1000 --warmup 20 --no-cpu > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?"
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 1000 --warmup 20 --no-cpu > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "rc=$?"
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $n --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref_n$n.json 2> gpurun_out/bench_ref_n$n.err; echo "ref rc=$?"; cut -c1-200 gpurun_out/bench_ref_n$n.json
  fi
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n$n.json").read().strip().splitlines()[-1])
    print({k:d[k] for k in ("value","n_gpus","ms_per_step","gpu_launches")}, d["roofline"]["frac"], d["e2e"]["value"], d["clocks"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/bench_n$n.err").read()[-1500:])
PY
done
