#!/bin/bash
# bench at N GPUs only (our arm), printing the per-rank breakdown
set -u
mkdir -p gpurun_out
n=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 1000 --warmup 20 --no-cpu ${2:-} > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "rc=$?"
tail -2 gpurun_out/bench_n$n.err
python tools/show_bench.py gpurun_out/bench_n$n.json
python -c "
import json; d=json.loads(open('gpurun_out/bench_n$n.json').read().strip().splitlines()[-1]); print(d.get('per_rank'))"
