#!/bin/bash
# e2e scaling with and without NUMA pinning of the ranks (8-GPU box)
set -u
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
lscpu | grep -iE "numa|socket|^cpu\(s\)|model name" >> gpurun_out/topo.txt
for n in 8 4; do
  for aff in "" "--no-affinity"; do
    tag="n${n}${aff:+_noaff}"
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 300 --warmup 10 --no-cpu $aff > gpurun_out/aff_$tag.json 2> gpurun_out/aff_$tag.err; echo "$tag rc=$?"
    python -c "
import json; d=json.loads(open('gpurun_out/aff_$tag.json').read().strip().splitlines()[-1]); print('$tag', 'value', d['value'], 'e2e', d['e2e']['value'], d['config']['cpu_affinity'])"
  done
done
cat gpurun_out/topo.txt
