#!/bin/bash
# Bench + ncu evidence for profiles/: the driver-shaped bench line, the reference arm, the launch list of the same
# command and one full capture of the dominant kernel of every BASELINE configuration.
# Every ncu run follows a plain run of the same command (&&).  Logs land in gpurun_out/.
set -u
mkdir -p gpurun_out
echo "== bench (driver-shaped: --steps 20 --warmup 5)"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_driver.json 2> gpurun_out/bench_driver.err; echo "bench rc=$?"
cut -c1-300 gpurun_out/bench_driver.json; tail -3 gpurun_out/bench_driver.err
echo "== reference arm"
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-300 gpurun_out/bench_ref.json
echo "== ncu launch list"
BENCH="python bench.py --steps 20 --warmup 5 --no-cpu --no-e2e --no-configs"
timeout 300 $BENCH > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_launch.log 2>&1; echo "launch list rc=$?"
capture () {   # name, config, kernel regex, launches to skip
  CFG="python bench.py --configs-only $2"
  timeout 300 $CFG > gpurun_out/plain_$1.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$3 -s $4 -c 1 -f -o gpurun_out/prof_$1 $CFG > gpurun_out/ncu_$1.log 2>&1; echo "ncu $1 rc=$?"
}
capture c1 c1 tb_xf_generic 8
capture c2 c2 tb_xf_stream 8
capture c3 c3 tb_xf_stream 8
capture c4 c4 tb_resize_ring 8
capture c4_b256 c4_b256 tb_resize_ring 8
capture c5 c5 tb_metrics_row 4
capture c5_bf16 c5_bf16 tb_metrics_row 4
