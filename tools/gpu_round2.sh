#!/bin/bash
# GPU round with the throughput kernel profiled as well: tests, probe, ncu --set full of (a) the 512-robot batched
# launch (sample-per-thread kernel) and (b) the single 10k tick (four-lanes kernel).  Outputs under gpurun_out/.
set -u
TAG=${1:-r1}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_gpu_$TAG.log
python tools/probe_latency.py phases > gpurun_out/probe_$TAG.log 2>&1; tail -7 gpurun_out/probe_$TAG.log
CMD="python bench.py --steps 12 --warmup 3 --multi-robots 512 --no-cpu-baseline"
if timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
  echo "ncu list rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:tick_kernel_quad -s 6 -c 1 -f -o gpurun_out/prof_quad_$TAG $CMD > gpurun_out/ncu_full_quad_$TAG.log 2>&1
  echo "ncu full (quad) rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:tick_kernel<" -s 2 -c 1 -f -o gpurun_out/prof_batched_$TAG $CMD > gpurun_out/ncu_full_batched_$TAG.log 2>&1
  echo "ncu full (batched) rc=$?"
else
  echo "plain run failed"; tail -20 gpurun_out/plain_$TAG.log
fi
ls -la gpurun_out | tail -12
