#!/bin/bash
# One gpurun call: GPU parity tests, full bench line, then (same command line, plain run first) ncu launch list and
# one `--set full` capture of the tick kernel.  Outputs under gpurun_out/.
set -u
TAG=${1:-r1}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_gpu_$TAG.log
timeout 600 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; cat gpurun_out/bench_$TAG.json
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2>&1; cat gpurun_out/bench_ref_$TAG.json
CMD="python bench.py --steps 12 --warmup 3 --multi-robots 8 --no-cpu-baseline --no-sweep"
if timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
  echo "ncu list rc=$?"
  # the single 10k-sample tick (whichever variant qpmpc_create picked) ...
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:tick_kernel -s 6 -c 1 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
  echo "ncu full rc=$?"
  # ... and the batched throughput launch (512 robots x 10k)
  CMDB="python bench.py --steps 6 --warmup 3 --multi-robots 512 --no-cpu-baseline --no-sweep"
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^tick_kernel$" -s 2 -c 1 -f -o gpurun_out/prof_batched_$TAG $CMDB > gpurun_out/ncu_full_batched_$TAG.log 2>&1
  echo "ncu full batched rc=$?"
else
  echo "plain run failed"; tail -20 gpurun_out/plain_$TAG.log
fi
ls -la gpurun_out | tail -20
