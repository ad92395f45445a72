#!/bin/bash
# One gpurun call: GPU parity tests, full bench line + reference arm, then (same command lines, plain run first) the ncu
# launch list of the bench command and `--set full` captures of the three kernels that matter: the single 10k-sample tick,
# the batched throughput launch, the (tick, tail) pair of BASELINE config 3.  Outputs under gpurun_out/.
set -u
TAG=${1:-r2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 | tee gpurun_out/pytest_gpu_$TAG.log
timeout 600 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/bench_$TAG.json
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref_$TAG.json 2>&1; cut -c1-300 gpurun_out/bench_ref_$TAG.json
CMD="python bench.py --steps 12 --warmup 3 --multi-robots 8 --no-cpu-baseline --no-sweep --no-configs"
if timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
  echo "ncu list rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:tick_kernel -s 6 -c 1 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
  echo "ncu full rc=$?"
else
  echo "plain run failed"; tail -20 gpurun_out/plain_$TAG.log
fi
if timeout 120 python tools/ncu_one.py 128 10000 1 > /dev/null 2>&1; then
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^tick_kernel$" -s 1 -c 1 -f -o gpurun_out/prof_batched_$TAG python tools/ncu_one.py 128 10000 1 > gpurun_out/ncu_full_batched_$TAG.log 2>&1
  echo "ncu full batched rc=$?"
fi
if timeout 120 python tools/ncu_one.py 1 50000 0 4 config2 > /dev/null 2>&1; then
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:tick_kernel|tail_kernel" -s 2 -c 2 -f -o gpurun_out/prof_config2_$TAG python tools/ncu_one.py 1 50000 0 4 config2 > gpurun_out/ncu_full_config2_$TAG.log 2>&1
  echo "ncu full config2 rc=$?"
fi
# summarise on the box (the reports together exceed what gpurun copies back); keep the summaries, drop the reports
for f in prof_$TAG prof_batched_$TAG prof_config2_$TAG; do
  if [ -f gpurun_out/$f.ncu-rep ]; then
    python tools/ncu_summary.py gpurun_out/$f.ncu-rep gpurun_out/$f --top 40 > /dev/null 2>&1
    python tools/ncu_lines.py gpurun_out/$f.ncu-rep 40 gpurun_out/${f}_lines.csv > /dev/null 2>&1
    rm -f gpurun_out/$f.ncu-rep
  fi
done
ls -la gpurun_out | tail -30
