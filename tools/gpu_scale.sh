#!/bin/bash
# Scaling run as the driver does it (one box, N = 1, 2, 4, 8 back to back) + the sample-sharded controller check.
# Usage (8-GPU box): gpurun --gpus 8 -- 'bash tools/gpu_scale.sh r1h'
set -u
TAG=${1:-r1}
G=${2:-8}
mkdir -p gpurun_out
timeout 300 python bench.py --gpus 1 --steps 300 --warmup 20 > gpurun_out/scale_${TAG}_1gpu.json 2> gpurun_out/scale_${TAG}_1gpu.err
echo "N=1 rc=$?"; cut -c1-300 gpurun_out/scale_${TAG}_1gpu.json
for N in 2 4 8; do
  [ $N -le $G ] || continue
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
      bench.py --gpus $N --steps 300 --warmup 20 > gpurun_out/scale_${TAG}_${N}gpu.json 2> gpurun_out/scale_${TAG}_${N}gpu.err
  echo "N=$N rc=$?"; cut -c1-300 gpurun_out/scale_${TAG}_${N}gpu.json
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29533 \
    tools/multi_gpu_check.py --exchange peer --ticks 300 --out gpurun_out/sharded_${TAG}_${G}gpu_peer.jsonl > gpurun_out/sharded_${TAG}_${G}gpu_peer.log 2>&1
echo "sharded rc=$?"; tail -3 gpurun_out/sharded_${TAG}_${G}gpu_peer.log | cut -c1-400
# the reference arm as the driver launches it at N > 1 (rank 0 alone runs it, on all host cores, N controllers per step)
timeout 300 python bench.py --impl reference --gpus 1 --steps 10 --warmup 3 > gpurun_out/scale_${TAG}_ref_1gpu.json 2>/dev/null; cut -c1-200 gpurun_out/scale_${TAG}_ref_1gpu.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29540 \
    bench.py --impl reference --gpus $G --steps 10 --warmup 3 > gpurun_out/scale_${TAG}_ref_${G}gpu.json 2>/dev/null; cut -c1-200 gpurun_out/scale_${TAG}_ref_${G}gpu.json
