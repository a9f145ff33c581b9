#!/bin/bash
# GPU-box script: bench.py under torchrun on N GPUs.  usage: run_multi_gpu.sh N [1080p|uw4k|both]
N=${1:-8}; WHAT=${2:-1080p}
mkdir -p gpurun_out
TR="timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
if [ "$WHAT" != "uw4k" ]; then
  $TR bench.py --gpus $N > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "1080p rc=$?"
  tail -c 400 gpurun_out/bench_${N}gpu.err
fi
if [ "$WHAT" != "1080p" ]; then
  $TR bench.py --gpus $N --config uw4k --no-train --no-profile --no-cpu-baseline > gpurun_out/bench_uw4k_${N}gpu.json 2> gpurun_out/bench_uw4k_${N}gpu.err; echo "uw4k rc=$?"
fi
