#!/bin/bash
# GPU-box script: shards-per-GPU sweep of the headline bench (no training / profile / CPU legs)
mkdir -p gpurun_out
for S in 2 3 4 5; do
  ZT_SHARDS_PER_GPU=$S python bench.py --no-train --no-profile --no-cpu-baseline > gpurun_out/bench_S$S.json 2> gpurun_out/bench_S$S.err; echo "S=$S rc=$?"
done
