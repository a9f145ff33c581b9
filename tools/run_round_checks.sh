#!/bin/bash
# GPU-box script: the whole GPU suite, smoke, the default bench, the 4K config on one GPU, a CTA-share sweep for 3 shards.
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?"; tail -3 gpurun_out/t_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench rc=$?"
python bench.py --config uw4k --no-train --no-profile --no-cpu-baseline > gpurun_out/bench_uw4k_1gpu.json 2> gpurun_out/bench_uw4k_1gpu.err; echo "uw4k rc=$?"
for C in 40 60 74; do
  ZT_SHARD_CTAS=$C python bench.py --no-train --no-profile --no-cpu-baseline > gpurun_out/bench_C$C.json 2> gpurun_out/bench_C$C.err; echo "C=$C rc=$?"
done
