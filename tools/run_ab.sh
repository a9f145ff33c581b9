#!/bin/bash
# GPU-box script: conv_tma / whole-frame tests, then the bench without training / CPU legs; optional second arm with env B
mkdir -p gpurun_out
python -m pytest tests/test_gpu_conv_tma.py tests/test_gpu_e2e.py -x -q -m gpu > gpurun_out/t_ct.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_ct.log
python bench.py --no-train --no-cpu-baseline > gpurun_out/bench_A.json 2> gpurun_out/bench_A.err; echo "A rc=$?"
if [ -n "$1" ]; then env $1 python bench.py --no-train --no-cpu-baseline > gpurun_out/bench_B.json 2> gpurun_out/bench_B.err; echo "B ($1) rc=$?"; fi
