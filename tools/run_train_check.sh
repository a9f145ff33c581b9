#!/bin/bash
# GPU-box script: training tests, then the bench with the training leg (no CPU leg)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_train.py -x -q -m gpu > gpurun_out/t_train.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_train.log
python bench.py --no-cpu-baseline > gpurun_out/bench_T.json 2> gpurun_out/bench_T.err; echo "bench rc=$?"
