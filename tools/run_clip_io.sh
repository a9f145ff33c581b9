#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_clip_io.py tests/test_gpu_e2e.py tests/test_dropin.py -x -q -m gpu > gpurun_out/t_clipio.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_clipio.log
timeout 150 python tools/clip_io_bench.py 16 > gpurun_out/clip_io_bench.txt 2>&1; echo "bench rc=$?"; tail -6 gpurun_out/clip_io_bench.txt
