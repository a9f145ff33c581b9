#!/bin/bash
# GPU-box script: the whole GPU suite, the default bench (both arms), the launch list of two steady frames.
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?"; tail -3 gpurun_out/t_gpu.log
python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python tools/prof_steady.py 2 > gpurun_out/prof_steady.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_steady.csv \
      python tools/prof_steady.py 2 > gpurun_out/ncu_steady.log 2>&1; echo "ncu rc=$?"
