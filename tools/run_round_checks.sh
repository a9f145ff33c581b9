#!/bin/bash
# GPU-box script: new no-alloc test alone, the whole GPU suite, the default bench (both arms), one ncu capture.
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/t_noalloc.log 2>&1; echo "noalloc rc=$?"
python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?"; tail -3 gpurun_out/t_gpu.log
python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python tools/prof_raft_update.py > gpurun_out/prof_ru.log 2>&1 && \
  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:raft_update -c 1 \
      -o gpurun_out/r8_raft_update -f python tools/prof_raft_update.py > gpurun_out/ncu_ru.log 2>&1; echo "ncu rc=$?"
