cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_raft_update.py tests/test_gpu_ops.py -q 2>&1 | tail -40) > gpurun_out/r3_raft_update.log 2>&1
timeout 300 python tools/raft_iter_time.py > gpurun_out/r3_raft_iter.log 2>&1
ZT_RAFT_FUSED_BN=64,64,64,64,64,64,64,64,64 timeout 300 python tools/raft_iter_time.py > gpurun_out/r3_raft_iter_bn64.log 2>&1
ZT_RAFT_FUSED_BN=32,32,32,32,32,32,32,32,32 timeout 300 python tools/raft_iter_time.py > gpurun_out/r3_raft_iter_bn32.log 2>&1
(timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_raft_update.py --deselect tests/test_gpu_ops.py 2>&1 | tail -30) > gpurun_out/r3_pytest.log 2>&1
timeout 900 python bench.py --steps 30 --warmup 5 --no-train --no-cpu-baseline > gpurun_out/r3_bench.json 2> gpurun_out/r3_bench.err
ZT_SHARDS_PER_GPU=1 timeout 900 python bench.py --steps 30 --warmup 5 --no-train --no-cpu-baseline --no-profile > gpurun_out/r3_bench_s1.json 2> gpurun_out/r3_bench_s1.err
ZT_SHARDS_PER_GPU=2 timeout 900 python bench.py --steps 30 --warmup 5 --no-train --no-cpu-baseline --no-profile > gpurun_out/r3_bench_s2.json 2> gpurun_out/r3_bench_s2.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r3_launches.csv python bench.py --steps 6 --warmup 3 --no-train --no-cpu-baseline --no-profile > gpurun_out/r3_ncu.log 2>&1
