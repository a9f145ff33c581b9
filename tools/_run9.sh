cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for S in 1 2 3; do
ZT_SHARDS_PER_GPU=$S timeout 900 python bench.py --steps 30 --warmup 5 --no-train --no-cpu-baseline --no-profile > gpurun_out/r9_bench_s$S.json 2> gpurun_out/r9_bench_s$S.err
done
ZT_RAFT_FUSED=0 ZT_SHARDS_PER_GPU=3 timeout 900 python bench.py --steps 30 --warmup 5 --no-train --no-cpu-baseline --no-profile > gpurun_out/r9_bench_unfused_s3.json 2> gpurun_out/r9_bench_unfused_s3.err
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) > gpurun_out/r9_pytest.log 2>&1
