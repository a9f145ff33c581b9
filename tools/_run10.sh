cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for cfg in "2 74" "2 60" "3 49" "3 74"; do
set -- $cfg
ZT_SHARDS_PER_GPU=$1 ZT_RAFT_FUSED_CTAS=$2 timeout 900 python bench.py --steps 30 --warmup 5 --no-train --no-cpu-baseline --no-profile > gpurun_out/r10_bench_s$1_g$2.json 2> gpurun_out/r10_bench_s$1_g$2.err
done
