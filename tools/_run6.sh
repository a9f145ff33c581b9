cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_raft_update.py -q 2>&1 | tail -15) > gpurun_out/r6_unit.log 2>&1
for c in 0 119; do timeout 300 python tools/raft_update_trace.py $c > gpurun_out/r6_trace_cta$c.log 2>&1; done
timeout 300 python tools/raft_iter_time.py > gpurun_out/r6_raft_iter.log 2>&1
