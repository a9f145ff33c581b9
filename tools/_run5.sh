cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for c in 0 61 119; do timeout 300 python tools/raft_update_trace.py $c > gpurun_out/r5_trace_cta$c.log 2>&1; done
(timeout 600 python -m pytest tests/test_gpu_conv_tc.py -q 2>&1 | tail -5) > gpurun_out/r5_unit.log 2>&1
timeout 600 python tools/grad_check.py rlv > gpurun_out/r5_gradcheck_full.log 2>&1
