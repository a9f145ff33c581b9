cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python tools/raft_update_trace.py 0 > gpurun_out/r8_trace_cta0.log 2>&1
