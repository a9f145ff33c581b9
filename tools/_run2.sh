cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 300 python -m pytest tests/test_gpu_raft_update.py -x -q 2>&1 | tail -40) > gpurun_out/r2_raft_update.log 2>&1
if grep -q "passed" gpurun_out/r2_raft_update.log && ! grep -q "failed\|error" gpurun_out/r2_raft_update.log; then
  echo "fused ok" >> gpurun_out/r2_raft_update.log
else
  export ZT_RAFT_FUSED=0
  echo "FUSED DISABLED for the rest of this run" >> gpurun_out/r2_raft_update.log
fi
timeout 300 python tools/raft_iter_time.py > gpurun_out/r2_raft_iter.log 2>&1
(timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_raft_update.py 2>&1 | tail -30) > gpurun_out/r2_pytest.log 2>&1
timeout 900 python bench.py --steps 30 --warmup 5 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err
timeout 600 python tools/grad_conditioning.py rlv > gpurun_out/r2_gradcond_rlv.txt 2>&1
timeout 600 python tools/grad_conditioning.py 540x960 > gpurun_out/r2_gradcond_540.txt 2>&1
