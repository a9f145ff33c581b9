cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python tools/raft_update_trace.py > gpurun_out/r4_trace.log 2>&1
ZT_RAFT_FUSED_BN=64,64,64,64,64,64,64,64,64 timeout 300 python tools/raft_update_trace.py > gpurun_out/r4_trace_bn64.log 2>&1
(timeout 900 python -m pytest tests/test_gpu_conv_tc.py tests/test_gpu_raft_update.py tests/test_gpu_ops.py -q 2>&1 | tail -30) > gpurun_out/r4_unit.log 2>&1
# mode 4 (scaled fp16 split) end to end
(ZT_TC_RAFT=4 timeout 900 python -m pytest tests/test_gpu_raft_update.py tests/test_gpu_e2e.py tests/test_gpu_recurrence.py -q -k "not 1080" 2>&1 | tail -30) > gpurun_out/r4_mode4_e2e.log 2>&1
cp gpurun_out/recurrence_390x400_64frames.txt gpurun_out/r4_recurrence_390x400_mode4.txt
for mode in "ZT_TC_RAFT=4 ZT_TC_TRAIN_D1=4 ZT_TC_TRAIN_ENH=4 ZT_TC_TRAIN_D2=4" "ZT_TC_RAFT=4" "ZT_TC_TRAIN_D1=4 ZT_TC_TRAIN_ENH=4 ZT_TC_TRAIN_D2=4"; do
  echo "=== mode: $mode" >> gpurun_out/r4_gradcheck.log
  env $mode timeout 600 python tools/grad_check.py rlv 2>&1 | grep -v Warning | tail -22 >> gpurun_out/r4_gradcheck.log
done
echo "=== 540x960 all mode 4" >> gpurun_out/r4_gradcheck.log
ZT_TC_RAFT=4 ZT_TC_TRAIN_D1=4 ZT_TC_TRAIN_ENH=4 ZT_TC_TRAIN_D2=4 timeout 600 python tools/grad_check.py 540x960 2>&1 | grep -v Warning | tail -22 >> gpurun_out/r4_gradcheck.log
ZT_TC_RAFT=4 timeout 300 python tools/raft_iter_time.py > gpurun_out/r4_raft_iter_mode4.log 2>&1
