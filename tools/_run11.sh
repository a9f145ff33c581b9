cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests/test_gpu_train.py tests/test_gpu_conv_tc.py -x -q 2>&1 | tail -12) > gpurun_out/r11_train_tests.log 2>&1
timeout 600 python tools/train_profile.py > gpurun_out/r11_train_profile.log 2>&1
