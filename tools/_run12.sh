cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_train.py -x -q -k "wgrad" 2>&1 | tail -25) > gpurun_out/r12_wgrad.log 2>&1
(timeout 1200 python -m pytest tests/test_gpu_train.py -q 2>&1 | tail -12) > gpurun_out/r12_train_tests.log 2>&1
timeout 600 python tools/train_profile.py > gpurun_out/r12_train_profile.log 2>&1
