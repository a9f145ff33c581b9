cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(timeout 1800 python -m pytest tests -m gpu -q 2>&1 | tail -8) > gpurun_out/rA_pytest.log 2>&1
timeout 900 python tools/corr_warp_sweep.py > gpurun_out/rA_corr_warp_sweep.txt 2>&1
timeout 300 python tools/raft_update_trace.py 0 > gpurun_out/rA_trace.txt 2>&1
python tools/prof_steady.py 2 > gpurun_out/rA_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/rA_launches.csv python tools/prof_steady.py 2 > gpurun_out/rA_ncu.log 2>&1
