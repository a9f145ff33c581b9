cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python tools/prof_raft_update.py > gpurun_out/r7_plain.log 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:raft_update -c 1 -f -o gpurun_out/r7_raft_update python tools/prof_raft_update.py > gpurun_out/r7_ncu.log 2>&1
