#!/bin/bash
mkdir -p gpurun_out
python tools/prof_steady.py 1 > gpurun_out/prof_steady.log 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:conv_tma -c 9 \
    -o gpurun_out/r10_conv_tma -f python tools/prof_steady.py 1 > gpurun_out/ncu_ct.log 2>&1; echo "ncu rc=$?"
