#!/bin/bash
# round 2, call 18 (1 GPU): ncu --set full with source of the k = 31 bucket sort (unpipelined loop)
mkdir -p gpurun_out
PDG_ML_PIPE=0 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'ml_bucket_sort' --launch-skip 1 -c 1 -f -o gpurun_out/r02_full_k31_bucket python tools/profile_step.py 100 31 > gpurun_out/r2_18_ncu.log 2>&1
echo "exit $?"; tail -n 3 gpurun_out/r2_18_ncu.log | cut -c1-300; ls -la gpurun_out/*.ncu-rep
