#!/bin/bash
# round 2, call 37 (1 GPU): ncu --set full of the k = 31 bucket-sort launches (three classes, CTA sort, rows) of the second, warm build
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'ml_class_sort|ml_mid_bucket|ml_rows' --launch-skip 5 -c 5 -f -o gpurun_out/r02_full_k31_classes python tools/profile_step.py 100 31 > gpurun_out/r2_37_ncu.log 2>&1
echo "ncu exit $?"; ls -la gpurun_out/r02_full_k31_classes.ncu-rep
