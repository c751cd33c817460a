#!/bin/bash
# round 2, call 24 (1 GPU): launch list of one k = 31 build (three bucket-size classes)
mkdir -p gpurun_out
PDG_ML_CLASSES=3 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_k31_classes.csv python tools/profile_step.py 100 31 > gpurun_out/r2_24_ncu.log 2>&1
echo "exit $?"; tail -n 2 gpurun_out/r2_24_ncu.log | cut -c1-200
