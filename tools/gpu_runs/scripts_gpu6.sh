#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/profile_step.py 20 > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k 'regex:m3_pass|m3_hist' -s 4 -c 4 -f -o gpurun_out/prof_pass python tools/profile_step.py 20 > gpurun_out/ncu2.log 2>&1
tail -n 3 gpurun_out/plain.log gpurun_out/ncu2.log | cut -c1-600
