#!/bin/bash
mkdir -p gpurun_out
export PDG_LIBPDGPU=$PWD/prairiedog_b200/build/variants/libpdgpu_${VAR}.so
timeout 300 python tools/profile_step.py 20 > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k 'regex:m3_pass' -s 3 -c 3 -f -o gpurun_out/prof_var python tools/profile_step.py 20 > gpurun_out/ncu2.log 2>&1
tail -n 2 gpurun_out/ncu2.log | cut -c1-300
