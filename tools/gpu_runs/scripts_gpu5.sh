#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.log 2>&1; echo "exit $?" >> gpurun_out/bench_full.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "exit $?" >> gpurun_out/bench_ref.log
timeout 300 python tools/profile_step.py 20 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python tools/profile_step.py 20 > gpurun_out/ncu1.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -s 22 -c 22 -f -o gpurun_out/prof_step python tools/profile_step.py 20 > gpurun_out/ncu2.log 2>&1
tail -n 3 gpurun_out/bench_full.log gpurun_out/bench_ref.log gpurun_out/plain.log gpurun_out/ncu2.log | cut -c1-1500
