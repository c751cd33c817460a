#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.log 2>&1; echo "exit $?" >> gpurun_out/bench_full.log
timeout 300 python tools/profile_step.py 20 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python tools/profile_step.py 20 > gpurun_out/ncu1.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:onesweep|traverse|radix_hist|g_rows|g_head' -s 7 -c 7 -f -o gpurun_out/prof_build python tools/profile_step.py 20 > gpurun_out/ncu2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:fasta' -s 60 -c 3 -f -o gpurun_out/prof_ingest python tools/profile_step.py 20 > gpurun_out/ncu3.log 2>&1
tail -n 3 gpurun_out/pytest_gpu.log gpurun_out/bench_full.log gpurun_out/plain.log gpurun_out/ncu2.log
