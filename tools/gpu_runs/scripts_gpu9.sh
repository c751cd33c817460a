#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.log 2>&1; echo "exit $?" >> gpurun_out/bench_full.log
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "exit $?" >> gpurun_out/bench_ref.log
timeout 900 python bench.py --k 31 --genomes 100 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_k31.log 2>&1; echo "exit $?" >> gpurun_out/bench_k31.log
timeout 900 python tools/bench_queries.py --genomes 100 --k 11 > gpurun_out/queries_k11.log 2>&1; echo "exit $?" >> gpurun_out/queries_k11.log
timeout 900 python tools/bench_queries.py --genomes 8 --k 11 --check > gpurun_out/queries_k11_check.log 2>&1; echo "exit $?" >> gpurun_out/queries_k11_check.log
timeout 900 python tools/bench_queries.py --genomes 40 --k 31 > gpurun_out/queries_k31.log 2>&1; echo "exit $?" >> gpurun_out/queries_k31.log
timeout 300 python tools/profile_step.py 20 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python tools/profile_step.py 20 > gpurun_out/ncu1.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -c 60 -f -o gpurun_out/prof_step python tools/profile_step.py 20 > gpurun_out/ncu2.log 2>&1
tail -n 2 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench_ref.log gpurun_out/bench_k31.log gpurun_out/queries_k11.log gpurun_out/queries_k11_check.log gpurun_out/queries_k31.log | cut -c1-1200
