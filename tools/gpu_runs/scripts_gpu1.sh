#!/bin/bash
# first GPU call: tests, smoke, small + full bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/nproc.txt; lscpu | grep "Model name" >> gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
timeout 300 python bench.py --genomes 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_small.log 2>&1; echo "exit $?" >> gpurun_out/bench_small.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.log 2>&1; echo "exit $?" >> gpurun_out/bench_full.log
tail -3 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench_small.log gpurun_out/bench_full.log
