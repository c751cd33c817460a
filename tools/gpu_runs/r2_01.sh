#!/bin/bash
# round 2, call 1: pipe-throughput microbenchmarks + baseline of the round-1 state on today's box
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2_01_smi.log 2>&1
timeout 300 tools/pipelab > gpurun_out/r2_01_pipelab.log 2>&1; echo "exit $?" >> gpurun_out/r2_01_pipelab.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_01_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_01_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_01_bench.log 2>&1; echo "exit $?" >> gpurun_out/r2_01_bench.log
cat gpurun_out/r2_01_pipelab.log; tail -n 3 gpurun_out/r2_01_pytest.log; cut -c1-600 gpurun_out/r2_01_bench.log
